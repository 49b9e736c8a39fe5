#!/bin/bash
# r3k: the new list pass built like the shipped library (r3j's copy had been built with --split-compile: 5 us slower)
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
for i in 1 2 3; do
for v in notouch list2; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3k_${v}_$i.json 2> $O/r3k_${v}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3k_${v}_$i.json")); print("$v #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("$v failed", e)
PY
done
done
