#!/bin/bash
# r3h: lengths touched (real loads, L2 only, values discarded) before the dependency wait - A/B + trace
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > $O/r3h_trace.txt 2>&1; sed -n 1,4p $O/r3h_trace.txt
for i in 1 2 3; do
for v in notouch touch; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3h_${v}_$i.json 2> $O/r3h_${v}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3h_${v}_$i.json")); print("$v #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("$v failed", e)
PY
done
done
