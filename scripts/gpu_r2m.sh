#!/bin/bash
# r2m: spin-up A/B, alternating, 6 samples each
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
for i in 1 2 3 4 5 6; do
  for s in 0 30 100; do
    B200MEL_LIB=$V/libspec.so timeout 300 python bench.py --no-cpu-baseline --e2e-steps 2 --steps 20 --warmup 3 --spin-up-ms $s > $O/r2m_${s}_$i.json 2> $O/r2m_${s}_$i.err
    python - <<PY
import json
try:
    d=json.load(open("$O/r2m_${s}_$i.json")); print("spin $s #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("failed", e)
PY
  done
done
