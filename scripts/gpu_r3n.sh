#!/bin/bash
# r3n: MFCC second stage compiled for 3 vs 4 resident blocks per SM
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
for i in 1 2 3; do
for v in mf3 mf4; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python bench.py --config 4k --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $O/r3n_${v}_$i.json 2> $O/r3n_${v}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3n_${v}_$i.json")); print("$v #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("$v failed", e)
PY
done
done
B200MEL_LIB=$V/libmf4.so timeout 600 python -m pytest tests -m gpu -q -x -k "mfcc or no_alloc or shipped" > $O/r3n_pytest.log 2>&1; tail -n 2 $O/r3n_pytest.log
