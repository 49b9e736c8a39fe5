#!/bin/bash
# r3d: lengths pulled into L2 before the dependency wait - A/B against the same source without it
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
for i in 1 2 3; do
for v in base2 lenpf; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3d_${v}_$i.json 2> $O/r3d_${v}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3d_${v}_$i.json")); print("$v #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("$v failed", e)
PY
done
done
B200MEL_LIB=$V/liblenpf.so timeout 600 python -m pytest tests/test_parity_fullsize_gpu.py tests/test_guard_bands_gpu.py -m gpu -x -q > $O/r3d_pytest.log 2>&1; tail -n 2 $O/r3d_pytest.log
