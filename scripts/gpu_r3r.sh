#!/bin/bash
# r3r: last variants of the headline kernel: two quads held at the TMEM release (TC_TAIL_K = 2), column splits 6/11 and 5/10
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
for v in base3 k2 s6_11 s5_10 base3; do
  [ -f $V/lib$v.so ] || { echo "$v: not built"; continue; }
  B200MEL_LIB=$V/lib$v.so timeout 300 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -x -q -k "config2 or libri" > $O/r3r_${v}_pytest.log 2>&1
  echo "$v pytest: $(tail -n 1 $O/r3r_${v}_pytest.log)"
  for i in 1 2; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3r_${v}_$i.json 2> $O/r3r_${v}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3r_${v}_$i.json")); print("$v #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("$v failed", e)
PY
  done
done
