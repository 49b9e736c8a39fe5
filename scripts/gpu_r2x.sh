#!/bin/bash
# r2x: MFCC second stage with the DCT folded by its mirror symmetry
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "mfcc or no_alloc or guard or shipped" > $O/r2x_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2x_pytest.log
tail -n 6 $O/r2x_pytest.log
timeout 600 python experiments/sanitize_driver.py > $O/r2x_driver.log 2>&1; echo "driver exit $?"; grep mfcc $O/r2x_driver.log
for i in 1 2; do
  timeout 300 python bench.py --config 4k --steps 50 --warmup 5 --e2e-steps 2 > $O/r2x_4k_$i.json 2> $O/r2x_4k_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2x_4k_$i.json")); print("4k: %.1f us  frac %.4f  parity %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], (d.get("cpu_baseline") or {}).get("parity_max_abs_err")))
except Exception as e: print("4k failed", e)
PY
done
