#!/bin/bash
# r2v: float32 producers with the OR-accumulated exactness test
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "float32 or f32 or guard or shipped or sanit" > $O/r2v_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2v_pytest.log
tail -n 8 $O/r2v_pytest.log
timeout 600 python experiments/sanitize_driver.py > $O/r2v_driver.log 2>&1; echo "driver exit $?"; tail -n 3 $O/r2v_driver.log
timeout 900 python experiments/config_sweep.py > $O/r2v_configs.jsonl 2> $O/r2v_configs.err; echo "sweep exit $?"
B200MEL_F32_TWO_PASS=1 timeout 900 python experiments/config_sweep.py 2>/dev/null | grep float32 > $O/r2v_configs_two_pass.jsonl
python - <<PY
import json
for f in ("$O/r2v_configs.jsonl", "$O/r2v_configs_two_pass.jsonl"):
    for l in open(f):
        try:
            d=json.loads(l); print("%-80s %-8s %8.1f us  %.3f" % (d["case"][:80], d["kernel"], d["us_per_batch"], d["frac_of_hbm_peak"]))
        except Exception as e: print(l.strip()[:120])
PY
tail -n 5 $O/r2v_configs.err
