#!/bin/bash
# r2w: producers load interior groups without clamping (one base address); float32 groups as ten loads in a row
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > $O/r2w_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2w_pytest.log
tail -n 5 $O/r2w_pytest.log
for i in 1 2; do
for k in "20 3" "100 10"; do
  set -- $k
  timeout 300 python bench.py --steps $1 --warmup $2 --no-cpu-baseline --e2e-steps 2 > $O/r2w_bench_$1_$i.json 2> $O/r2w_bench_$1_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2w_bench_$1_$i.json")); print("bench $1/$2: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("bench failed", e)
PY
done
done
timeout 900 python experiments/config_sweep.py > $O/r2w_configs.jsonl 2> $O/r2w_configs.err; echo "sweep exit $?"
python - <<PY
import json
for l in open("$O/r2w_configs.jsonl"):
    try:
        d=json.loads(l); print("%-80s %-8s %8.1f us  %.3f" % (d["case"][:80], d["kernel"], d["us_per_batch"], d["frac_of_hbm_peak"]))
    except Exception as e: print(l.strip()[:120])
PY
