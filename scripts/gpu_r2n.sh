#!/bin/bash
# r2n: bound steps + spin-up sweep on the default build, then the full GPU suite
mkdir -p gpurun_out
O=gpurun_out
for i in 1 2 3; do
  for s in 0 8 15 30; do
    timeout 300 python bench.py --no-cpu-baseline --e2e-steps 2 --steps 20 --warmup 3 --spin-up-ms $s > $O/r2n_${s}_$i.json 2> $O/r2n_${s}_$i.err
    python - <<PY
import json
try:
    d=json.load(open("$O/r2n_${s}_$i.json")); print("spin $s #$i: %.2f us  frac %.4f host %.1f us %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["roofline"]["host_enqueue_us_per_step"], d["clocks"]))
except Exception as e: print("failed", e)
PY
  done
done
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 2 --steps 20 --warmup 3 --trace-steps > $O/r2n_trace.json 2>$O/r2n_trace.err
python -c "
import json; d=json.load(open('$O/r2n_trace.json')); print(d['step_us'])"
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > $O/r2n_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2n_pytest.log
tail -n 4 $O/r2n_pytest.log
