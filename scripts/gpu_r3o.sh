#!/bin/bash
# r3o: C = 3 copy-out (deltas) handed to the producer warps, one tile behind the epilogue
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libpco.so timeout 900 python -m pytest tests/test_parity_fullsize_gpu.py tests/test_guard_bands_gpu.py tests/test_parity_gpu.py -m gpu -q -x --timeout 300 > $O/r3o_pytest.log 2>&1; echo "pytest exit $?" >> $O/r3o_pytest.log; tail -n 6 $O/r3o_pytest.log
for c in 3d 5; do
  B200MEL_LIB=$V/libpco.so timeout 600 python bench.py --config $c --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $O/r3o_config_$c.json 2> $O/r3o_config_$c.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3o_config_$c.json")); print("config $c: %.1f us  %.3e audio-s/s frac %.4f" % (d["ms_per_step"]*1e3, d["value"], d["roofline"]["frac"]))
except Exception as e: print("config $c failed", e)
PY
done
B200MEL_LIB=$V/libpco.so timeout 900 python experiments/config_sweep.py 2>/dev/null | grep -E "C=3|C=1 " > $O/r3o_configs.jsonl
python - <<PY
import json
for l in open("$O/r3o_configs.jsonl"):
    try:
        d=json.loads(l); print("%-80s %-8s %8.1f us  %.3f" % (d["case"][:80], d["kernel"], d["us_per_batch"], d["frac_of_hbm_peak"]))
    except Exception as e: print(l.strip()[:120])
PY
