#!/bin/bash
# r3s: generic kernel without per-element divisions (FFT index of power-of-two passes, row-wise copy-out): tests + A/B sweep
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libgen2.so timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands_gpu.py -m gpu -q -x --timeout 300 > $O/r3s_pytest.log 2>&1; echo "pytest exit $?" >> $O/r3s_pytest.log; tail -n 3 $O/r3s_pytest.log
for v in gen1 gen2; do
  B200MEL_LIB=$V/lib$v.so timeout 900 python experiments/config_sweep.py 2>/dev/null | grep -E '"kernel": "generic"' > $O/r3s_$v.jsonl
done
python - <<PY
import json
a=[json.loads(l) for l in open("$O/r3s_gen1.jsonl")]; b=[json.loads(l) for l in open("$O/r3s_gen2.jsonl")]
for x,y in zip(a,b): print("%-70s %8.1f -> %8.1f us  (%.3f -> %.3f)" % (x["case"][:70], x["us_per_batch"], y["us_per_batch"], x["frac_of_hbm_peak"], y["frac_of_hbm_peak"]))
PY
