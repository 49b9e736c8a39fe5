#!/bin/bash
# r2t: MFCC second stage v6 (f16 m16n8k16 x3, 128-row tiles, one resident wave)
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "mfcc or no_alloc or guard or shipped" > $O/r2t_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2t_pytest.log
tail -n 6 $O/r2t_pytest.log
for rev in 1 0 1 0; do
  B200MEL_MFCC_REVERSE=$rev timeout 300 python bench.py --config 4k --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $O/r2t_4k_rev$rev.json 2> $O/r2t_4k_rev$rev.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2t_4k_rev$rev.json")); print("4k reverse=$rev: %.1f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("4k rev=$rev failed", e)
PY
done
