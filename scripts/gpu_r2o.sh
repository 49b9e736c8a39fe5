#!/bin/bash
# r2o: MFCC second stage v2 (fragments straight from global memory, 256-thread blocks): tests + tiles-per-block sweep
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "mfcc or no_alloc or guard or shipped" > $O/r2o_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2o_pytest.log
tail -n 6 $O/r2o_pytest.log
for tpb in 0 1 2 3 4 6; do
  B200MEL_MFCC_TILES_PER_BLOCK=$tpb timeout 300 python bench.py --config 4k --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $O/r2o_4k_tpb$tpb.json 2> $O/r2o_4k_tpb$tpb.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2o_4k_tpb$tpb.json")); print("4k tpb=$tpb: %.1f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("4k tpb=$tpb failed", e)
PY
done
