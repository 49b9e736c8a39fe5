#!/bin/bash
# r2i: MFCC second stage on mma.sync (TF32 x3), streaming f32->int16 conversion, no-alloc entry point
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python __graft_entry__.py smoke > $O/r2i_smoke.log 2>&1; echo "smoke exit $?" | tee -a $O/r2i_smoke.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x --timeout 300 > $O/r2i_noalloc_alone.log 2>&1; echo "noalloc alone exit $?" | tee -a $O/r2i_noalloc_alone.log
tail -n 3 $O/r2i_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 --durations=8 > $O/r2i_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2i_pytest.log
tail -n 16 $O/r2i_pytest.log
for tpb in 0 1 2 3 5 7 13; do
  B200MEL_MFCC_TILES_PER_BLOCK=$tpb timeout 300 python bench.py --config 4k --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $O/r2i_4k_tpb$tpb.json 2> $O/r2i_4k_tpb$tpb.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2i_4k_tpb$tpb.json")); print("4k tpb=$tpb: %.1f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("4k tpb=$tpb failed", e)
PY
done
timeout 900 python experiments/config_sweep.py > $O/r2i_configs.jsonl 2> $O/r2i_configs.err; echo "sweep exit $?"
python - <<PY
import json
for l in open("$O/r2i_configs.jsonl"):
    try:
        d=json.loads(l); print("%-46s %-8s %8.1f us  %.3f" % (d["case"], d["kernel"], d["us_per_batch"], d["frac_of_hbm_peak"]))
    except Exception as e: print(l.strip()[:120])
PY
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r2i_bench_20.json 2> $O/r2i_bench_20.err; echo "bench exit $?"; cat $O/r2i_bench_20.json | cut -c1-900
