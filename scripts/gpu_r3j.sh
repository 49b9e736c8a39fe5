#!/bin/bash
# r3j: tile list with one utterance per thread (warp sums through shared memory): trace, tests, A/B against notouch (= previous list pass)
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > $O/r3j_trace.txt 2>&1; sed -n 1,4p $O/r3j_trace.txt
B200MEL_LIB=$V/liblist2.so timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > $O/r3j_pytest.log 2>&1; echo "pytest exit $?" >> $O/r3j_pytest.log; tail -n 4 $O/r3j_pytest.log
for i in 1 2 3; do
for v in notouch list2; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3j_${v}_$i.json 2> $O/r3j_${v}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3j_${v}_$i.json")); print("$v #$i: %.2f us  frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("$v failed", e)
PY
done
done
