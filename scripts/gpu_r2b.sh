#!/bin/bash
# r2b: first run of the three-group epilogue kernel: smoke, full GPU test-suite, variants A/B, bench
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2b_smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/r2b_smoke.log
tail -n 4 gpurun_out/r2b_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 --durations=8 > gpurun_out/r2b_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2b_pytest.log
tail -n 14 gpurun_out/r2b_pytest.log
bench() {  # name lib steps warmup
  B200MEL_LIB=$2 timeout 300 python bench.py --steps $3 --warmup $4 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2b_bench_$1.json 2> gpurun_out/r2b_bench_$1.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2b_bench_$1.json")); print("$1: %.1f us  frac %.4f  host %.1f us  %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["roofline"]["host_enqueue_us_per_step"], d["clocks"]))
except Exception as e: print("$1: bench failed", e)
PY
}
V=$PWD/experiments/_variants
bench r1 $V/libr1.so 100 10
bench new "" 100 10
for v in e112p72 e104p80 s5_11 s4_10; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -x -q -k "config2 or libri" > gpurun_out/r2b_${v}_pytest.log 2>&1
  echo "$v pytest: $(tail -n 1 gpurun_out/r2b_${v}_pytest.log)"
  bench $v $V/lib$v.so 100 10
done
bench new20a "" 20 3
bench new20b "" 20 3
bench new200 "" 200 20
bench r1_20 $V/libr1.so 20 3
