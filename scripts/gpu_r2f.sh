#!/bin/bash
# r2f: two fat epilogue groups with the tile's last K quads prefetched into registers (early TMEM release)
mkdir -p gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libg2k3.so timeout 900 python -m pytest tests/test_parity_fullsize_gpu.py tests/test_parity_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/r2f_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2f_pytest.log
tail -n 5 gpurun_out/r2f_pytest.log
bench() {  # name lib steps warmup
  B200MEL_LIB=$2 timeout 300 python bench.py --steps $3 --warmup $4 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2f_bench_$1.json 2> gpurun_out/r2f_bench_$1.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2f_bench_$1.json")); print("$1: %.1f us  frac %.4f  host %.1f us  %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["roofline"]["host_enqueue_us_per_step"], d["clocks"]))
except Exception as e: print("$1: bench failed", e)
PY
}
bench r1 $V/libr1.so 100 10
bench g3k1 "" 100 10
for v in g3k2 g2k1 g2k2 g2k3 g2k4 g2k5 g2s9k3; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -x -q -k "config2 or libri" > gpurun_out/r2f_${v}_pytest.log 2>&1
  echo "$v pytest: $(tail -n 1 gpurun_out/r2f_${v}_pytest.log)"
  bench $v $V/lib$v.so 100 10
done
B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > gpurun_out/r2f_trace.txt 2>&1; cat gpurun_out/r2f_trace.txt
