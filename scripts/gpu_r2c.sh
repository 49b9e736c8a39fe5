#!/bin/bash
# r2c: restored 3-group register producer; register splits; role trace
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_fullsize_gpu.py tests/test_parity_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/r2c_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2c_pytest.log
tail -n 5 gpurun_out/r2c_pytest.log
bench() {  # name lib steps warmup
  B200MEL_LIB=$2 timeout 300 python bench.py --steps $3 --warmup $4 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2c_bench_$1.json 2> gpurun_out/r2c_bench_$1.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2c_bench_$1.json")); print("$1: %.1f us  frac %.4f  host %.1f us  %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["roofline"]["host_enqueue_us_per_step"], d["clocks"]))
except Exception as e: print("$1: bench failed", e)
PY
}
V=$PWD/experiments/_variants
bench r1 $V/libr1.so 100 10
bench new "" 100 10
for v in e88p104 e80p120; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -x -q -k "config2 or libri" > gpurun_out/r2c_${v}_pytest.log 2>&1
  echo "$v pytest: $(tail -n 1 gpurun_out/r2c_${v}_pytest.log)"
  bench $v $V/lib$v.so 100 10
done
B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > gpurun_out/r2c_trace.txt 2>&1; cat gpurun_out/r2c_trace.txt
bench new20 "" 20 3
