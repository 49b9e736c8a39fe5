#!/bin/bash
# r2g: eps folded into the accumulators; poll back-off, register split and column split variants; the full
# GPU test-suite on the default build; first bench --config lines
mkdir -p gpurun_out
V=$PWD/experiments/_variants
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r2g_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2g_pytest.log
tail -n 5 gpurun_out/r2g_pytest.log
bench() {  # name lib steps warmup
  B200MEL_LIB=$2 timeout 300 python bench.py --steps $3 --warmup $4 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2g_bench_$1.json 2> gpurun_out/r2g_bench_$1.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2g_bench_$1.json")); print("$1: %.1f us  frac %.4f  host %.1f us  %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["roofline"]["host_enqueue_us_per_step"], d["clocks"]))
except Exception as e: print("$1: bench failed", e)
PY
}
bench r1 $V/libr1.so 100 10
bench new "" 100 10
for v in sleep32 sleep100 e80p120 s5_11 s4_10 s5_9 s4_9; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -x -q -k "config2 or libri" > gpurun_out/r2g_${v}_pytest.log 2>&1
  echo "$v pytest: $(tail -n 1 gpurun_out/r2g_${v}_pytest.log)"
  bench $v $V/lib$v.so 100 10
done
bench new2 "" 100 10
bench new20 "" 20 3
for c in 3 3d 4k 5; do
  timeout 600 python bench.py --config $c --steps 50 --warmup 5 --e2e-steps 4 > gpurun_out/r2g_config_$c.json 2> gpurun_out/r2g_config_$c.err; echo "config $c exit $?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2g_config_$c.json")); print("config $c: %.1f us  %.3e audio-s/s frac %.4f cpu %s id %s" % (d["ms_per_step"]*1e3, d["value"], d["roofline"]["frac"], d.get("cpu_baseline"), d.get("mask_identity")))
except Exception as e: print("config $c failed", e)
PY
done
