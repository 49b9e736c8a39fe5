#!/bin/bash
# r2k: guard-band tests (default build) + kernel variants: reversed group order, later CFREE, speculative first tile
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
timeout 900 python -m pytest tests/test_guard_bands_gpu.py -m gpu -q -x --timeout 300 > $O/r2k_guard.log 2>&1; echo "guard exit $?" >> $O/r2k_guard.log
tail -n 25 $O/r2k_guard.log
bench() {  # name lib steps warmup
  B200MEL_LIB=$2 timeout 300 python bench.py --steps $3 --warmup $4 --no-cpu-baseline --e2e-steps 2 > $O/r2k_bench_$1.json 2> $O/r2k_bench_$1.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2k_bench_$1.json")); print("$1: %.2f us  frac %.4f  host %.1f us" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["roofline"]["host_enqueue_us_per_step"]))
except Exception as e: print("$1: bench failed", e)
PY
}
for v in base rev cf2 spec all base; do
  B200MEL_LIB=$V/lib$v.so timeout 300 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -x -q -k "config2 or libri or config5" > $O/r2k_${v}_pytest.log 2>&1
  echo "$v pytest: $(tail -n 1 $O/r2k_${v}_pytest.log)"
  bench ${v}_100 $V/lib$v.so 100 10
  bench ${v}_20 $V/lib$v.so 20 3
done
