#!/bin/bash
# r2l: why is the 20-step bench slower per step than the 100-step one?  per-step events, warm-up length, spin-up
mkdir -p gpurun_out
O=gpurun_out
V=$PWD/experiments/_variants
run() { # name args...
  n=$1; shift
  B200MEL_LIB=$V/libspec.so timeout 300 python bench.py --no-cpu-baseline --e2e-steps 2 "$@" > $O/r2l_$n.json 2> $O/r2l_$n.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r2l_$n.json")); print("$n: %.2f us  frac %.4f clocks %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["clocks"])); print("   ", d.get("step_us"))
except Exception as e: print("$n failed", e)
PY
}
run a_20_3 --steps 20 --warmup 3 --trace-steps
run b_20_3 --steps 20 --warmup 3
run c_20_50 --steps 20 --warmup 50 --trace-steps
run d_20_3_spin --steps 20 --warmup 3 --spin-up-ms 50 --trace-steps
run e_20_3_spin --steps 20 --warmup 3 --spin-up-ms 50
run f_100_10 --steps 100 --warmup 10 --trace-steps
run g_20_3 --steps 20 --warmup 3
