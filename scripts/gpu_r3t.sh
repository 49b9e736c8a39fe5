#!/bin/bash
# r3l (re-run as r3t after the generic-kernel change): last check of the committed tree: GPU tests, smoke, bench
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > $O/r3t_pytest.log 2>&1; echo "pytest exit $?" >> $O/r3t_pytest.log; tail -n 4 $O/r3t_pytest.log
timeout 300 python __graft_entry__.py smoke > $O/r3t_smoke.log 2>&1; echo "smoke exit $?"
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r3t_bench_20.json 2> $O/r3t_bench_20.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('$O/r3t_bench_20.json')); print('%.2f us %.3e frac %.4f e2e %.3e cpu %s launches %d' % (d['ms_per_step']*1e3, d['value'], d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'], d['gpu_launches']))"
