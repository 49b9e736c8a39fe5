#!/bin/bash
# r2j: state check of the restored tree: GPU tests, 20-step bench, ncu launch list + one full capture of the
# headline kernel, reference arm, then one compute-sanitizer memcheck pass over experiments/sanitize_driver.py
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $O/r2j_smi.csv 2>&1
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 --durations=8 > $O/r2j_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2j_pytest.log
tail -n 14 $O/r2j_pytest.log
timeout 300 python __graft_entry__.py smoke > $O/r2j_smoke.log 2>&1; echo "smoke exit $?" | tee -a $O/r2j_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r2j_bench_20.json 2> $O/r2j_bench_20.err; echo "bench exit $?"; cut -c1-1200 $O/r2j_bench_20.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2j_reference_arm.json 2> $O/r2j_reference_arm.err; echo "ref exit $?"; cut -c1-800 $O/r2j_reference_arm.json
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
timeout 300 $CMD > $O/r2j_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/r2j_tc_launches.csv $CMD > $O/r2j_ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_frontend -s 5 -c 1 -f -o $O/r2j_prof $CMD > $O/r2j_ncu_full.log 2>&1
echo "ncu full exit $?"
bash scripts/gpu_sanitize.sh memcheck r2j
ls -la $O | tail -n 30
