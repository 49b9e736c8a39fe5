#!/bin/bash
# r3b: final state of the round: GPU tests, smoke, the driver's bench form, reference arm, launch list and one
# full ncu capture of the headline kernel, config lines
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $O/r3b_smi.csv 2>&1
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 --durations=5 > $O/r3b_pytest.log 2>&1; echo "pytest exit $?" >> $O/r3b_pytest.log
tail -n 10 $O/r3b_pytest.log
timeout 300 python __graft_entry__.py smoke > $O/r3b_smoke.log 2>&1; echo "smoke exit $?" | tee -a $O/r3b_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r3b_bench_20.json 2> $O/r3b_bench_20.err; echo "bench exit $?"; cut -c1-300 $O/r3b_bench_20.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r3b_reference_arm.json 2> $O/r3b_reference_arm.err; echo "ref exit $?"
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
timeout 300 $CMD > $O/r3b_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r3b_tc_launches.csv $CMD > $O/r3b_ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_frontend -s 120 -c 1 -f -o $O/r3b_prof $CMD > $O/r3b_ncu_full.log 2>&1
echo "ncu full exit $?"
for c in 3 3d 4k; do
  timeout 600 python bench.py --config $c --steps 20 --warmup 3 > $O/r3b_config_$c.json 2> $O/r3b_config_$c.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r3b_config_$c.json")); print("config $c: %.1f us  %.3e audio-s/s frac %.4f cpu %s" % (d["ms_per_step"]*1e3, d["value"], d["roofline"]["frac"], (d.get("cpu_baseline") or {}).get("value")))
except Exception as e: print("config $c failed", e)
PY
done
timeout 600 python experiments/aux_bench.py > $O/r3b_aux.json 2> $O/r3b_aux.err; echo "aux exit $?"; cat $O/r3b_aux.json | cut -c1-600
