#!/bin/bash
# Run on the GPU box (via gpurun): GPU tests, smoke, bench, then ncu launch list + one full capture.
# Usage: scripts/gpu_round.sh <tag> [kernel-regex]
TAG=${1:-r1}
KREGEX=${2:-generic_frontend}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > gpurun_out/${TAG}_smi.csv 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -n 25 gpurun_out/${TAG}_pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/${TAG}_smoke.log
tail -n 5 gpurun_out/${TAG}_smoke.log
python bench.py --steps 200 --warmup 20 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json; tail -n 5 gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>&1; cat gpurun_out/${TAG}_bench_ref.json
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:${KREGEX} -s 5 -c 2 -f -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
python experiments/config_sweep.py > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err; echo "config sweep exit $?"
python experiments/aux_bench.py > gpurun_out/${TAG}_aux.json 2> gpurun_out/${TAG}_aux.err; echo "aux bench exit $?"; cat gpurun_out/${TAG}_aux.json
ls -la gpurun_out | tail -n 20
