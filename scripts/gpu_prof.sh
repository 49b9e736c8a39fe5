#!/bin/bash
# On the GPU box: bench (plain), then one ncu --set full capture of the TC kernel.  Usage: gpu_prof.sh <tag>
TAG=${1:-p}
mkdir -p gpurun_out
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && tail -c 600 gpurun_out/${TAG}_plain.log | grep -o '"value": [0-9.e+]*' | head -1
ncu --set full --clock-control none --import-source on -k regex:tc_frontend -s 5 -c 1 -f -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
