#!/bin/bash
# r2p: ncu capture of the MFCC second stage
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --config 4k --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 2"
timeout 300 $CMD > $O/r2p_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mfcc_post_mma -s 4 -c 1 -f -o $O/r2p_mfcc $CMD > $O/r2p_ncu.log 2>&1
echo "ncu exit $?"
