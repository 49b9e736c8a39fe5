#!/bin/bash
# r3u: ncu launch lists of the two-kernel MFCC step (config 4k) and the SpecAugment + delta step (config 3d)
mkdir -p gpurun_out
O=gpurun_out
for c in 4k 3d; do
  CMD="python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 2 --spin-up-ms 0"
  timeout 300 $CMD > $O/r3u_plain_$c.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $O/r3u_launches_$c.csv $CMD > $O/r3u_ncu_$c.log 2>&1
  echo "config $c ncu exit $?"
done
