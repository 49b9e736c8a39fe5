#!/bin/bash
# r3c: the full ncu capture of r3b again (the skip count was larger than the number of launches under ncu)
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
timeout 300 $CMD > $O/r3c_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_frontend -s 12 -c 1 -f -o $O/r3b_prof $CMD > $O/r3b_ncu_full.log 2>&1
echo "ncu full exit $?"; tail -3 $O/r3b_ncu_full.log | cut -c1-200
