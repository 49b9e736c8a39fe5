#!/bin/bash
# r2e: ncu capture of the three-group kernel (rolled loop) and of the round-1 kernel for comparison
mkdir -p gpurun_out
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
$CMD > gpurun_out/r2e_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:tc_frontend -s 8 -c 2 -f -o gpurun_out/r2e_new $CMD > gpurun_out/r2e_ncu_new.log 2>&1
echo "ncu new exit $?"; tail -n 3 gpurun_out/r2e_plain.log
export B200MEL_LIB=$PWD/experiments/_variants/libr1.so
$CMD > gpurun_out/r2e_plain_r1.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:tc_frontend -s 8 -c 2 -f -o gpurun_out/r2e_r1 $CMD > gpurun_out/r2e_ncu_r1.log 2>&1
echo "ncu r1 exit $?"
ls -la gpurun_out/*.ncu-rep
