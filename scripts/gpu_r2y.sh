#!/bin/bash
# r2y: role trace of the C = 1 kernel with and without SpecAugment (what do the masks cost per tile?)
mkdir -p gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > gpurun_out/r2y_trace_plain.txt 2>&1
TRACE_SA=1 B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > gpurun_out/r2y_trace_sa.txt 2>&1
echo "== plain"; sed -n 1,3p gpurun_out/r2y_trace_plain.txt; sed -n '/tile 10/,/tile 20/p' gpurun_out/r2y_trace_plain.txt
echo "== SpecAugment"; sed -n 1,8p gpurun_out/r2y_trace_sa.txt; sed -n '/tile 10/,/tile 20/p' gpurun_out/r2y_trace_sa.txt
