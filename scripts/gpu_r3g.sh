#!/bin/bash
# r3g: where do the ~6 k cycles between the dependency wait and the first tile go?  (stamps inside the tile-list pass)
mkdir -p gpurun_out
V=$PWD/experiments/_variants
B200MEL_LIB=$V/libtrace.so timeout 120 python experiments/trace_roles.py > gpurun_out/r3g_trace.txt 2>&1
sed -n 1,12p gpurun_out/r3g_trace.txt
