#!/bin/bash
# Build a kernel variant: experiments/build_variant.sh <name> [-DFLAG ...]  ->  experiments/_variants/lib<name>.so
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
mkdir -p experiments/_variants
# SPLIT="a,b": another column split of the three epilogue groups (regenerated tables)
if [ -n "$SPLIT" ]; then
  A=${SPLIT%,*}; B=${SPLIT#*,}
  g++ -O1 -std=c++17 -DTC_SPLIT_A=$A -DTC_SPLIT_B=$B -o experiments/_variants/gen_$NAME speech_recognition_b200/csrc/gen_tc_tables.cc
  experiments/_variants/gen_$NAME > experiments/_variants/tables_$NAME.inc
  rm -f experiments/_variants/gen_$NAME
  set -- "$@" "-DTC_TABLES_INC=\"$PWD/experiments/_variants/tables_$NAME.inc\""
fi
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -cudart static \
  -DB200MEL_WITH_TC "$@" -Xptxas -v -o experiments/_variants/lib${NAME}.so speech_recognition_b200/csrc/b200mel_api.cu 2>&1 \
  | grep -A2 "tc_frontend_kernelILi1" | grep -E "Used|spill" || true
