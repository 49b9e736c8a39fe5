#!/bin/bash
# Build a kernel variant: experiments/build_variant.sh <name> [-DFLAG ...]  ->  experiments/_variants/lib<name>.so
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
mkdir -p experiments/_variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -cudart static \
  -DB200MEL_WITH_TC "$@" -Xptxas -v -o experiments/_variants/lib${NAME}.so speech_recognition_b200/csrc/b200mel_api.cu 2>&1 \
  | grep -A2 "tc_frontend_kernelILi1" | grep -E "Used|spill" || true
