#!/bin/bash
# On the GPU box: parity-test + bench every variant in experiments/_variants.  Usage: run_variants.sh <tag> [names...]
TAG=$1; shift
mkdir -p gpurun_out
for so in "$@"; do
  L=$PWD/experiments/_variants/lib${so}.so
  B200MEL_LIB=$L timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tc or golden or libri" > gpurun_out/${TAG}_${so}_pytest.log 2>&1
  echo "$so pytest: $(tail -n 1 gpurun_out/${TAG}_${so}_pytest.log)"
  B200MEL_LIB=$L timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 2 > gpurun_out/${TAG}_${so}_bench.json 2> gpurun_out/${TAG}_${so}_bench.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_${so}_bench.json")); print("${so}: %.3e audio-s/s  %.1f us  frac %.3f" % (d["value"], d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print("${so}: bench failed", e)
PY
done
