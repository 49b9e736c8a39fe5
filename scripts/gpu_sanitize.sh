#!/bin/bash
# One compute-sanitizer pass (ONE tool per gpurun call) over a small invocation of every kernel through the C ABI.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_sanitize.sh memcheck r2'
# The same command first runs WITHOUT the tool; the tool pass only starts when that exited 0.
TOOL=${1:-memcheck}; TAG=${2:-r2}
mkdir -p gpurun_out
O=gpurun_out/${TAG}_sanitize_${TOOL}
CMD="python experiments/sanitize_driver.py"
timeout 600 $CMD > $O.plain.log 2>&1
rc=$?; echo "plain exit $rc" | tee -a $O.plain.log
[ $rc -eq 0 ] || { tail -n 20 $O.plain.log; exit $rc; }
timeout 1300 compute-sanitizer --tool $TOOL --print-limit 40 --error-exitcode 9 $CMD > $O.log 2>&1
echo "sanitizer exit $?" | tee -a $O.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Invalid|Uninitialized|hazard|sanitizer exit|ok:" $O.log | head -60
