#!/bin/bash
# r3q: flakiness check - the GPU suite three times in a row, the full-size parity file five more times
mkdir -p gpurun_out
O=gpurun_out
for i in 1 2 3; do
  timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > $O/r3q_pytest_$i.log 2>&1; echo "run $i: exit $? $(tail -n 1 $O/r3q_pytest_$i.log)"
done
for i in 1 2 3 4 5; do
  timeout 600 python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -q -x -p no:cacheprovider > $O/r3q_full_$i.log 2>&1; echo "fullsize $i: exit $? $(tail -n 1 $O/r3q_full_$i.log)"
done
