#!/bin/bash
# r2z: SpecAugment time warp (row N4): GPU tests
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_time_warp.py tests/test_data_config.py -m gpu -q -x --timeout 600 > $O/r2z_pytest.log 2>&1; echo "pytest exit $?" >> $O/r2z_pytest.log
tail -n 30 $O/r2z_pytest.log
