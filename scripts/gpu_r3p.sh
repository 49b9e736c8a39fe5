#!/bin/bash
# r3p: FeatureFrontend.bind test
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_host_pipeline.py -m gpu -q -x > gpurun_out/r3p_pytest.log 2>&1; tail -n 12 gpurun_out/r3p_pytest.log
