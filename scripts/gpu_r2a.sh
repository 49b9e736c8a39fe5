#!/bin/bash
# r2a: the new full-size parity tests on the round-1 kernel + where the 20-step bench loses time
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_smi.csv 2>&1
python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/r2a_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2a_pytest.log
tail -n 30 gpurun_out/r2a_pytest.log
for i in 1 2 3; do python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2a_bench20_$i.json 2> gpurun_out/r2a_bench20_$i.err; python -c "
import json;d=json.load(open('gpurun_out/r2a_bench20_$i.json'));print('20-step',d['ms_per_step'],d['roofline']['frac'],d['roofline']['host_enqueue_us_per_step'],d['clocks'])"; done
python bench.py --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 2 > gpurun_out/r2a_bench200.json 2> gpurun_out/r2a_bench200.err; python -c "
import json;d=json.load(open('gpurun_out/r2a_bench200.json'));print('200-step',d['ms_per_step'],d['roofline']['frac'],d['roofline']['host_enqueue_us_per_step'],d['clocks'])"
