#!/bin/bash
# r3m: ncu captures of the round's other kernels: MFCC second stage (f16 mma.sync) and the float32-input tensor-core kernel
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --config 4k --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 2"
timeout 300 $CMD > $O/r3m_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mfcc_post_mma -s 6 -c 1 -f -o $O/r3m_mfcc $CMD > $O/r3m_ncu_mfcc.log 2>&1
echo "ncu mfcc exit $?"
cat > /tmp/f32run.py <<'PY'
import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
from speech_recognition_b200 import FeatureFrontend
fe = FeatureFrontend("log-mel-spectrogram", 16000, 320, 160, 320, 80, None, 80.0, 7600.0)
rng = np.random.default_rng(0)
x = np.clip(np.rint(np.clip(rng.standard_normal((256, 240000), dtype=np.float32) * 0.1, -1, 1) * 32768), -32768, 32767).astype(np.int16)
xf = torch.from_numpy(x).cuda().to(torch.float32) / 32768.0
out = torch.empty((256, 1499, 80, 1), device="cuda")
for _ in range(6):
    fe(xf, out=out)
torch.cuda.synchronize()
print("ok")
PY
timeout 300 python /tmp/f32run.py > $O/r3m_f32_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_frontend -s 3 -c 1 -f -o $O/r3m_f32 python /tmp/f32run.py > $O/r3m_ncu_f32.log 2>&1
echo "ncu f32 exit $?"
