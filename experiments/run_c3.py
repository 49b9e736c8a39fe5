"""Config 2 + delta (C = 3) in a loop: the target of an ncu capture of tc_frontend_kernel<3, *>."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_recognition_b200 import FeatureFrontend
fe = FeatureFrontend("log-mel-spectrogram", 16000, 320, 160, 320, 80, None, 80.0, 7600.0, use_delta_accelerate=True)
rng = np.random.default_rng(0)
pcm = [torch.from_numpy(rng.integers(-3000, 3000, (256, 240000), dtype=np.int16)).cuda() for _ in range(3)]
outs = [torch.empty((256, 1499, 80, 3), dtype=torch.float32, device="cuda") for _ in range(3)]
for i in range(12):
    fe(pcm[i % 3], out=outs[i % 3])
torch.cuda.synchronize()
print("ok")
