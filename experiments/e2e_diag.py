import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import bench
from speech_recognition_b200 import FeatureFrontend
mode = sys.argv[1]
fe = FeatureFrontend("log-mel-spectrogram", **bench.CFG)
B, L, T = 256, 240000, 1499
if mode in ("gauss", "gauss_resident"):
    ins = [torch.from_numpy(bench.synth_batch(5000 + i)).pin_memory() for i in range(2)]
else:
    rng = np.random.default_rng(0)
    ins = [torch.from_numpy(rng.integers(-3000, 3000, (B, L), dtype=np.int16)).pin_memory() for _ in range(2)]
outs = [torch.empty((B, T, 80, 1), dtype=torch.float32).pin_memory() for _ in range(2)]
if mode == "gauss_resident":
    d = [x.cuda() for x in ins]
    o = torch.empty((B, T, 80, 1), device="cuda")
    for i in range(200):
        fe(d[i % 2], out=o)
    torch.cuda.synchronize()
for i in range(3):
    fe.forward_host(ins[i % 2], out=outs[i % 2])
n = 20
t0 = time.perf_counter()
for i in range(n):
    fe.forward_host(ins[i % 2], out=outs[i % 2])
print(mode, "sync %.3f ms/step" % ((time.perf_counter() - t0) / n * 1e3))
