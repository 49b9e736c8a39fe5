"""e2e through host buffers: synchronous call vs submit / wait on two slots, per-step wall times."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_recognition_b200 import FeatureFrontend
fe = FeatureFrontend("log-mel-spectrogram", 16000, 320, 160, 320, 80, None, 80.0, 7600.0)
B, L = 256, 240000
rng = np.random.default_rng(0)
ins = [torch.from_numpy(rng.integers(-3000, 3000, (B, L), dtype=np.int16)).pin_memory() for _ in range(2)]
outs = [torch.empty((B, 1499, 80, 1), dtype=torch.float32).pin_memory() for _ in range(2)]
for i in range(3):
    fe.forward_host(ins[i % 2], out=outs[i % 2])
n = 20
t0 = time.perf_counter()
for i in range(n):
    fe.forward_host(ins[i % 2], out=outs[i % 2])
t_sync = (time.perf_counter() - t0) / n
t0 = time.perf_counter()
stamps = []
fe.forward_host_submit(ins[0], out=outs[0], slot=0)
for i in range(n):
    a = time.perf_counter()
    if i + 1 < n:
        fe.forward_host_submit(ins[(i + 1) % 2], out=outs[(i + 1) % 2], slot=(i + 1) % 2)
    b = time.perf_counter()
    fe.forward_host_wait(i % 2)
    c = time.perf_counter()
    stamps.append((b - a, c - b))
t_pipe = (time.perf_counter() - t0) / n
print("sync %.3f ms/step (%.2f M audio-s/s)   pipelined %.3f ms/step (%.2f M audio-s/s)" % (
    t_sync * 1e3, 3840 / t_sync / 1e6, t_pipe * 1e3, 3840 / t_pipe / 1e6))
print("submit / wait ms per step:", " ".join("%.2f/%.2f" % (s * 1e3, w * 1e3) for s, w in stamps[:8]))
