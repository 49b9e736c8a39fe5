"""Developer tool: per-launch time of the TC kernel vs batch size (fixed cost vs per-tile cost)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_recognition_b200 import FeatureFrontend
cfg = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
           lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
fe = FeatureFrontend("log-mel-spectrogram", **cfg, device="cuda:0")
for B in (37, 74, 148, 256, 512, 1024):
    xs = [torch.randint(-3000, 3000, (B, 240000), dtype=torch.int16, device="cuda") for _ in range(3)]
    outs = [torch.empty((B, 1499, 80, 1), dtype=torch.float32, device="cuda") for _ in range(3)]
    lens = torch.full((B,), 240000, dtype=torch.int32, device="cuda")
    for i in range(5):
        fe(xs[i % 3], lengths=lens, out=outs[i % 3], kernel="tc")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 50
    for i in range(n):
        fe(xs[i % 3], lengths=lens, out=outs[i % 3], kernel="tc")
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    tiles = B * 12
    print("B=%4d tiles=%5d rounds=%.2f  %.1f us/launch  %.2f us/round  %.2fM audio-s/s" % (B, tiles, tiles / 148, us, us / (tiles / 148), B * 15 / us))
