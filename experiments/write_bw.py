"""HBM probe: pure write, pure read-ish (sum), copy, and a 1-read : 3-write mix (the C = 3 traffic shape)."""
import torch
x = torch.empty(1 << 29, dtype=torch.float32, device="cuda")  # 2 GB
y = torch.empty(1 << 29, dtype=torch.float32, device="cuda")
def t(fn, n=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e-3
s = t(lambda: x.zero_()); print("memset  2 GB: %.0f GB/s" % (x.numel() * 4 / s / 1e9))
s = t(lambda: y.copy_(x)); print("copy 2+2 GB: %.0f GB/s" % (2 * x.numel() * 4 / s / 1e9))
q = x[: 1 << 27]; o = y[: 3 << 27].view(3, -1)
s = t(lambda: torch.add(q.unsqueeze(0), 1.0, out=o) if False else o.copy_(q.unsqueeze(0).expand(3, -1))); print("1r:3w mix: %.0f GB/s" % (4 * q.numel() * 4 / s / 1e9))
