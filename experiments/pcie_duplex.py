"""Developer probe: what the box's PCIe link gives for 123 MB up + 123 MB down at once (the e2e bound)."""
import torch, time
n = 122880000
hu = torch.empty(n, dtype=torch.uint8).pin_memory(); hd = torch.empty(n, dtype=torch.uint8).pin_memory()
du = torch.empty(n, dtype=torch.uint8, device="cuda"); dd = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(up, down, reps=10):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1): du.copy_(hu, non_blocking=True)
        if down:
            with torch.cuda.stream(s2): hd.copy_(dd, non_blocking=True)
        torch.cuda.synchronize()
    return (time.perf_counter() - t) / reps * 1e3
for _ in range(2): run(True, True, 2)
print("H2D alone %.2f ms  D2H alone %.2f ms  both at once %.2f ms (123 MB each)" % (run(True, False), run(False, True), run(True, True)))
