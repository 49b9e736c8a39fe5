"""Developer probe: what does the host side give when N GPUs copy at once?

Launched like the bench (python, or torchrun for N > 1; gloo barrier only).  Every rank copies 122.9 MB up
and 122.8 MB down (the e2e step of BASELINE config 2) between pinned host memory and its GPU, alone
directions and both at once, all ranks started together; rank 0 prints one JSON line with the per-rank
GB/s (min / median / max over ranks) for each pattern.  Options probe the usual suspects:
  --bind      pin the process to a slice of the host cores (cores / world each) before allocating
  --wc        write-combined pinned memory for the upload buffer (cudaHostAllocWriteCombined)
  --chunks K  cut each direction into K cudaMemcpyAsync calls
"""
import argparse
import ctypes
import json
import os
import time

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--bind", action="store_true")
ap.add_argument("--wc", action="store_true")
ap.add_argument("--chunks", type=int, default=1)
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
if args.bind:
    cores = sorted(os.sched_getaffinity(0))
    per = max(1, len(cores) // world)
    os.sched_setaffinity(0, cores[rank * per:(rank + 1) * per] or cores)
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("gloo")

NU, ND = 122880000, 122798080
cudart = ctypes.CDLL("libcudart.so.12") if args.wc else None


def pinned(n, wc):
    if not wc:
        return torch.empty(n, dtype=torch.uint8).pin_memory(), None
    p = ctypes.c_void_p()
    rc = cudart.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n), ctypes.c_uint(0x04))   # cudaHostAllocWriteCombined
    assert rc == 0, rc
    return None, p


hu_t, hu_p = pinned(NU, args.wc)
hd_t, _ = pinned(ND, False)
du = torch.empty(NU, dtype=torch.uint8, device="cuda")
dd = torch.empty(ND, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]


def copy(dst, src, n, kind, stream):
    k = args.chunks
    step = (n + k - 1) // k
    for o in range(0, n, step):
        m = min(step, n - o)
        rc = rt.cudaMemcpyAsync(dst + o, src + o, m, kind, ctypes.c_void_p(stream.cuda_stream))
        assert rc == 0, rc


hu_addr = hu_p.value if hu_p is not None else hu_t.data_ptr()


def run(up, down, reps):
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t = time.perf_counter()
    for _ in range(reps):
        if up:
            copy(du.data_ptr(), hu_addr, NU, 1, s1)
        if down:
            copy(hd_t.data_ptr(), dd.data_ptr(), ND, 2, s2)
        torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / reps
    return dt


res = {}
run(True, True, 2)
for name, (u, d) in {"h2d": (True, False), "d2h": (False, True), "both": (True, True)}.items():
    dt = run(u, d, args.reps)
    gbs = ((NU if u else 0) + (ND if d else 0)) / dt / 1e9
    t = torch.tensor([gbs, dt * 1e3], dtype=torch.float64)
    if dist is not None:
        allv = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
    else:
        allv = [t]
    g = sorted(float(v[0]) for v in allv)
    ms = sorted(float(v[1]) for v in allv)
    res[name] = {"GBps_per_rank_min_med_max": [round(g[0], 1), round(g[len(g) // 2], 1), round(g[-1], 1)],
                 "GBps_total": round(sum(g), 1), "ms_max": round(ms[-1], 2)}
if rank == 0:
    print(json.dumps({"probe": "pcie_concurrent", "n_gpus": world, "bind": args.bind, "wc": args.wc,
                      "chunks": args.chunks, "affinity_cores": len(os.sched_getaffinity(0)), **res}), flush=True)
if dist is not None:
    dist.barrier()
    dist.destroy_process_group()
