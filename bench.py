#!/usr/bin/env python
"""bench.py - headline benchmark of the log-mel front-end (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this framework (CUDA, sm_100a)
    python bench.py --impl reference --gpus N ...            # the reference's CPU path (oracle port)

One "step" = one pass of the hot path over one synthetic batch of BASELINE config 2:
256 utterances x 15 s of 16 kHz int16 PCM -> log-mel [256, 1499, 80, 1] float32 (libri_config.yml
parameters).  N > 1 (torchrun, one rank per GPU): every rank runs the same per-GPU workload on its
own utterances (weak scaling, no collective on the data path - SURVEY 8e); `value` = audio-seconds
all ranks processed / max-over-ranks device time.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      achieved algorithmic GB/s of the fused kernel (bytes from DESIGN.md / SURVEY 8d)
                against the measured HBM copy peak in MEASURED_PEAKS.json
  cpu_baseline  the oracle's C restatement of the reference's TF CPU path on this box's host cores
  e2e           same metric through the host-buffer C-ABI call (H2D + kernel + D2H every step)
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "audio_seconds_per_second_log_mel"
UNIT = "audio-s/s"
CFG = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
           lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
B, SECONDS, SR = 256, 15, 16000
L = SECONDS * SR                       # 240000 samples
T = 1 + (L - 320) // 160               # 1499 frames
AUDIO_S_PER_STEP = B * SECONDS         # 3840 audio-seconds per batch per GPU
ALGO_BYTES_PER_STEP = B * (L * 2 + T * 80 * 4)   # int16 in + f32 out = 245.68 MB
RING = 5                               # distinct batches cycled through: 5 x 245.7 MB >> 126 MB L2
WORKLOAD = "synthetic 16 kHz LibriSpeech-shaped batch: 256 utterances x 15 s, 80 mel bins (BASELINE config 2)"


def synth_batch(seed, n_utt=B):
    """BASELINE config 2 audio: clip(N(0, 0.1^2), -1, 1) quantised to int16."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n_utt, L), dtype=np.float32) * np.float32(0.1)
    np.clip(x, -1.0, 1.0, out=x)
    return np.clip(np.rint(x * 32768.0), -32768, 32767).astype(np.int16)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if any."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("dram_bytes_per_launch")
    return None


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML.  The thread is started BEFORE the warm-up steps
    (NVML initialisation and thread start-up take milliseconds: between the warm-up and the timed region
    they would leave the GPU idle long enough to drop its clocks, which a 2 ms timed region then pays
    for); `mark()` / `stop()` bracket the timed region and only samples taken inside it are reported."""

    PERIOD_S = 0.0005

    def __init__(self, index):
        self.samples, self.max_mhz = [], None      # (time, sm_mhz, reason mask)
        self._stop = threading.Event()
        self._thread = None
        self.index = index
        self.t0 = self.t1 = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, mask))
            except Exception:
                pass
            time.sleep(self.PERIOD_S)

    def start(self):
        if self.nv is not None and not os.environ.get("B200MEL_BENCH_NO_SAMPLER"):
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def mark(self):
        self.t0 = time.perf_counter()

    def stop(self):
        self.t1 = time.perf_counter()
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        nv = self.nv
        names = {} if nv is None else {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        inside = [s for s in self.samples if self.t0 is not None and self.t0 <= s[0] <= self.t1]
        if not inside and self.samples and self.t0 is not None:   # (region shorter than one period: the nearest sample)
            inside = [min(self.samples, key=lambda s: abs(s[0] - 0.5 * (self.t0 + self.t1)))]
        reasons = set()
        for _, _, mask in inside:
            for bit, name in names.items():
                if mask & bit:
                    reasons.add(name)
        med = float(np.median([s[1] for s in inside])) if inside else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(reasons), "samples": len(inside)}


_JSON_FD = []  # the real stdout, saved while fd 1 points at stderr (multi-rank runs)


def emit_line(line):
    text = json.dumps(line) + "\n"
    if _JSON_FD:
        sys.stdout.flush()
        os.write(_JSON_FD[0], text.encode())
    else:
        sys.stdout.write(text)
        sys.stdout.flush()


def cpu_reference_run(n_utt, steps, warmup, threads=0):
    """Time the oracle's C restatement of the reference's TF CPU path (kind = "port": TensorFlow
    cannot be installed here) on `n_utt` utterances of the same workload per step."""
    from oracle.ref_c import RefPlan

    if threads <= 0:
        # every host core this process may use - explicitly, because torchrun exports
        # OMP_NUM_THREADS=1 and the OpenMP default would then time a single thread
        try:
            threads = len(os.sched_getaffinity(0))
        except AttributeError:
            threads = os.cpu_count() or 1
    plan = RefPlan("log-mel-spectrogram", **CFG)
    pcm = synth_batch(1234, n_utt)
    out = np.empty((n_utt, T, 80, 1), np.float32)
    for _ in range(warmup):
        plan(pcm, out=out, n_threads=threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        plan(pcm, out=out, n_threads=threads)
    dt = time.perf_counter() - t0
    return n_utt * SECONDS * steps / dt, dt / steps, plan.threads_used


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_utt = 64
    value, per_step, cores = cpu_reference_run(n_utt, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "step": "%d utterances x 15 s per step (bounded sample of the batch)" % n_utt,
                   "note": "TensorFlow (the reference's engine) is not installable here; this is the oracle's C "
                           "restatement of data.py:145-189 on all host cores"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d x 15 s utterances per step, %d steps" % (n_utt, args.steps)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "tc"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = min(steps, 20)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch

    from speech_recognition_b200 import FeatureFrontend

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the framework has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        # libnccl prints its version banner on fd 1 at the first communicator: keep rank 0's stdout to the
        # one JSON line by pointing fd 1 at stderr until the line is printed
        sys.stdout.flush()
        _JSON_FD.append(os.dup(1))
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)

    fe = FeatureFrontend("log-mel-spectrogram", **CFG, device=dev)
    warmup = max(args.warmup, 3)
    # ring of distinct batches, resident in HBM before the timed region starts
    ring_in = [torch.from_numpy(synth_batch(1000 * rank + i)).to(dev) for i in range(RING)]
    ring_out = [torch.empty((B, T, 80, 1), dtype=torch.float32, device=dev) for _ in range(RING)]
    stream = torch.cuda.current_stream(dev)
    d_lengths = torch.full((B,), L, dtype=torch.int32, device=dev)   # resident: no per-step H2D
    d_nframes = torch.empty((B,), dtype=torch.int32, device=dev)

    def step(i):
        fe(ring_in[i % RING], lengths=d_lengths, out=ring_out[i % RING], n_frames=d_nframes, kernel=args.kernel)

    # staging: every ring slot goes through the path once (first touch of the output pages, the kernel's
    # status word checked on real data) before the W warm-up steps
    status = torch.zeros((1,), dtype=torch.int32, device=dev)
    for i in range(RING):
        _, _, st_i = fe(ring_in[i], lengths=d_lengths, out=ring_out[i], kernel=args.kernel, return_status=True)
        status |= st_i
    torch.cuda.synchronize()
    if int(status.item()) != 0:
        raise SystemExit("bench.py: kernel status word %#x on the staged batches (0x100 = pipeline wait timed out)" % int(status.item()))
    sampler = ClockSampler(torch.cuda._get_nvml_device_index(dev) if hasattr(torch.cuda, "_get_nvml_device_index")
                           else local_rank)
    sampler.start()
    launches0 = None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(warmup):
        step(i)
    launches0 = fe.launch_count
    # nothing but the barrier + synchronize stands between the warm-up steps and the timed region
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.mark()
    ev0.record(stream)
    t_host0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    host_us_per_call = (time.perf_counter() - t_host0) * 1e6 / max(args.steps, 1)   # enqueue cost, GPU running behind
    ev1.record(stream)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    clocks = sampler.stop()
    launches = fe.launch_count - launches0
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * args.steps * AUDIO_S_PER_STEP / (ms_max * 1e-3)

    # ---- e2e: host buffers through the C ABI; H2D + kernel + D2H inside the timed region --------
    e2e_steps = args.e2e_steps or min(args.steps, 20)
    host_in = [torch.from_numpy(synth_batch(5000 + 10 * rank + i)).pin_memory() for i in range(2)]
    host_out = [torch.empty((B, T, 80, 1), dtype=torch.float32).pin_memory() for _ in range(2)]
    for i in range(2):   # untimed: staging buffers of both slots are allocated here, both forms run once
        fe.forward_host(host_in[i % 2], out=host_out[i % 2], kernel=args.kernel)
    for i in range(2):
        fe.forward_host_submit(host_in[i % 2], out=host_out[i % 2], kernel=args.kernel, slot=i % 2)
    for i in range(2):
        fe.forward_host_wait(i % 2)
    # (a) synchronous call per step: upload, kernels, download, return
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        fe.forward_host(host_in[i % 2], out=host_out[i % 2], kernel=args.kernel)
    e2e_sync_s = time.perf_counter() - t0
    # (b) the streaming form of the same call (submit / wait on two staging slots): step i+1 is submitted
    #     before step i is waited for, so its uploads overlap step i's downloads.  Every step's input is
    #     copied from pinned host memory and its result is read on the host inside the timed region.
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    checksum = 0.0
    fe.forward_host_submit(host_in[0], out=host_out[0], kernel=args.kernel, slot=0)
    for i in range(e2e_steps):
        if i + 1 < e2e_steps:
            fe.forward_host_submit(host_in[(i + 1) % 2], out=host_out[(i + 1) % 2], kernel=args.kernel, slot=(i + 1) % 2)
        fe.forward_host_wait(i % 2)
        checksum += float(host_out[i % 2][0, 0, 0, 0])   # the step's result is read on the host
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_stream_value = world * e2e_steps * AUDIO_S_PER_STEP / float(te.item())
    ts = torch.tensor([e2e_sync_s], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(ts, op=dist.ReduceOp.MAX)
    e2e_sync_value = world * e2e_steps * AUDIO_S_PER_STEP / float(ts.item())
    # both forms move every step's input up and its result down inside the timed region; the line's value is
    # the faster one on this box (the streaming form wins on 1-2 GPUs, the synchronous call when 8 ranks share
    # the host's memory system), both are listed
    e2e_value = max(e2e_stream_value, e2e_sync_value)
    e2e_form = "submit/wait on two staging slots" if e2e_stream_value >= e2e_sync_value else "synchronous call"

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    per_launch_s = ms * 1e-3 / max(launches, 1)          # rank 0's own launches
    achieved = ALGO_BYTES_PER_STEP / per_launch_s / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "input": "int16 PCM resident in HBM", "output": "[256,1499,80,1] f32",
                   "frame_length": 320, "frame_step": 160, "fft_length": 320, "num_mel_bins": 80,
                   "batch_per_gpu": B, "global_batch": B * world, "parallelism": "utterance-sharded x%d, no collective" % world,
                   "l2_policy": "ring of %d distinct batches (%.0f MB touched per cycle > 126 MB L2)"
                                % (RING, RING * ALGO_BYTES_PER_STEP / 1e6),
                   "kernel": args.kernel, "tc_supported": bool(fe.tc_supported)},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": ALGO_BYTES_PER_STEP,
                     "launch_us": per_launch_s * 1e6, "host_enqueue_us_per_step": host_us_per_call},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * L * 2, "d2h_bytes_per_step": B * T * 80 * 4,
                "steps": e2e_steps,
                "timer": "host wall clock around the loop of C-ABI calls, max over ranks",
                "form": e2e_form, "stream_value": e2e_stream_value, "sync_call_value": e2e_sync_value,
                "checksum": checksum},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        try:
            n_utt = 64
            v1, per_step, cores = cpu_reference_run(n_utt, 1, 1)
            reps = int(max(1, min(40, 10.0 / max(per_step * cores, 1e-3))))   # ~10 s of CPU work
            v, per_step, cores = cpu_reference_run(n_utt, reps, 0)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "%d x 15 s utterances x %d passes (oracle C restatement of the TF CPU "
                                              "path; TensorFlow not installable here)" % (n_utt, reps)}
        except Exception as e:  # the baseline is informational; never lose the GPU numbers to it
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": "failed: %r" % (e,)}
    emit_line(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
