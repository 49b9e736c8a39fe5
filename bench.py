#!/usr/bin/env python
"""bench.py - headline benchmark of the log-mel front-end (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this framework (CUDA, sm_100a)
    python bench.py --impl reference --gpus N ...            # the reference's CPU path (oracle port)
    python bench.py --config {2,3,3d,4k,5} ...               # the other BASELINE.json configs (default: 2)

Default (config 2, the configuration BASELINE.json's metric is quoted on): one "step" = one pass of the hot
path over one synthetic batch of 256 utterances x 15 s of 16 kHz int16 PCM -> log-mel [256, 1499, 80, 1]
float32 (libri_config.yml parameters).  N > 1 (torchrun, one rank per GPU): every rank runs the same
per-GPU workload on its own utterances (weak scaling, no collective on the data path - SURVEY 8e); `value`
= audio-seconds all ranks processed / max-over-ranks device time.

Other configs (same line format, `config.workload` names them):
  3    config 2 + SpecAugment masks (F=27, m_F=2, T=100, p=1.0, m_T=2; seed 43; draws keyed by the GLOBAL
       utterance id) -> [256,1499,80,1]; the line carries `mask_identity`: one fixed global batch of 256
       utterances sharded over the N ranks, per-utterance zero rows / columns gathered, hashed and checked
       against the oracle's draws - the sha1 must be the same at N = 1, 2, 4, 8
  3d   libri_config.yml as shipped: SpecAugment + delta / delta-delta -> [256,1499,80,3]
  4k   kspon_config.yml / clovacall_config.yml: MFCC-40 + delta -> [256,1499,40,3]
  5    10 h of variable-length utterances (durations clip(N(12.3, 4.5^2), 1, 20.48) s, default_rng(1)),
       batches of 256 in arrival order, ragged int16 in, [B,Tmax,80,3] out; every batch is split over the N
       ranks in contiguous utterance ranges balanced by samples (strong scaling); one step = one pass
       over the 10 h; only real (unpadded) audio seconds count

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      achieved algorithmic GB/s of the fused kernel (bytes from DESIGN.md / SURVEY 8d)
                against the measured HBM copy peak in MEASURED_PEAKS.json
  cpu_baseline  the oracle's C restatement of the reference's TF CPU path on this box's host cores, on
                a bounded sample of the same workload; the GPU result for that sample is compared with it
                on the way (`parity_max_abs_err`: the bench fails loudly if the stated tolerance is broken)
  e2e           same metric through the host-buffer C-ABI call (H2D + kernel + D2H every step)
"""
import argparse
import hashlib
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
SA = dict(enable=True, W=None, F=27, m_F=2, T=100, p=1.0, m_T=2)      # libri_config.yml:28-35
SEED = 43                                                              # train_config_sample.yml:17
B, SECONDS, SR = 256, 15, 16000
L = SECONDS * SR                       # 240000 samples
T = 1 + (L - 320) // 160               # 1499 frames
SPIN_UP_MS = 8.0                       # (--spin-up-ms) untimed passes before the W warm-up steps: see the note at its use
RING = 5                               # distinct batches cycled through: 5 x 245.7 MB >> 126 MB L2

CONFIGS = {
    "2": dict(ftype="log-mel-spectrogram", extra={}, delta=False, sa=False, F=80,
              workload="synthetic 16 kHz LibriSpeech-shaped batch: 256 utterances x 15 s, 80 mel bins (BASELINE config 2)"),
    "3": dict(ftype="log-mel-spectrogram", extra={}, delta=False, sa=True, F=80,
              workload="BASELINE config 3: config 2 + SpecAugment masking (F=27, m_F=2, T=100, p=1.0, m_T=2, seed 43, "
                       "draws keyed by global utterance id), 256 x 15 s"),
    "3d": dict(ftype="log-mel-spectrogram", extra={}, delta=True, sa=True, F=80,
               workload="BASELINE config 3 as libri_config.yml ships it: SpecAugment + delta/delta-delta, 256 x 15 s -> [256,1499,80,3]"),
    "4k": dict(ftype="mfcc", extra=dict(num_mfcc=40), delta=True, sa=False, F=40,
               workload="BASELINE config 4: kspon_config.yml / clovacall_config.yml (MFCC-40 + delta), 256 x 15 s -> [256,1499,40,3]"),
    "5": dict(ftype="log-mel-spectrogram", extra={}, delta=True, sa=False, F=80,
              workload="BASELINE config 5: 10 h of variable-length utterances (1-20.48 s), batches of 256 in arrival order, "
                       "ragged int16 in, [B,Tmax,80,3] out, batches split over the ranks by samples"),
}


def synth_batch(seed, n_utt=B):
    """BASELINE config 2 audio: clip(N(0, 0.1^2), -1, 1) quantised to int16."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n_utt, L), dtype=np.float32) * np.float32(0.1)
    np.clip(x, -1.0, 1.0, out=x)
    return np.clip(np.rint(x * 32768.0), -32768, 32767).astype(np.int16)


def config5_lengths():
    """durations clip(N(12.3, 4.5^2), 1.0, 20.48) s until 10 h, default_rng(1)."""
    rng = np.random.default_rng(1)
    durs = []
    while sum(durs) < 36000.0:
        durs.append(float(np.clip(rng.normal(12.3, 4.5), 1.0, 20.48)))
    return np.array([int(d * SR) for d in durs], dtype=np.int64)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(config):
    """dram bytes per launch of the dominant kernel from the committed ncu capture of THIS config, if any
    (a profiler figure cannot be taken inside a timed run; the file names the capture it comes from)."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        if config == "2":
            return d.get("dram_bytes_per_launch")
        return (d.get("configs") or {}).get(config)
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


def host_threads():
    # every host core this process may use - explicitly, because torchrun exports OMP_NUM_THREADS=1 and the
    # OpenMP default would then time a single thread
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def ref_plan(config):
    """The oracle's C restatement of the reference's CPU path for this config (test infrastructure: used here
    as the timed CPU baseline / reference arm and as the checker of the GPU result, never by the product)."""
    from oracle.ref_c import RefPlan

    spec = CONFIGS[config]
    return RefPlan(spec["ftype"], **CFG, **spec["extra"], use_delta_accelerate=spec["delta"])


def ref_sample(config, n_utt):
    """(pcm, lengths, offsets, audio_seconds) of the CPU sample: n_utt utterances of the workload."""
    if config == "5":
        lens = config5_lengths()[:n_utt]
        rng = np.random.default_rng(4242)
        flat = np.clip(np.rint(np.clip(rng.standard_normal(int(lens.sum()), dtype=np.float32) * np.float32(0.1), -1, 1) * 32768.0),
                       -32768, 32767).astype(np.int16)
        offs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        return flat, lens.astype(np.int32), offs, float(lens.sum()) / SR
    return synth_batch(1234, n_utt), None, None, float(n_utt * SECONDS)


def cpu_reference_run(config, n_utt, steps, warmup, threads=0):
    """Time the oracle's C restatement of the reference's TF CPU path (kind = "port": TensorFlow cannot be
    installed here) on `n_utt` utterances of the workload per step.  SpecAugment masking is not part of the
    CPU arm (a few multiplies per masked element next to an FFT per frame)."""
    threads = threads if threads > 0 else host_threads()
    plan = ref_plan(config)
    pcm, lens, offs, audio_s = ref_sample(config, n_utt)
    for _ in range(warmup):
        plan(pcm, lengths=lens, offsets=offs, n_threads=threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        plan(pcm, lengths=lens, offsets=offs, n_threads=threads)
    dt = time.perf_counter() - t0
    return audio_s * steps / dt, dt / steps, plan.threads_used


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_utt = B     # the full batch of the GPU arm's step (same config, same step)
    value, per_step, cores = cpu_reference_run(args.config, n_utt, args.steps, args.warmup)
    spec = CONFIGS[args.config]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.config == "5" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": spec["workload"], "config": args.config,
                   "step": "%d utterances per step%s" % (n_utt, " (one batch of the 10 h: a bounded sample)" if args.config == "5"
                                                         else " (the GPU arm's batch)"),
                   "note": "TensorFlow (the reference's engine) is not installable here; this is the oracle's C "
                           "restatement of data.py:122-328 on all host cores (OpenMP over utterances)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d utterances per step, %d steps" % (n_utt, args.steps)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
def parity_against_cpu(config, fe, dev, torch, n_utt, kernel):
    """The bounded CPU sample once more on the GPU, compared with the C restatement: max abs error of
    channel 0 (log-mel: log domain where the energy is above the stated floor; MFCC: coefficients), padding
    exactly zero.  Raises if the stated tolerance is broken.  SpecAugment is left out of this comparison
    (the masks have their own check: `mask_identity`)."""
    from speech_recognition_b200 import FeatureFrontend
    from tests.helpers import assert_logmel_close

    spec = CONFIGS[config]
    pcm, lens, offs, _ = ref_sample(config, n_utt)
    want, nf_want = ref_plan(config)(pcm, lengths=lens, offsets=offs, n_threads=host_threads())
    fe_plain = fe if not spec["sa"] else FeatureFrontend(spec["ftype"], **CFG, **spec["extra"],
                                                          use_delta_accelerate=spec["delta"], device=dev)
    got, nf = fe_plain(torch.from_numpy(pcm).to(dev), lengths=lens, offsets=offs, kernel=kernel)
    got, nf = got.cpu().numpy(), nf.cpu().numpy()
    if not np.array_equal(nf, nf_want):
        raise SystemExit("bench.py: frame counts differ from the CPU restatement")
    err = 0.0
    for b in range(got.shape[0]):
        t = int(nf[b])
        g, w = got[b, :t, :, :1], want[b, :t, :, :1]
        if spec["ftype"] == "mfcc":
            e = float(np.abs(g - w).max()) if t else 0.0
            if e > 1e-3 * np.sqrt(160.0):
                raise SystemExit("bench.py: MFCC parity broken on utterance %d: %.3e" % (b, e))
        else:
            try:
                assert_logmel_close(g, w, what="bench parity, utterance %d" % b)
            except AssertionError as exc:
                raise SystemExit("bench.py: %s" % exc)
            above = np.exp(w.astype(np.float64)) >= 1e-7
            e = float(np.abs(g - w)[above].max()) if above.any() else 0.0
        err = max(err, e)
        if got[b, t:].any():
            raise SystemExit("bench.py: padding of utterance %d is not zero" % b)
    return err


def mask_identity(fe, dev, torch, dist, rank, world, kernel):
    """BASELINE config 3: masks must be identical at 1/2/4/8 GPUs.  One fixed global batch of 256 utterances
    (ids 0..255) is split over the ranks in contiguous ranges; each rank reports, per utterance, which
    frames and which bins are entirely zero; rank 0 checks them against the oracle's draws for
    (seed, global id) and hashes them - the sha1 is the same for every N."""
    from oracle import logmel_oracle as O
    from speech_recognition_b200.sharding import shard_ranges

    n_glob = 256
    begin, end = shard_ranges([L] * n_glob, world)[rank]
    pcm = synth_batch(777, n_glob)[begin:end]
    rows = torch.zeros((n_glob, T), dtype=torch.uint8, device=dev)
    cols = torch.zeros((n_glob, fe.frequency_dim), dtype=torch.uint8, device=dev)
    if end > begin:
        out, _ = fe(torch.from_numpy(np.ascontiguousarray(pcm)).to(dev), seed=SEED, first_utterance_id=begin, kernel=kernel)
        z = out[..., 0] == 0
        rows[begin:end] = z.all(dim=2).to(torch.uint8)
        cols[begin:end] = z.all(dim=1).to(torch.uint8)
    if dist is not None:
        dist.all_reduce(rows, op=dist.ReduceOp.MAX)
        dist.all_reduce(cols, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    rows, cols = rows.cpu().numpy().astype(bool), cols.cpu().numpy().astype(bool)
    ok = True
    for g in range(n_glob):
        tbl = O.specaug_draw_masks(T, fe.frequency_dim, SA["F"], SA["m_F"], SA["T"], SA["p"], SA["m_T"], seed=SEED, utterance_id=g)
        wr, wc = np.zeros(T, bool), np.zeros(fe.frequency_dim, bool)
        for f0, f in tbl[:SA["m_F"]]:
            wc[f0:f0 + f] = True
        for t0, t in tbl[SA["m_F"]:]:
            wr[t0:t0 + t] = True
        if wc.all() or wr.all():     # (everything masked: every row and column is zero)
            wr[:], wc[:] = True, True
        ok &= bool(np.array_equal(rows[g], wr) and np.array_equal(cols[g], wc))
    sha = hashlib.sha1(np.packbits(rows).tobytes() + np.packbits(cols).tobytes()).hexdigest()
    return {"global_batch": n_glob, "ranks": world, "sha1": sha, "oracle_draws_match": ok,
            "what": "per utterance: frames and bins that are entirely zero, global ids 0..255 split over the ranks"}


# ----------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="2", choices=sorted(CONFIGS))
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "tc"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = min(steps, 20)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--trace-steps", action="store_true",
                    help="diagnostic: one CUDA event per timed step, per-step durations in the line's step_us")
    ap.add_argument("--spin-up-ms", type=float, default=SPIN_UP_MS,
                    help="untimed passes of the step before the W warm-up steps (clock / power-state ramp)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch

    from speech_recognition_b200 import FeatureFrontend
    from speech_recognition_b200.sharding import shard_ranges

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

    spec = CONFIGS[args.config]
    C = 3 if spec["delta"] else 1
    F = spec["F"]
    fe = FeatureFrontend(spec["ftype"], **CFG, **spec["extra"], use_delta_accelerate=spec["delta"],
                         spec_augment=SA if spec["sa"] else None, device=dev)
    warmup = max(args.warmup, 3)
    stream = torch.cuda.current_stream(dev)
    status = torch.zeros((1,), dtype=torch.int32, device=dev)

    if args.config != "5":
        # ring of distinct batches, resident in HBM before the timed region starts
        ring_in = [torch.from_numpy(synth_batch(1000 * rank + i)).to(dev) for i in range(RING)]
        ring_out = [torch.empty((B, T, F, C), dtype=torch.float32, device=dev) for _ in range(RING)]
        d_lengths = torch.full((B,), L, dtype=torch.int32, device=dev)   # resident: no per-step H2D
        d_nframes = torch.empty((B,), dtype=torch.int32, device=dev)
        audio_s_per_step = float(B * SECONDS)                            # per rank
        algo_bytes_per_step = B * (L * 2 + T * F * C * 4)                # per rank
        launches_note = "one fused kernel per step" if spec["ftype"] != "mfcc" else "log-mel stage + DCT kernel per step"

        # the steady-state form of the public call (FeatureFrontend.bind: arguments checked once, the step
        # is the C-ABI call alone); the staging pass below goes through FeatureFrontend.__call__
        bound = [fe.bind(ring_in[k], d_lengths, ring_out[k], d_nframes, seed=SEED, first_utterance_id=rank * B,
                         status=status, kernel=args.kernel) for k in range(RING)]

        def step(i, st=None):
            if st is None:
                bound[i % RING]()
                return
            r = fe(ring_in[i % RING], lengths=d_lengths, out=ring_out[i % RING], n_frames=d_nframes, seed=SEED,
                   first_utterance_id=rank * B, kernel=args.kernel, return_status=True)
            st |= r[2]

        n_stage = RING
    else:
        # 10 h, batches of 256 in arrival order; every batch split over the ranks by samples
        lens_all = config5_lengths()
        batches = []
        real_samples = 0
        out_bytes = 0
        g = torch.Generator(device=dev).manual_seed(11 + rank)
        for s in range(0, len(lens_all), B):
            ln_full = lens_all[s:s + B]
            tmax = fe.num_frames(int(ln_full.max()))
            b0, b1 = shard_ranges(ln_full.tolist(), world)[rank]
            ln = ln_full[b0:b1]
            if len(ln) == 0:
                continue
            starts = np.concatenate([[0], np.cumsum((ln + 7) // 8 * 8)[:-1]]).astype(np.int64)  # 16-byte aligned starts
            total = int(starts[-1] + (ln[-1] + 7) // 8 * 8)
            buf = (torch.randn((total,), device=dev, generator=g) * 0.1).clamp_(-1, 1).mul_(32767).round_().to(torch.int16)
            batches.append((buf, torch.from_numpy(starts).to(dev), torch.from_numpy(ln.astype(np.int32)).to(dev), tmax,
                            torch.empty((len(ln), tmax, F, C), dtype=torch.float32, device=dev),
                            torch.empty((len(ln),), dtype=torch.int32, device=dev), s + b0))
            real_samples += int(ln.sum())
            out_bytes += sum(fe.num_frames(int(x)) for x in ln) * F * C * 4
        audio_s_per_step = real_samples / SR                             # this rank's share of the 10 h
        algo_bytes_per_step = real_samples * 2 + out_bytes               # real (unpadded) bytes only
        launches_note = "%d batches per pass on this rank" % len(batches)

        def step(i, st=None):
            for buf, off, ln, tmax, out, nfr, first in batches:
                r = fe(buf, lengths=ln, offsets=off, t_max=tmax, out=out, n_frames=nfr, first_utterance_id=first,
                       kernel=args.kernel, return_status=st is not None)
                if st is not None:
                    st |= r[2]

        n_stage = 1

    # staging: every buffer goes through the path once (first touch of the output pages, the kernel's
    # status word checked on real data) before the W warm-up steps
    for i in range(n_stage):
        step(i, status)
    torch.cuda.synchronize()
    if int(status.item()) != 0:
        raise SystemExit("bench.py: kernel status word %#x on the staged batches (0x100 = pipeline wait timed out)" % int(status.item()))
    sampler = ClockSampler(torch.cuda._get_nvml_device_index(dev) if hasattr(torch.cuda, "_get_nvml_device_index")
                           else local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # clock / power-state ramp: NVML start-up above leaves the GPU idle for tens of ms; untimed passes of the
    # same step bring SM and HBM clocks back to their loaded state before the W warm-up steps
    n_spin = 0
    if dist is not None and not os.environ.get("B200MEL_BENCH_NO_ALIGN"):
        # align the ranks BEFORE the untimed steps: the mandated barrier in front of the timed region is then
        # reached by all ranks within microseconds, and no GPU sits idle for milliseconds right before it is timed
        dist.barrier()
        torch.cuda.synchronize()
    if args.spin_up_ms > 0:
        t_spin = time.perf_counter()
        while (time.perf_counter() - t_spin) * 1e3 < args.spin_up_ms:
            step(n_spin)
            n_spin += 1
            if n_spin % 8 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
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
    step_events = []
    for i in range(args.steps):
        step(i)
        if args.trace_steps:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            step_events.append(e)
    host_us_per_call = (time.perf_counter() - t_host0) * 1e6 / max(args.steps, 1)   # enqueue cost, GPU running behind
    ev1.record(stream)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    clocks = sampler.stop()
    launches = fe.launch_count - launches0
    if int(status.item()) != 0:   # (the bound steps OR into the same word: a 0x100 pipeline time-out cannot be silent)
        raise SystemExit("bench.py: kernel status word %#x after the timed steps" % int(status.item()))
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    total_audio = torch.tensor([audio_s_per_step], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(total_audio, op=dist.ReduceOp.SUM)
    ms_max = float(t.item())
    value = float(total_audio.item()) * args.steps / (ms_max * 1e-3)

    identity = None
    if spec["sa"]:
        identity = mask_identity(fe, dev, torch, dist, rank, world, args.kernel)

    # ---- e2e: host buffers through the C ABI; H2D + kernel + D2H inside the timed region --------
    e2e = None
    if args.config != "5":
        e2e_steps = args.e2e_steps or min(args.steps, 20)
        host_in = [torch.from_numpy(synth_batch(5000 + 10 * rank + i)).pin_memory() for i in range(2)]
        host_out = [torch.empty((B, T, F, C), dtype=torch.float32).pin_memory() for _ in range(2)]
        kw = dict(kernel=args.kernel, seed=SEED, first_utterance_id=rank * B)
        for i in range(2):   # untimed: staging buffers of both slots are allocated here, both forms run once
            fe.forward_host(host_in[i % 2], out=host_out[i % 2], **kw)
        for i in range(2):
            fe.forward_host_submit(host_in[i % 2], out=host_out[i % 2], slot=i % 2, **kw)
        for i in range(2):
            fe.forward_host_wait(i % 2)
        # (a) synchronous call per step: upload, kernels, download, return
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            fe.forward_host(host_in[i % 2], out=host_out[i % 2], **kw)
        e2e_sync_s = time.perf_counter() - t0
        # (b) the streaming form of the same call (submit / wait on two staging slots): step i+1 is submitted
        #     before step i is waited for, so its uploads overlap step i's downloads.  Every step's input is
        #     copied from pinned host memory and its result is read on the host inside the timed region.
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        checksum = 0.0
        fe.forward_host_submit(host_in[0], out=host_out[0], slot=0, **kw)
        for i in range(e2e_steps):
            if i + 1 < e2e_steps:
                fe.forward_host_submit(host_in[(i + 1) % 2], out=host_out[(i + 1) % 2], slot=(i + 1) % 2, **kw)
            fe.forward_host_wait(i % 2)
            checksum += float(host_out[i % 2][0, 0, 0, 0])   # the step's result is read on the host
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        te = torch.tensor([e2e_s, e2e_sync_s], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_stream_value = world * e2e_steps * audio_s_per_step / float(te[0].item())
        e2e_sync_value = world * e2e_steps * audio_s_per_step / float(te[1].item())
        # both forms move every step's input up and its result down inside the timed region; the line's value is
        # the faster one on this box, both are listed
        e2e = {"value": max(e2e_stream_value, e2e_sync_value), "unit": UNIT, "h2d_bytes_per_step": B * L * 2,
               "d2h_bytes_per_step": B * T * F * C * 4, "steps": e2e_steps,
               "timer": "host wall clock around the loop of C-ABI calls, max over ranks",
               "form": "submit/wait on two staging slots" if e2e_stream_value >= e2e_sync_value else "synchronous call",
               "stream_value": e2e_stream_value, "sync_call_value": e2e_sync_value, "checksum": checksum}

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    per_step_s = ms * 1e-3 / max(args.steps, 1)          # rank 0's own steps
    achieved = algo_bytes_per_step / per_step_s / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "strong" if args.config == "5" else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": spec["workload"], "config": args.config, "input": "int16 PCM resident in HBM",
                   "output": "[%s,%s,%d,%d] f32" % ("B" if args.config == "5" else B, "Tmax" if args.config == "5" else T, F, C),
                   "frame_length": 320, "frame_step": 160, "fft_length": 320, "num_mel_bins": 80,
                   "batch_per_gpu": B if args.config != "5" else "256 / %d, by samples" % world,
                   "global_batch": B * world if args.config != "5" else B,
                   "parallelism": "utterance-sharded x%d, no collective" % world,
                   "l2_policy": ("ring of %d distinct batches (%.0f MB touched per cycle > 126 MB L2)" % (RING, RING * algo_bytes_per_step / 1e6))
                                if args.config != "5" else "one pass touches %.0f MB of distinct buffers (> 126 MB L2)" % (algo_bytes_per_step / 1e6),
                   "kernel": args.kernel, "tc_supported": bool(fe.tc_supported), "launches": launches_note},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(args.config), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": algo_bytes_per_step, "launch_us": per_step_s * 1e6,
                     "host_enqueue_us_per_step": host_us_per_call},
        "gpu_launches": int(launches),
        **({"step_us": [round(1e3 * a.elapsed_time(b), 2) for a, b in zip([ev0] + step_events[:-1], step_events)]}
           if step_events else {}),
        **({"spin_up": {"ms": args.spin_up_ms, "untimed_steps": n_spin}} if n_spin else {}),
        "clocks": clocks,
    }
    if e2e is not None:
        line["e2e"] = e2e
    if identity is not None:
        line["mask_identity"] = identity
        if not identity["oracle_draws_match"]:
            raise SystemExit("bench.py: SpecAugment zero set differs from the oracle's draws: %r" % (identity,))
    if not args.no_cpu_baseline:
        try:
            n_utt = 64
            v1, per_step, cores = cpu_reference_run(args.config, n_utt, 1, 1)
            reps = int(max(1, min(40, 10.0 / max(per_step, 1e-3))))   # ~10 s of wall time on all cores
            v, per_step, cores = cpu_reference_run(args.config, n_utt, reps, 0)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "%d utterances x %d passes (oracle C restatement of the TF CPU "
                                              "path; TensorFlow not installable here)" % (n_utt, reps)}
            line["cpu_baseline"]["parity_max_abs_err"] = parity_against_cpu(args.config, fe, dev, torch, n_utt, args.kernel)
        except SystemExit:
            raise
        except Exception as e:  # the baseline is informational; never lose the GPU numbers to it
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": "failed: %r" % (e,)}
    emit_line(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
