"""Device-resident throughput of BASELINE configs 2-5 through the public
FeatureFrontend API (= the C ABI), CUDA events, two distinct buffer sets per case (> L2 per cycle).
Prints one JSON line per case; `python experiments/config_sweep.py > gpurun_out/configs.jsonl`.
Config 2 C=1 is the headline bench.py line; everything else here is secondary evidence."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_recognition_b200 import FeatureFrontend

PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))).get("hbm_gbs", 6546.9) \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6546.9
LIBRI = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
             lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
SA = dict(enable=True, W=None, F=27, m_F=2, T=100, p=1.0, m_T=2)
dev = torch.device("cuda:0")


def timed(fn, n_warm=4, n=30):
    for i in range(n_warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n  # us per call


def report(name, fe, us, audio_s, in_bytes, out_bytes, kernel, extra=None):
    row = {"case": name, "kernel": kernel, "us_per_batch": round(us, 1), "audio_s_per_s": round(audio_s / (us * 1e-6)),
           "algorithmic_bytes": int(in_bytes + out_bytes), "GBps": round((in_bytes + out_bytes) / us / 1e3, 1),
           "frac_of_hbm_peak": round((in_bytes + out_bytes) / us / 1e3 / PEAK, 4),
           "out_shape_FC": [fe.frequency_dim, fe.feature_dim]}
    if extra:
        row.update(extra)
    print(json.dumps(row), flush=True)


def dense_case(name, ftype, cfg, B, L, dtype=torch.int16, delta=False, sa=None, seed=0, n=30, exact=False):
    fe = FeatureFrontend(ftype, **cfg, use_delta_accelerate=delta, spec_augment=sa, device=dev)
    g = torch.Generator(device="cuda").manual_seed(0)
    xs = []
    for _ in range(2):
        x = (torch.randn((B, L), device=dev, generator=g) * 0.1).clamp_(-1, 1)
        if dtype == torch.int16:
            xs.append((x * 32767).round().to(torch.int16))
        elif exact:
            xs.append((x * 32767).round() / 32768.0)   # what load_audio_file produces: int16 / 32768
        else:
            xs.append(x)
        del x
    T = fe.num_frames(L)
    outs = [torch.empty((B, T, fe.frequency_dim, fe.feature_dim), dtype=torch.float32, device=dev) for _ in range(2)]
    lens = torch.full((B,), L, dtype=torch.int32, device=dev)
    us = timed(lambda i: fe(xs[i % 2], lengths=lens, out=outs[i % 2], seed=seed, first_utterance_id=0), n=n)
    kern = "tc" if (fe.tc_supported and (dtype == torch.int16 or exact)) else "generic"
    report(name, fe, us, B * L / cfg["sample_rate"], B * L * xs[0].element_size(),
           B * T * fe.frequency_dim * fe.feature_dim * 4, kern, {"B": B, "L": L, "T": T})


# ---- config 2 (+ deltas), config 3 (SpecAugment, seed 43) -------------------------------------
dense_case("config2 log-mel C=1", "log-mel-spectrogram", LIBRI, 256, 240000)
dense_case("config2 log-mel C=3 (delta)", "log-mel-spectrogram", LIBRI, 256, 240000, delta=True)
dense_case("config3 log-mel C=1 + SpecAugment", "log-mel-spectrogram", LIBRI, 256, 240000, sa=SA, seed=43)
dense_case("config3 log-mel C=3 + SpecAugment (libri_config.yml as shipped)", "log-mel-spectrogram", LIBRI, 256, 240000,
           delta=True, sa=SA, seed=43)
dense_case("config2 float32 PCM in (int16 / 32768, as load_audio_file gives), C=1", "log-mel-spectrogram", LIBRI, 256, 240000,
           dtype=torch.float32, n=10, exact=True)
dense_case("config2 float32 PCM in (arbitrary floats), C=1", "log-mel-spectrogram", LIBRI, 256, 240000, dtype=torch.float32, n=10)
# ---- config 4 (i): kspon / clovacall MFCC-40 + delta -------------------------------------------
dense_case("config4 kspon mfcc-40 C=3", "mfcc", dict(LIBRI, num_mfcc=40), 256, 240000, delta=True, n=10)
# ---- config 4 (ii): the reference's own parameter grid (tests/test_data.py:84-87) + 400/160/512 ----
for (sr, lw, h, nfft, m, lo, hi) in [(22050, 1024, 1024, 1024, 80, 10.0, 10000.0), (16000, 128, 64, 256, 123, 12.0, 88.0),
                                     (32000, 128, 80, 128, 321, 32.0, 16000.0), (44100, 512, 512, 256, 333, 333.0, 3333.0),
                                     (16000, 400, 160, 512, 80, 80.0, 7600.0)]:
    cfg = dict(sample_rate=sr, frame_length=lw, frame_step=h, fft_length=nfft, num_mel_bins=m, lower_edge_hertz=lo,
               upper_edge_hertz=hi)
    dense_case("config4 grid sr=%d Lw=%d H=%d N=%d M=%d" % (sr, lw, h, nfft, m), "log-mel-spectrogram", cfg, 64, 15 * sr, n=6)
dense_case("spectrogram 320/160/320 (A9)", "spectrogram", dict(sample_rate=16000, frame_length=320, frame_step=160,
                                                               fft_length=320), 64, 240000, n=6)

# ---- config 5: 10 h of variable-length utterances, batches of 256 in arrival order --------------
rng = np.random.default_rng(1)
durs = []
while sum(durs) < 36000.0:
    durs.append(float(np.clip(rng.normal(12.3, 4.5), 1.0, 20.48)))
lens_all = np.array([int(d * 16000) for d in durs], dtype=np.int64)
for aligned in (True, False):
    fe = FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True, device=dev)
    batches = []
    for s in range(0, len(lens_all), 256):
        ln = lens_all[s:s + 256]
        starts = np.zeros(len(ln), dtype=np.int64)
        pos = 0
        for i, l in enumerate(ln):
            starts[i] = pos
            pos += int(l)
            if aligned:
                pos = (pos + 7) // 8 * 8  # utterances start on 16-byte boundaries in the packed buffer
        buf = (torch.randn((pos + 8,), device=dev) * 0.1).clamp_(-1, 1).mul_(32767).round_().to(torch.int16)
        tmax = fe.num_frames(int(ln.max()))
        batches.append((buf, torch.from_numpy(starts).to(dev), torch.from_numpy(ln.astype(np.int32)).to(dev), tmax,
                        torch.empty((len(ln), tmax, 80, 3), dtype=torch.float32, device=dev), s))

    def run_all(_):
        for buf, off, ln, tmax, out, first in batches:
            fe(buf, lengths=ln, offsets=off, t_max=tmax, out=out, first_utterance_id=first)

    us = timed(run_all, n_warm=2, n=5)
    frames = sum(fe.num_frames(int(l)) for l in lens_all)
    report("config5 ragged 10 h, %d utterances, %d batches, [B,Tmax,80,3] (%s offsets)" % (
        len(lens_all), len(batches), "16-byte aligned" if aligned else "packed, unaligned"), fe, us,
        float(lens_all.sum()) / 16000.0, int(lens_all.sum()) * 2, frames * 80 * 3 * 4, "tc",
        {"real_audio_s": round(float(lens_all.sum()) / 16000.0, 1),
         "padded_out_bytes": int(sum(b[4].numel() * 4 for b in batches))})
    del batches
