"""Parity tests proper: the CUDA path (through the C ABI) against the oracle on the same inputs.

Tolerance (stated in BASELINE.md section 6): max |delta log-mel| <= 1e-3 against the float64
oracle where the mel energy >= 1e-7, linear-domain bound below it; masks / padding / delta given
the features are bit-exact.
"""
import os

import numpy as np
import pytest
import torch

from oracle import logmel_oracle as O

from .helpers import (CONFIG_DIR, GOLDEN_DIR, LIBRI, REF_LOGMEL_GRID, REF_MFCC_GRID, REF_SPECTROGRAM_GRID,
                      assert_logmel_close, coloured_int16, noise_int16, tone_int16)

pytestmark = pytest.mark.gpu

KERNELS = ["generic", "auto"]   # auto = the tensor-core kernel for the shipped parameter set (int16 PCM)


@pytest.fixture(scope="module")
def fe_mod(built_lib):
    import speech_recognition_b200 as pkg

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return pkg


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def run_logmel(pkg, pcm_list, kernel="generic", dtype=np.int16, **kw):
    cfg = dict(LIBRI)
    cfg.update(kw)
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **cfg)
    L = max(len(p) for p in pcm_list)
    batch = np.zeros((len(pcm_list), L), dtype=dtype)
    for i, p in enumerate(pcm_list):
        batch[i, :len(p)] = p if dtype == np.int16 else p.astype(np.float32) / 32768.0
    out, nf = fe(dev(batch), lengths=[len(p) for p in pcm_list], kernel=kernel)
    torch.cuda.synchronize()
    return out.cpu().numpy(), nf.cpu().numpy(), fe


# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kernel", KERNELS)
def test_reference_golden_fixture(fe_mod, kernel):
    """reference tests/test_data.py:53-57: silence -> log(1e-12) = -27.631021 exactly, [412, 80, 1]."""
    g = np.load(os.path.join(GOLDEN_DIR, "ref_wav_dataset_tfrecord.npz"))
    pcm = g["wav_pcm_int16"]
    out, nf, _ = run_logmel(fe_mod, [pcm, pcm], kernel)
    assert out.shape == (2, 412, 80, 1) and list(nf) == [412, 412]
    for i in range(2):
        assert np.array_equal(out[i], g["feature_%d" % i])
    # float32 entry point (the reference's own input type) gives the same
    outf, _, _ = run_logmel(fe_mod, [pcm], kernel, dtype=np.float32)
    assert np.array_equal(outf[0], g["feature_0"])


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("signal", ["noise", "quiet_noise", "coloured", "tone_on_bin", "tone_off_bin", "square"])
def test_logmel_libri_signals(fe_mod, kernel, signal):
    n = 48000
    if signal == "noise":
        pcm = noise_int16(n, 0, 0.1)
    elif signal == "quiet_noise":
        pcm = noise_int16(n, 1, 1e-3)
    elif signal == "coloured":
        pcm = coloured_int16(n, 2)
    elif signal == "tone_on_bin":
        pcm = tone_int16(n, 1000.0)     # 1000 Hz = bin 20 exactly
    elif signal == "tone_off_bin":
        pcm = tone_int16(n, 1234.5)
    else:
        pcm = (np.where((np.arange(n) // 37) % 2 == 0, 32767, -32768)).astype(np.int16)
    out, nf, _ = run_logmel(fe_mod, [pcm], kernel)
    want = O.make_log_mel_spectrogram(pcm.astype(np.float64) / 32768.0, **LIBRI, dtype=np.float64)
    assert out.shape == (1,) + want.shape
    assert_logmel_close(out[0], want, what="%s/%s" % (kernel, signal))


@pytest.mark.parametrize("kernel", KERNELS)
def test_f32_entry_matches_int16_entry(fe_mod, kernel):
    pcm = noise_int16(32000, 5)
    a, _, _ = run_logmel(fe_mod, [pcm], kernel, dtype=np.int16)
    b, _, _ = run_logmel(fe_mod, [pcm], kernel, dtype=np.float32)
    if kernel == "generic":
        assert np.array_equal(a, b)     # int16 * 2^-15 is exact, so both entries see the same floats
    else:                               # auto: exact float32 samples take the tensor-core kernel too
        assert np.array_equal(a, b)


@pytest.mark.parametrize("ftype,extra", [("log-mel-spectrogram", {}), ("mfcc", {"num_mfcc": 40})])
def test_f32_entry_is_gated_on_exactness(fe_mod, ftype, extra):
    """float32 PCM with kernel = auto: samples that are exact int16 / 32768 (the reference's
    load_audio_file output, data.py:98-116) take the tensor-core path and give bit-for-bit the int16
    entry's result; as soon as one valid sample is not, the generic kernel takes the whole batch
    (bit-for-bit the generic result).  Samples beyond lengths[b] are not looked at.  No host
    round trip: both kernels are launched, a device flag gates them."""
    sa = dict(enable=True, W=None, F=15, m_F=2, T=40, p=1.0, m_T=2)
    fe = fe_mod.FeatureFrontend(ftype, **dict(LIBRI, **extra), use_delta_accelerate=True, spec_augment=sa)
    lengths = [32000, 20000, 16123, 100]
    i16 = np.zeros((4, 32000), np.int16)
    for i, n in enumerate(lengths):
        i16[i, :n] = noise_int16(n, 90 + i)
    i16[0, 7] = -32768
    f32 = i16.astype(np.float32) / 32768.0
    f32[1, 20000:] = np.nan                               # garbage behind the utterance must not matter
    kw = dict(lengths=lengths, seed=3, first_utterance_id=50)
    want_tc, nf = fe(dev(i16), kernel="tc", **kw)
    got, nf2 = fe(dev(f32), kernel="auto", **kw)
    assert torch.equal(got, want_tc) and torch.equal(nf, nf2)
    inexact = f32.copy()
    inexact[2, 5000] += 1e-6
    want_gen, _ = fe(dev(inexact), kernel="generic", **kw)
    got2, nf3 = fe(dev(inexact), kernel="auto", **kw)
    assert torch.equal(got2, want_gen) and torch.equal(nf, nf3)
    over = f32.copy()
    over[1, 10] = 1.0                                     # +32768 is not an int16
    got3, _ = fe(dev(over), kernel="auto", **kw)
    want3, _ = fe(dev(over), kernel="generic", **kw)
    assert torch.equal(got3, want3)
    # a bad sample that no frame uses (utterance 3 has 100 samples = no frame at all): the tensor-core kernel,
    # which reads the floats itself, never meets it (the two-pass form with the conversion kernel looks at
    # every sample below lengths[b] and hands the batch to the generic kernel: either result is the features
    # of the same valid frames)
    unused = f32.copy()
    unused[3, 10] = 1.0
    got4, _ = fe(dev(unused), kernel="auto", **kw)
    want4, _ = fe(dev(unused), kernel="generic", **kw)
    assert torch.equal(got4, want_tc) or torch.equal(got4, want4)


@pytest.mark.parametrize("kernel", KERNELS)
def test_edge_lengths_and_padding(fe_mod, kernel):
    lengths = [0, 319, 320, 479, 480, 1000, 66150 // 8, 5000]
    pcms = [noise_int16(n, 10 + i) for i, n in enumerate(lengths)]
    out, nf, fe = run_logmel(fe_mod, pcms, kernel)
    assert list(nf) == [O.frame_count(n, 320, 160) for n in lengths] == [fe.num_frames(n) for n in lengths]
    for i, p in enumerate(pcms):
        want = O.make_log_mel_spectrogram(p.astype(np.float64) / 32768.0, **LIBRI, dtype=np.float64)
        t = want.shape[0]
        if t:
            assert_logmel_close(out[i, :t], want, what="len %d" % lengths[i])
        pad = out[i, t:]
        assert np.all(pad == 0.0) and not np.any(np.signbit(pad))       # exactly +0.0


@pytest.mark.parametrize("sr,fl,fs,n,m,lo,hi", REF_LOGMEL_GRID + [(16000, 400, 160, 512, 80, 80.0, 7600.0),
                                                                    (16000, 320, 160, 320, 40, 0.0, 8000.0),
                                                                    (8000, 200, 80, 200, 23, 100.0, 3800.0),
                                                                    (16000, 15, 7, 15, 10, 100.0, 7000.0),
                                                                    (16000, 14, 5, 22, 10, 100.0, 7000.0)])
def test_logmel_parameter_grid(fe_mod, sr, fl, fs, n, m, lo, hi):
    """The reference's own parametrisation (tests/test_data.py:81-89; shape-only there) plus
    non-power-of-two, odd and prime-factor FFT lengths, with values checked."""
    pcm = noise_int16(20000, 7)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", sr, fl, fs, n, m, None, float(lo), float(hi))
    out, nf = fe(dev(pcm[None]))
    want = O.make_log_mel_spectrogram(pcm.astype(np.float64) / 32768.0, sr, fl, fs, n, m, lo, hi, dtype=np.float64)
    assert out.shape[1:] == want.shape == ((20000 - fl + fs) // fs, m, 1)
    assert_logmel_close(out[0].cpu().numpy(), want, what=str((sr, fl, fs, n, m)))
    Wlib = fe.mel_weights()
    Wora = O.linear_to_mel_weight_matrix(m, n // 2 + 1, sr, lo, hi)
    assert np.allclose(Wlib, Wora, rtol=0, atol=2e-6)
    assert np.array_equal(Wlib > 0, Wora > 0) or np.abs(Wlib - Wora).max() < 2e-6


@pytest.mark.parametrize("fl,fs,n", REF_SPECTROGRAM_GRID + [(320, 160, 320), (400, 160, 512), (21, 9, 21)])
def test_spectrogram_grid(fe_mod, fl, fs, n):
    pcm = noise_int16(20000, 8)
    fn = fe_mod.make_spectrogram(fl, fs, n)
    got = fn(pcm).cpu().numpy()
    want = O.make_spectrogram(pcm.astype(np.float64) / 32768.0, fl, fs, n, np.float64)
    nn = n if n is not None else O.enclosing_power_of_two(fl)
    assert got.shape == want.shape == ((20000 - fl + fs) // fs, nn // 2 + 1, 1)
    scale = np.abs(want).max()
    assert np.abs(got - want).max() <= 2e-6 * scale + 1e-7


@pytest.mark.parametrize("sr,fl,fs,n,m,c,lo,hi", REF_MFCC_GRID + [(16000, 320, 160, 320, 80, 40, 80.0, 7600.0)])
def test_mfcc_grid(fe_mod, sr, fl, fs, n, m, c, lo, hi):
    """make_mfcc (data.py:192-241) over the reference's grid (tests/test_data.py:114-122), as two tight
    statements instead of one loose one: (i) the log-mel the coefficients are made of holds the stated
    log-mel tolerance against the float64 oracle (linear-domain rule below the floor - the rows with
    empty or nearly empty filters, M = 123 / 321 / 333); (ii) the coefficients are the oracle's
    float64 DCT-II * (2M)^-1/2 of THAT log-mel, up to float32 rounding of an M-term sum."""
    pcm = noise_int16(20000, 9)
    x64 = pcm.astype(np.float64) / 32768.0
    fn = fe_mod.make_mfcc(sr, fl, fs, n, m, c, float(lo), float(hi))
    got = fn(pcm.astype(np.float32) / 32768.0).cpu().numpy()
    want = O.make_mfcc(x64, sr, fl, fs, n, m, c, lo, hi, dtype=np.float64)
    assert got.shape == want.shape == ((20000 - fl + fs) // fs, c, 1)
    lm_fn = fe_mod.make_log_mel_spectrogram(sr, fl, fs, n, m, float(lo), float(hi))
    lm = lm_fn(pcm.astype(np.float32) / 32768.0).cpu().numpy()
    lm_want = O.make_log_mel_spectrogram(x64, sr, fl, fs, n, m, lo, hi, dtype=np.float64)
    assert_logmel_close(lm, lm_want, what="mfcc grid log-mel %s" % ((sr, fl, fs, n, m),))           # (i)
    dct_of_own = lm[..., 0].astype(np.float64) @ O.dct2_matrix(m, c)                                   # (ii)
    # float32 rounding of an M-term dot product: sqrt(M) * 2^-24 * sum |terms| (random-walk model), 4x margin
    bound = 4.0 * np.sqrt(m) * 2.0 ** -24 * (np.abs(lm[..., 0].astype(np.float64)) @ np.abs(O.dct2_matrix(m, c))) + 1e-6
    assert (np.abs(got[..., 0] - dct_of_own) <= bound).all(), (np.abs(got[..., 0] - dct_of_own) / bound).max()
    if (np.exp(lm_want) - 1e-12 >= 1e-7).all():   # every bin above the floor: the direct bound follows from (i)
        assert np.abs(got - want).max() <= 1e-3 * np.sqrt(2.0 * m)


# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kernel", KERNELS)
def test_delta_and_masks_fused(fe_mod, kernel):
    """feature -> spec_augment -> delta in the reference's stage order (run/train.py:88-116):
    given the kernel's own base features, masks and deltas must be bit-exact."""
    sa = dict(enable=True, W=None, F=27, m_F=2, T=100, p=1.0, m_T=2)   # libri_config.yml:28-35
    lengths = [48000, 31000, 16500, 20000, 100]      # >= 100 frames (no TF draw-error case), plus an empty one
    pcms = [noise_int16(n, 20 + i) for i, n in enumerate(lengths)]
    L = max(lengths)
    batch = np.zeros((len(pcms), L), np.int16)
    for i, p in enumerate(pcms):
        batch[i, :len(p)] = p
    base_fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    full_fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True, spec_augment=sa)
    base, nf = base_fe(dev(batch), lengths=lengths, kernel=kernel)
    full, nf2, status = full_fe(dev(batch), lengths=lengths, seed=43, first_utterance_id=1000, kernel=kernel,
                                return_status=True)
    base, full, nf = base.cpu().numpy(), full.cpu().numpy(), nf.cpu().numpy()
    assert full.shape == (5, 299, 80, 3) and np.array_equal(nf, nf2.cpu().numpy())
    for i in range(len(pcms)):
        t = int(nf[i])
        x = base[i, :t]
        tbl = O.specaug_draw_masks(t, 80, 27, 2, 100, 1.0, 2, seed=43, utterance_id=1000 + i) if t > 0 else None
        if t == 0:
            assert np.all(full[i] == 0)
            continue
        want = O.delta_accelerate(O.spec_augment_apply(x, tbl[:2], tbl[2:]))
        assert np.array_equal(full[i, :t], want)
        assert np.array_equal(np.signbit(full[i, :t, :, 0]), np.signbit(want[:, :, 0]))   # -0.0 where masked
        assert np.all(full[i, t:] == 0.0) and not np.any(np.signbit(full[i, t:]))


@pytest.mark.parametrize("m_F,m_T", [(2, 2), (5, 4), (8, 8)])
def test_tc_mask_paths_agree_with_the_oracle_draws(fe_mod, m_F, m_T):
    """The tensor-core kernel draws the tables itself for <= 8 masks per utterance and takes them
    from a pre-pass kernel for 9..16 (the library's maximum): both must apply exactly the oracle's
    draws (C = 1, the TMA-store path)."""
    sa = dict(enable=True, W=None, F=11, m_F=m_F, T=40, p=1.0, m_T=m_T)
    lengths = [48000, 31000, 16500, 20000]
    batch = np.zeros((4, 48000), np.int16)
    for i, n in enumerate(lengths):
        batch[i, :n] = noise_int16(n, 60 + i)
    base, nf = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI)(dev(batch), lengths=lengths, kernel="tc")
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, spec_augment=sa)
    got, _ = fe(dev(batch), lengths=lengths, seed=7, first_utterance_id=123, kernel="auto")
    base, got, nf = base.cpu().numpy(), got.cpu().numpy(), nf.cpu().numpy()
    for i in range(4):
        t = int(nf[i])
        tbl = O.specaug_draw_masks(t, 80, 11, m_F, 40, 1.0, m_T, seed=7, utterance_id=123 + i)
        want = O.spec_augment_apply(base[i, :t], tbl[:m_F], tbl[m_F:])
        assert np.array_equal(got[i, :t], want)          # same kernel, same base features
        assert np.array_equal(np.signbit(got[i, :t]), np.signbit(want))
        assert np.all(got[i, t:] == 0.0)


def test_tc_chunked_launches_match_a_single_launch():
    """A batch with more tiles than one launch's per-CTA tile lists can hold is processed in utterance
    ranges; B200MEL_TC_MAX_TILES forces that path at a small size (separate process: the limit is
    read once)."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import numpy as np, torch
        from speech_recognition_b200 import FeatureFrontend
        cfg = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
                   lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
        rng = np.random.default_rng(3)
        lengths = rng.integers(0, 48000, size=37).astype(np.int32)
        lengths[5] = 0
        pcm = torch.from_numpy(rng.integers(-3000, 3000, size=(37, 48000)).astype(np.int16)).cuda()
        sa = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
        for delta in (False, True):
            fe = FeatureFrontend("log-mel-spectrogram", **cfg, use_delta_accelerate=delta, spec_augment=sa)
            out, nf = fe(pcm, lengths=lengths, seed=43, first_utterance_id=9, kernel="tc")
            print(delta, fe.launch_count, float(out.double().abs().sum()), int(nf.sum()))
    """)
    import os
    outs = []
    for cap in ("0", "7"):
        env = dict(os.environ, B200MEL_TC_MAX_TILES=cap, PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        outs.append([l.split() for l in r.stdout.strip().splitlines()])
    whole, chunked = outs
    for a, b in zip(whole, chunked):
        assert a[0] == b[0] and a[2:] == b[2:]          # identical checksums and frame counts
        assert int(b[1]) > int(a[1])                     # ... from more launches


@pytest.mark.parametrize("num_mfcc", [13, 40, 80])
def test_mfcc_two_stage_path_matches_generic_kernel(fe_mod, num_mfcc):
    """kspon / clovacall parameters (mfcc-40 + SpecAugment on the coefficients + delta): the
    tensor-core log-mel stage followed by the DCT / mask / delta kernel must agree with the generic
    kernel - identical mask zero-set and padding, values within the MFCC tolerance - ragged lengths
    and an injected mask table included."""
    sa = dict(enable=True, W=None, F=15, m_F=2, T=40, p=1.0, m_T=2)
    cfg = dict(LIBRI, num_mfcc=num_mfcc)
    sa = dict(sa, F=min(15, num_mfcc - 1))
    lengths = [48000, 31000, 16500, 20000, 100, 320]
    batch = np.zeros((len(lengths), 48000), np.int16)
    for i, n in enumerate(lengths):
        batch[i, :n] = noise_int16(n, 80 + i)
    fe = fe_mod.FeatureFrontend("mfcc", **cfg, use_delta_accelerate=True, spec_augment=sa)
    assert fe.tc_supported
    tol = 1e-3 * np.sqrt(160.0)
    for kwargs in (dict(seed=11, first_utterance_id=500),
                   dict(mask_table=torch.tensor([[3, 4, 8, 3, 10, 7, 50, 9]] * len(lengths), dtype=torch.int32))):
        a, nfa = fe(dev(batch), lengths=lengths, kernel="tc", **kwargs)
        g, nfg = fe(dev(batch), lengths=lengths, kernel="generic", **kwargs)
        a, g = a.cpu().numpy(), g.cpu().numpy()
        assert a.shape == g.shape == (6, 299, num_mfcc, 3) and torch.equal(nfa, nfg)
        assert np.array_equal(a[..., 0] == 0, g[..., 0] == 0)          # masks and padding
        assert np.abs(a[..., 0] - g[..., 0]).max() <= tol and np.abs(a - g).max() <= 4 * tol
        for i, n in enumerate(nfa.cpu().numpy()):
            assert np.all(a[i, n:] == 0.0) and not np.any(np.signbit(a[i, n:]))


def test_masks_are_sharding_invariant(fe_mod):
    """Masks are a pure function of (seed, global utterance id): any split of the batch over
    ranks gives the same result (BASELINE config 3)."""
    sa = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, spec_augment=sa)
    batch = np.stack([noise_int16(32000, 40 + i) for i in range(8)])
    whole, _ = fe(dev(batch), seed=43, first_utterance_id=0)
    for world in (2, 4, 8):
        per = 8 // world
        parts = [fe(dev(batch[r * per:(r + 1) * per]), seed=43, first_utterance_id=r * per)[0] for r in range(world)]
        assert torch.equal(torch.cat(parts), whole)


def test_injected_mask_table_and_draw_kernel(fe_mod):
    """Draws recorded elsewhere (e.g. from a live TF run) can be injected; the device draw
    kernel reproduces the oracle's Philox stream bit for bit."""
    from speech_recognition_b200 import frontend as F

    num_time = [299, 199, 150, 101, 1499]      # > T: no draw can hit TF's maxval <= minval case
    table, nf, nt, status = F.specaug_draw(num_time, 80, 27, 2, 100, 1.0, 2, 43, 7, "cuda")
    table = table.cpu().numpy()
    assert (nf, nt) == (2, 2) and int(status.item()) == 0
    for i, t in enumerate(num_time):
        want = O.specaug_draw_masks(t, 80, 27, 2, 100, 1.0, 2, seed=43, utterance_id=7 + i)
        assert [tuple(r) for r in table[i].reshape(-1, 2)] == want
    # p < 1 caps the total masked time
    table2, _, _, _ = F.specaug_draw([234] * 50, 80, 15, 2, 70, 0.2, 2, 1, 0, "cuda")
    t2 = table2.cpu().numpy().reshape(50, 4, 2)
    assert (t2[:, 2:, 1].sum(axis=1) <= int(np.float32(234) * np.float32(0.2))).all()
    for i in range(50):
        assert [tuple(r) for r in t2[i]] == O.specaug_draw_masks(234, 80, 15, 2, 70, 0.2, 2, seed=1, utterance_id=i)
    # TF's error case (maxval <= minval): T > num_time with p = 1 can draw t == num_time
    _, _, _, st = F.specaug_draw([3] * 200, 80, None, None, 100, 1.0, 1, 5, 0, "cuda")
    bad = 0
    for i in range(200):
        try:
            O.specaug_draw_masks(3, 80, T=100, p=1.0, m_T=1, seed=5, utterance_id=i)
        except ValueError:
            bad += 1
    assert (int(st.item()) != 0) == (bad > 0)
    # injected table
    sa = dict(enable=True, F=27, m_F=1, T=100, p=1.0, m_T=1)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, spec_augment=sa)
    base_fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    pcm = noise_int16(16000, 3)
    tbl = torch.tensor([[10, 5, 20, 30]], dtype=torch.int32)
    got, _ = fe(dev(pcm[None]), mask_table=tbl)
    base, _ = base_fe(dev(pcm[None]))
    want = O.spec_augment_apply(base[0].cpu().numpy(), [(10, 5)], [(20, 30)])
    assert np.array_equal(got[0].cpu().numpy(), want)


@pytest.mark.parametrize("W,F,m_F,T,p,m_T", [(None, 27, 1, 100, 1.0, 1), (None, 15, 2, 70, 0.2, 2)])
def test_spec_augment_wrapper_reference_invariants(fe_mod, W, F, m_F, T, p, m_T):
    """The reference's own SpecAugment test (tests/test_data.py:148-163), minus time warping."""
    num_time, num_frequency = 234, 80
    fe_mod.set_random_seed(1234)
    fn = fe_mod.spec_augment(num_frequency, W, F, m_F, T, p, m_T)
    data = torch.rand(num_time, num_frequency, 1, device="cuda") * 0.9 + 0.1
    changed = False
    for _ in range(5):
        aug = fn(data)
        is_zero = (aug == 0.0).all(dim=2)
        assert int(is_zero.all(dim=0).sum()) <= F * m_F
        assert int(is_zero.all(dim=1).sum()) <= T * m_T
        assert aug.shape == data.shape
        changed |= bool((aug != data).any())
    assert changed
    aug, text = fn(data, "hello")
    assert text == "hello"
    # same seed -> same stream
    fe_mod.set_random_seed(99)
    a = fn(data)
    fe_mod.set_random_seed(99)
    assert torch.equal(a, fn(data))


def test_per_example_wrappers_compose_like_the_reference(fe_mod):
    """dataset.map(feature_fn).map(spec_augment).map(delta_accelerate) == fused batched call."""
    pcm = noise_int16(40000, 77)
    audio = pcm.astype(np.float32) / 32768.0
    feature_fn = fe_mod.make_log_mel_spectrogram(16000, 320, 160, 320, 80, 80.0, 7600.0)
    sa_fn = fe_mod.spec_augment(80, None, 27, 2, 100, 1.0, 2)
    feat, text = feature_fn(audio, [1, 2, 3])
    assert text == [1, 2, 3] and feat.shape == (249, 80, 1)
    stepwise = fe_mod.delta_accelerate(sa_fn(feat, utterance_id=5, seed=43))
    assert stepwise.shape == (249, 80, 3)
    fused_fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True,
                                      spec_augment=dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2))
    # (the per-example float32 entry holds exact int16 / 32768 samples, so it runs on the same
    #  tensor-core kernel as the fused int16 call)
    fused, _ = fused_fe(dev(pcm[None]), seed=43, first_utterance_id=5, kernel="auto")
    assert torch.equal(stepwise, fused[0])
    want = O.delta_accelerate(feat.cpu().numpy())
    assert np.array_equal(fe_mod.delta_accelerate(feat).cpu().numpy(), want)


def test_shipped_configs_end_to_end(fe_mod):
    """libri (log-mel + delta), kspon / clovacall (mfcc-40 + delta) from the unmodified YAMLs."""
    pcm = noise_int16(24000, 4)
    x64 = pcm.astype(np.float64) / 32768.0
    for name in ["libri_config.yml", "kspon_config.yml", "clovacall_config.yml"]:
        cfg = fe_mod.DataConfig.from_yaml(os.path.join(CONFIG_DIR, name))
        fe = cfg.frontend(training=True)
        out, nf = fe(dev(pcm[None]))
        assert out.shape == (1, 149, cfg.frequency_dim, cfg.feature_dim) == (1, 149, cfg.frequency_dim, 3)
        want, _ = O.batched_features([x64], cfg.audio_feature_type, cfg.sample_rate, cfg.frame_length,
                                     cfg.frame_step, cfg.fft_length, cfg.num_mel_bins, cfg.num_mfcc,
                                     cfg.lower_edge_hertz, cfg.upper_edge_hertz, use_delta=True)
        got = out.cpu().numpy()
        tol = 1e-3 if cfg.audio_feature_type != "mfcc" else 1e-3 * np.sqrt(160.0)
        assert np.abs(got[..., 0] - want[..., 0]).max() <= tol
        assert np.abs(got - want).max() <= 4 * tol
        per_example = cfg.audio_feature_fn(x64.astype(np.float32))
        assert per_example.shape == (149, cfg.frequency_dim, 1)
        # per-example float32 entry (generic kernel) vs batched int16 entry (tensor-core kernel for libri)
        assert (per_example[..., 0] - out[0, :, :, 0]).abs().max().item() <= tol


@pytest.mark.parametrize("kernel", KERNELS)
def test_ragged_batch_and_tmax_slice(fe_mod, kernel):
    """Ragged int16 input (offsets + lengths) and the slice policy (t_max < frames)."""
    lengths = [16000, 48000, 9000, 32000]
    pcms = [noise_int16(n, 60 + i) for i, n in enumerate(lengths)]
    offsets = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.int64)
    flat = np.concatenate(pcms)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True)
    out, nf = fe(dev(flat), lengths=lengths, offsets=offsets, kernel=kernel)
    dense = np.zeros((4, 48000), np.int16)
    for i, p in enumerate(pcms):
        dense[i, :len(p)] = p
    out2, nf2 = fe(dev(dense), lengths=lengths, kernel=kernel)
    assert torch.equal(out, out2) and torch.equal(nf, nf2)
    sliced, nf3 = fe(dev(dense), lengths=lengths, t_max=150, kernel=kernel)
    assert sliced.shape == (4, 150, 80, 3) and torch.equal(sliced, out[:, :150]) and torch.equal(nf3, nf)


@pytest.mark.parametrize("kernel", KERNELS)
def test_forward_host_matches_device_path(fe_mod, kernel):
    lengths = [30000 - 997 * i for i in range(19)]
    L = max(lengths)
    batch = np.zeros((19, L), np.int16)
    for i, n in enumerate(lengths):
        batch[i, :n] = noise_int16(n, 100 + i)
    # (T = 70 < 74 frames of the shortest utterance: no draw can hit TF's maxval <= minval case)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True,
                                spec_augment=dict(enable=True, F=27, m_F=2, T=70, p=1.0, m_T=2))
    want, nf = fe(dev(batch), lengths=lengths, seed=43, first_utterance_id=11, kernel=kernel)
    got, nf_h = fe.forward_host(batch, lengths=lengths, seed=43, first_utterance_id=11, kernel=kernel)
    assert np.array_equal(got, want.cpu().numpy()) and np.array_equal(nf_h, nf.cpu().numpy())
    # ragged + pinned torch buffers
    offsets = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.int64)
    flat = torch.from_numpy(np.concatenate([batch[i, :n] for i, n in enumerate(lengths)])).pin_memory()
    got2, _ = fe.forward_host(flat, lengths=lengths, offsets=offsets, seed=43, first_utterance_id=11, kernel=kernel)
    assert np.array_equal(got2, got)


@pytest.mark.parametrize("kernel", KERNELS)
def test_invalid_draws_raise_on_the_host_path(fe_mod, kernel):
    """tf.random.uniform raises when maxval <= minval (data.py:284-285, 297).  The fused device call
    reports it in the status word (asynchronous); the host-buffer call reads the word back with the
    results and raises like the reference, after delivering the (unmasked) output."""
    sa = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, spec_augment=sa)
    lengths = [32000, 100, 16000]                      # utterance 1 has no frames: the time-mask draw is invalid
    batch = np.zeros((3, 32000), np.int16)
    for i, n in enumerate(lengths):
        batch[i, :n] = noise_int16(n, 200 + i)
    out, nf, status = fe(dev(batch), lengths=lengths, seed=43, kernel=kernel, return_status=True)
    assert int(status.item()) & 1
    host_out = np.full((3, 199, 80, 1), 7.0, np.float32)
    with pytest.raises(ValueError, match="maxval <= minval"):
        fe.forward_host(batch, lengths=lengths, seed=43, kernel=kernel, out=host_out)
    assert np.array_equal(host_out, out.cpu().numpy())   # the results were delivered before the error
    ok = [32000, 31000, 16000]
    fe.forward_host(batch, lengths=ok, seed=43, kernel=kernel)   # the next step on the slot is clean


def test_handles_do_not_lower_each_others_shared_memory_cap(fe_mod):
    """cudaFuncAttributeMaxDynamicSharedMemorySize is per (function, device): a handle created later with
    a smaller tile must not break an earlier handle's launches."""
    pcm = noise_int16(20000, 8)
    big = fe_mod.FeatureFrontend("spectrogram", 16000, 320, 160, 320)
    a0, _ = big(dev(pcm[None]), kernel="generic")
    small = fe_mod.FeatureFrontend("log-mel-spectrogram", 16000, 128, 64, 128, 20, None, 80.0, 7600.0)
    small(dev(pcm[None]), kernel="generic")
    mf = fe_mod.FeatureFrontend("mfcc", **dict(LIBRI, num_mfcc=40), use_delta_accelerate=True)
    m0, _ = mf(dev(pcm[None]))
    fe_mod.FeatureFrontend("mfcc", **dict(LIBRI, num_mfcc=13))(dev(pcm[None]))
    a1, _ = big(dev(pcm[None]), kernel="generic")
    m1, _ = mf(dev(pcm[None]))
    torch.cuda.synchronize()
    assert torch.equal(a0, a1) and torch.equal(m0, m1)


def test_errors_match_reference_stack(fe_mod):
    with pytest.raises(ValueError):        # TF: upper_edge_hertz above Nyquist
        fe_mod.FeatureFrontend("log-mel-spectrogram", 16000, 320, 160, 320, 80, None, 80.0, 9000.0)
    with pytest.raises(ValueError):
        fe_mod.FeatureFrontend("log-mel-spectrogram", 16000, 320, 160, 320, 80, None, 7600.0, 80.0)
    with pytest.raises(ValueError):
        fe_mod.FeatureFrontend("nonsense", 16000, 320, 160, 320)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    with pytest.raises(RuntimeError):      # host tensor to the device entry: no silent CPU path
        fe(torch.zeros(1, 1000, dtype=torch.int16))
    with pytest.raises(ValueError):
        fe(torch.zeros(1, 1000, dtype=torch.float64, device="cuda"))
    fn = fe_mod.make_log_mel_spectrogram(16000, 320, 160, 320, 80, 80.0, 9000.0)
    with pytest.raises(ValueError):        # raised on first call, like a tf.function trace
        fn(np.zeros(1000, np.float32))
    # spec_augment wider than the feature: TF raises at the draw
    fe_mod.set_random_seed(0)
    sa = fe_mod.spec_augment(80, None, None, None, 100, 1.0, 1)
    raised = False
    for _ in range(300):
        try:
            sa(torch.ones(3, 80, 1, device="cuda"))
        except ValueError:
            raised = True
            break
    assert raised


def test_large_batch_properties(fe_mod):
    """BASELINE config-2 shape at full size through size-independent properties: time-shift
    covariance (frame t of x[160:] == frame t+1 of x) and determinism."""
    B, L = 256, 240000
    rng = np.random.default_rng(0)
    base = np.clip(np.round(np.clip(rng.normal(0, 0.1, L + 160), -1, 1) * 32768), -32768, 32767).astype(np.int16)
    batch = np.empty((B, L), np.int16)
    for b in range(B):
        batch[b] = np.roll(base, 997 * b)[:L] if b % 2 == 0 else np.roll(base, 997 * (b - 1))[160:L + 160]
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    d = dev(batch)
    out, nf = fe(d)
    out_again, _ = fe(d)
    assert out.shape == (256, 1499, 80, 1) and bool((nf == 1499).all())
    assert torch.equal(out, out_again)
    assert torch.equal(out[0::2, 1:], out[1::2, :-1])          # shift by one hop == shift by one frame
    x = batch[37].astype(np.float64) / 32768.0
    assert_logmel_close(out[37].cpu().numpy(), O.make_log_mel_spectrogram(x, **LIBRI, dtype=np.float64), what="b=37")


def test_more_than_256_utterances_in_one_launch(fe_mod):
    """The tile list is built in chunks of 256 utterances (one per thread of the eight list-building warps, the
    warps' sums exchanged through shared memory): 700 ragged utterances = three chunks, the last one partial,
    with utterances that have no frame at all in between.  Tensor-core kernel against the generic kernel
    (same tolerance as the other tc-vs-generic checks), frame counts exact, padding exactly zero; C = 1 and 3."""
    rng = np.random.default_rng(77)
    B, L = 700, 9600
    x = np.clip(np.rint(rng.standard_normal((B, L)) * 3000.0), -32768, 32767).astype(np.int16)
    lengths = rng.integers(320, L + 1, B).astype(np.int32)
    lengths[[0, 255, 256, 511, 512, 699]] = [L, 319, L, 100, 4000, L]      # chunk borders; two utterances without a frame
    for delta in (False, True):
        fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=delta)
        a, nfa, st = fe(dev(x), lengths=lengths, kernel="tc", return_status=True)
        b, nfb = fe(dev(x), lengths=lengths, kernel="generic")
        assert int(st.item()) == 0
        want_nf = np.maximum(0, 1 + (lengths.astype(np.int64) - 320) // 160)
        assert np.array_equal(nfa.cpu().numpy(), want_nf) and torch.equal(nfa, nfb)
        assert (a - b).abs().max().item() <= (4e-2 if delta else 2e-2)
        a = a.cpu().numpy()
        for i in (0, 255, 256, 300, 511, 512, 699):
            assert not a[i, want_nf[i]:].any()
