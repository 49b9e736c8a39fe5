"""Full-size parity: EVERY shipped instantiation of the tensor-core kernel at the BASELINE config-2
shape (256 utterances x 15 s: 12 tiles per utterance, 20-21 tiles per CTA - the steady state of the
pipeline: TMEM ring wrap, barrier phase parity, staging-row reuse, more than one mask-table entry per
CTA) against the oracle, every utterance.

Checker: oracle/logmel_ref.c through RefPlan (float32 like TF's CPU kernels, OpenMP: 256 x 15 s in a
fraction of a second), itself pinned to the float64 numpy oracle in tests/test_oracle.py; a float64
numpy subset is compared here as well.  Stage order = run/train.py:88-116 (feature -> SpecAugment ->
delta), padding = padded_batch (run/train.py:192-197), golden silence value = tests/test_data.py:53-57.
"""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

from oracle import logmel_oracle as O
from oracle.ref_c import RefPlan

from .helpers import LIBRI, assert_logmel_close

pytestmark = pytest.mark.gpu

B, L, T = 256, 240000, 1499
SA = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)          # libri_config.yml:28-35
MFCC_TOL = 1e-3 * np.sqrt(160.0)   # |DCT row|_1 <= 2M / sqrt(2M): the bound a 1e-3 log-mel error allows


@pytest.fixture(scope="module")
def pkg(built_lib):
    import speech_recognition_b200 as p

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return p


@pytest.fixture(scope="module")
def batch():
    """BASELINE config 2 audio (bench.py's generator), with a few hard rows: silence (the reference's
    golden value), full-scale clipping, a very quiet utterance."""
    rng = np.random.default_rng(2024)
    x = rng.standard_normal((B, L), dtype=np.float32) * np.float32(0.1)
    x[5] *= 0.002
    x[6] *= 20.0
    np.clip(x, -1.0, 1.0, out=x)
    pcm = np.clip(np.rint(x * 32768.0), -32768, 32767).astype(np.int16)
    pcm[7] = 0
    return pcm


@pytest.fixture(scope="module")
def ref_logmel(batch):
    out, nf = RefPlan("log-mel-spectrogram", **LIBRI)(batch)
    assert out.shape == (B, T, 80, 1) and (nf == T).all()
    return out


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def close_in_chunks(got, want, what, step=16):
    for b0 in range(0, got.shape[0], step):
        assert_logmel_close(got[b0:b0 + step], want[b0:b0 + step], what="%s[%d:%d]" % (what, b0, b0 + step))


def mask_of(tables, t, v):
    """keep-mask [t, v] of one utterance from its (start, width) table, frequency masks first."""
    keep = np.ones((t, v), bool)
    for f0, f in tables[0]:
        keep[:, f0:f0 + f] = False
    for t0, w in tables[1]:
        keep[t0:t0 + w] = False
    return keep


def draw_tables(n_frames, v, seed, first_id):
    out = []
    for b, t in enumerate(n_frames):
        tbl = O.specaug_draw_masks(int(t), v, SA["F"], SA["m_F"], SA["T"], SA["p"], SA["m_T"], seed=seed,
                                   utterance_id=first_id + b)
        out.append((tbl[:SA["m_F"]], tbl[SA["m_F"]:]))
    return out


def check_masked(got0, ref0, n_frames, tables, tol_fn, what):
    """got0 / ref0: [B, T, v] channel 0.  Zero set and signs bit-exact, kept values within tolerance."""
    for b in range(got0.shape[0]):
        t = int(min(n_frames[b], got0.shape[1]))
        keep = mask_of(tables[b], t, got0.shape[2])
        g, r = got0[b, :t], ref0[b, :t]
        # every masked entry is a zero; an UNMASKED exact zero is legitimate only where the true value is
        # 0 within the tolerance (log(mel + eps) = 0 at mel = 1: about one such entry per 3e7)
        assert not g[~keep].any(), "%s: utterance %d: masked entries that are not zero at %s" % (
            what, b, np.argwhere(~keep & (g != 0.0))[:4].tolist())
        stray = keep & (g == 0.0)
        assert (np.abs(r[stray]) <= 1e-3).all(), "%s: utterance %d: zeros outside the oracle's masks at %s" % (
            what, b, np.argwhere(stray & (np.abs(r) > 1e-3))[:4].tolist())
        sure = ~keep & (np.abs(r) > 2e-3)        # masked negatives are -0.0 (audio *= mask, data.py:288,301)
        assert np.array_equal(np.signbit(g[sure]), np.signbit(r[sure])), "%s: sign of masked zeros, utterance %d" % (what, b)
        pad = got0[b, t:]
        assert not pad.any() and not np.signbit(pad).any(), "%s: padding of utterance %d is not +0.0" % (what, b)
        tol_fn(np.where(keep, g, 0.0), np.where(keep, r, 0.0), b)


def check_deltas_bit_exact(got, what, n_frames=None):
    """delta_accelerate (data.py:310-328) of the kernel's own channel 0, in float32: bit-exact.  The
    reference pads AFTER the deltas (padded_batch, run/train.py:192-197): rows >= n_frames[b] are +0.0
    in every channel."""
    x = got[..., 0]
    d1 = x.copy()
    d1[:, 1:] = x[:, 1:] - x[:, :-1]
    d2 = d1.copy()
    d2[:, 1:] = d1[:, 1:] - d1[:, :-1]
    if n_frames is not None:
        pad = np.arange(got.shape[1])[None, :] >= np.asarray(n_frames)[:, None]
        d1[pad] = 0.0
        d2[pad] = 0.0
    assert got[..., 1].tobytes() == d1.tobytes(), what + ": delta channel"
    assert got[..., 2].tobytes() == d2.tobytes(), what + ": delta-delta channel"


def logmel_tol(what):
    def fn(g, r, b):
        # (masked entries were replaced by 0 in both: exp(0) - eps is far above the floor and equal)
        assert_logmel_close(g[..., None], r[..., None], what="%s utterance %d" % (what, b))
    return fn


# ----------------------------------------------------------------------------------------------
def test_config2_every_utterance(pkg, batch, ref_logmel):
    """tc_frontend_kernel<1,false>: the headline instantiation, all 256 utterances."""
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    out, nf, status = fe(dev(batch), kernel="tc", return_status=True)
    got = out.cpu().numpy()
    assert int(status.item()) == 0 and (nf.cpu().numpy() == T).all()
    close_in_chunks(got, ref_logmel, "config 2")
    assert np.all(got[7] == np.float32(-27.631021))               # the reference's golden value, bit for bit
    for b in (0, 5, 6, 100, 255):                                # float64 numpy oracle on a subset
        want = O.make_log_mel_spectrogram(batch[b].astype(np.float64) / 32768.0, **LIBRI, dtype=np.float64)
        assert_logmel_close(got[b], want, what="config 2, float64 oracle, utterance %d" % b)


def test_config3_masks_every_utterance(pkg, batch, ref_logmel):
    """tc_frontend_kernel<1,true>: in-kernel Philox draws, seed 43 (train_config_sample.yml:17)."""
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, spec_augment=SA)
    out, nf, status = fe(dev(batch), seed=43, first_utterance_id=1000, kernel="tc", return_status=True)
    got = out.cpu().numpy()
    assert int(status.item()) == 0
    tables = draw_tables(np.full(B, T), 80, 43, 1000)
    check_masked(got[..., 0], ref_logmel[..., 0], np.full(B, T), tables, logmel_tol("config 3"), "config 3")
    # an injected table (e.g. draws recorded from a live TF run) takes the other table path of the kernel
    flat = np.array([[f0, f, f0b, fb, t0, t, t0b, tb] for ((f0, f), (f0b, fb)), ((t0, t), (t0b, tb)) in tables], np.int32)
    out2, _ = fe(dev(batch), mask_table=torch.from_numpy(flat), kernel="tc")
    assert torch.equal(out, out2)


@pytest.mark.parametrize("masks", [False, True])
def test_libri_as_shipped_every_utterance(pkg, batch, ref_logmel, masks):
    """tc_frontend_kernel<3,false> / <3,true>: delta + delta-delta (libri_config.yml:26), masks before
    the deltas (run/train.py:99-116)."""
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True, spec_augment=SA if masks else None)
    out, nf, status = fe(dev(batch), seed=43, first_utterance_id=0, kernel="tc", return_status=True)
    got = out.cpu().numpy()
    assert got.shape == (B, T, 80, 3) and int(status.item()) == 0
    if masks:
        tables = draw_tables(np.full(B, T), 80, 43, 0)
        check_masked(got[..., 0], ref_logmel[..., 0], np.full(B, T), tables, logmel_tol("libri"), "libri + masks")
    else:
        close_in_chunks(got[..., :1], ref_logmel, "libri")
    check_deltas_bit_exact(got, "libri masks=%s" % masks)


@pytest.mark.parametrize("masks", [False, True])
def test_kspon_clovacall_mfcc_every_utterance(pkg, batch, masks):
    """kspon_config.yml / clovacall_config.yml: MFCC-40 + delta (and SpecAugment on the coefficients,
    v = num_mfcc, run/train.py:100-110) against the ORACLE (not against the generic kernel)."""
    cfg = dict(LIBRI, num_mfcc=40)
    ref, _ = RefPlan("mfcc", **cfg)(batch)
    assert ref.shape == (B, T, 40, 1)
    fe = pkg.FeatureFrontend("mfcc", **cfg, use_delta_accelerate=True, spec_augment=SA if masks else None)
    assert fe.tc_supported
    out, nf, status = fe(dev(batch), seed=43, first_utterance_id=77, kernel="tc", return_status=True)
    got = out.cpu().numpy()
    assert got.shape == (B, T, 40, 3) and int(status.item()) == 0 and (nf.cpu().numpy() == T).all()

    def tol(g, r, b):
        err = np.abs(g - r)
        assert err.max() <= MFCC_TOL, "mfcc utterance %d: max |d| %.3e" % (b, err.max())
        assert np.sqrt((err ** 2).mean()) <= 1e-3, "mfcc utterance %d: rms %.3e" % (b, np.sqrt((err ** 2).mean()))

    tables = draw_tables(np.full(B, T), 40, 43, 77) if masks else [([], [])] * B
    if masks:
        check_masked(got[..., 0], ref[..., 0], np.full(B, T), tables, tol, "mfcc")
    else:
        for b in range(B):
            tol(got[b, :, :, 0], ref[b, :, :, 0], b)
    check_deltas_bit_exact(got, "mfcc masks=%s" % masks)
    for b in (0, 100):                                           # float64 numpy oracle
        want = O.make_mfcc(batch[b].astype(np.float64) / 32768.0, **cfg, dtype=np.float64)[..., 0]
        keep = mask_of(tables[b], T, 40)
        assert np.abs(np.where(keep, got[b, :, :, 0] - want, 0.0)).max() <= MFCC_TOL


def test_float32_entry_full_size(pkg, batch):
    """The reference's own input type (float32 = int16 / 32768, data.py:98-116) at full size: bit for
    bit the int16 entry's result, with and without deltas / masks."""
    f32 = dev(batch.astype(np.float32) / 32768.0)
    i16 = dev(batch)
    for kw in (dict(), dict(use_delta_accelerate=True, spec_augment=SA)):
        fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, **kw)
        a, nfa = fe(i16, seed=43, kernel="tc")
        b, nfb = fe(f32, seed=43, kernel="auto")
        assert torch.equal(a, b) and torch.equal(nfa, nfb)


# ----------------------------------------------------------------------------------------------
def config5_batch(n_utt=256, seed=1):
    """BASELINE config 5: durations clip(N(12.3, 4.5^2), 1.0, 20.48) s, default_rng(1)."""
    rng = np.random.default_rng(seed)
    dur = np.clip(rng.normal(12.3, 4.5, n_utt), 1.0, 20.48)
    lengths = (dur * 16000).astype(np.int32)
    lengths[3] = 0                       # an empty utterance and one below a frame: padded_batch rows of zeros
    lengths[4] = 319
    pcms = []
    for i, n in enumerate(lengths):
        x = rng.standard_normal(int(n), dtype=np.float32) * np.float32(0.1)
        pcms.append(np.clip(np.rint(np.clip(x, -1, 1) * 32768.0), -32768, 32767).astype(np.int16))
    return lengths, pcms


@pytest.mark.parametrize("layout", ["aligned", "packed"])
def test_config5_ragged_batch(pkg, layout):
    """One config-5 batch (256 variable-length utterances, [B, Tmax, 80, 3]): balanced tile partition
    across utterance boundaries, padding zero-fill by the producer warps, aligned (multiple-of-8
    offsets: the fast producer) and packed (arbitrary offsets: the element-wise producer) layouts."""
    lengths, pcms = config5_batch()
    if layout == "aligned":
        starts = np.concatenate([[0], np.cumsum((lengths + 7) // 8 * 8)[:-1]]).astype(np.int64)
        total = int(starts[-1] + (lengths[-1] + 7) // 8 * 8)
    else:
        starts = np.concatenate([[0], np.cumsum(lengths + 1)[:-1]]).astype(np.int64)   # odd offsets too
        total = int(starts[-1] + lengths[-1] + 1)
    flat = np.full(total, 12345, np.int16)               # garbage between the utterances must not matter
    for s, p in zip(starts, pcms):
        flat[s:s + len(p)] = p
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True)
    out, nf, status = fe(dev(flat), lengths=lengths, offsets=starts, kernel="tc", return_status=True)
    got = out.cpu().numpy()
    ref, nf_ref = RefPlan("log-mel-spectrogram", **LIBRI)(flat, lengths=lengths, offsets=starts)
    t_max = int(nf_ref.max())
    assert got.shape == (256, t_max, 80, 3) and np.array_equal(nf.cpu().numpy(), nf_ref) and int(status.item()) == 0
    assert nf_ref[3] == 0 and nf_ref[4] == 0
    for b in range(256):
        t = int(nf_ref[b])
        if t:
            assert_logmel_close(got[b, :t, :, :1], ref[b, :t], what="config 5 %s utterance %d" % (layout, b))
        pad = got[b, t:]
        assert not pad.any() and not np.signbit(pad).any(), "padding of utterance %d is not +0.0" % b
    check_deltas_bit_exact(got, "config 5 " + layout, nf_ref)
    # the dense layout of the same batch gives the same bits
    dense = np.zeros((256, int(lengths.max())), np.int16)
    for i, p in enumerate(pcms):
        dense[i, :len(p)] = p
    out2, _ = fe(dev(dense), lengths=lengths, kernel="tc")
    assert torch.equal(out, out2)


def test_full_size_batch_in_several_launches():
    """B200MEL_TC_MAX_TILES: the config-2 batch cut into utterance ranges (more than one launch per
    batch, per-launch tile lists, sub-batch mask ids) must give the bits of the single launch."""
    code = textwrap.dedent("""
        import hashlib, numpy as np, torch
        from speech_recognition_b200 import FeatureFrontend
        cfg = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
                   lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
        rng = np.random.default_rng(3)
        pcm = torch.from_numpy(rng.integers(-3000, 3000, size=(256, 240000)).astype(np.int16)).cuda()
        sa = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
        for delta in (False, True):
            fe = FeatureFrontend("log-mel-spectrogram", **cfg, use_delta_accelerate=delta, spec_augment=sa)
            out, nf, st = fe(pcm, seed=43, first_utterance_id=9, kernel="tc", return_status=True)
            torch.cuda.synchronize()
            print(delta, fe.launch_count, hashlib.sha1(out.cpu().numpy().tobytes()).hexdigest(), int(nf.sum()), int(st.item()))
    """)
    outs = []
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for cap in ("0", "1000"):
        env = dict(os.environ, B200MEL_TC_MAX_TILES=cap, PYTHONPATH=root)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        outs.append([l.split() for l in r.stdout.strip().splitlines()])
    whole, chunked = outs
    for a, b in zip(whole, chunked):
        assert a[0] == b[0] and a[2:] == b[2:] and a[4] == "0"
        assert int(b[1]) > int(a[1])
