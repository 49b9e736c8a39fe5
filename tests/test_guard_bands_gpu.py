"""Out-of-bounds evidence without compute-sanitizer (closed on this GPU pool, profiles/r2j_tc_kernel.md).

Every kernel of the library is called through the C ABI on buffers that sit inside a larger poisoned
allocation:

* the outputs (features, n_frames) have guard bands on both sides; after the call the bands must hold the
  poison pattern bit for bit (no write outside the result);
* the input is the tail of its allocation's used part, followed by a band whose content - and the content
  of every row's tail behind ``lengths[b]`` - is changed between two calls: the two results must be
  identical bit for bit (nothing outside ``[0, lengths[b])`` reaches a result).

Shapes: 200 ragged utterances of up to 3.2 s = several tiles per CTA of the persistent tensor-core kernel,
a last tile with few rows, zero-fill blocks behind short utterances.
"""
import numpy as np
import pytest
import torch

from .helpers import LIBRI

pytestmark = pytest.mark.gpu

B, L = 200, 51200
BAND = 1 << 16          # floats (256 KB) either side of the features
POISON_F = 0x7FC0DEAD   # a quiet NaN with a payload: no kernel produces it
POISON_I = 0x5A5A5A5A
SA = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
SA_MFCC = dict(enable=True, F=7, m_F=2, T=40, p=1.0, m_T=2)

CASES = {
    "tc C=1": ("log-mel-spectrogram", dict(LIBRI), "tc"),
    "tc C=1 + masks": ("log-mel-spectrogram", dict(LIBRI, spec_augment=SA), "tc"),
    "tc C=3": ("log-mel-spectrogram", dict(LIBRI, use_delta_accelerate=True), "tc"),
    "tc C=3 + masks": ("log-mel-spectrogram", dict(LIBRI, use_delta_accelerate=True, spec_augment=SA), "tc"),
    "mfcc-40 C=3 + masks (mma second stage)": ("mfcc", dict(LIBRI, num_mfcc=40, use_delta_accelerate=True,
                                                            spec_augment=SA_MFCC), "tc"),
    "mfcc-40 C=1": ("mfcc", dict(LIBRI, num_mfcc=40), "tc"),
    "mfcc-13 (scalar second stage)": ("mfcc", dict(LIBRI, num_mfcc=13), "tc"),
    "generic libri C=3 + masks": ("log-mel-spectrogram", dict(LIBRI, use_delta_accelerate=True, spec_augment=SA),
                                  "generic"),
    "generic 400/160/512 M=40": ("log-mel-spectrogram",
                                 dict(sample_rate=16000, frame_length=400, frame_step=160, fft_length=512,
                                      num_mel_bins=40, lower_edge_hertz=20.0, upper_edge_hertz=8000.0), "auto"),
    "spectrogram 320/160/320": ("spectrogram", dict(sample_rate=16000, frame_length=320, frame_step=160,
                                                    fft_length=320), "auto"),
    "spectrogram 128/64/256": ("spectrogram", dict(sample_rate=16000, frame_length=128, frame_step=64,
                                                   fft_length=256), "auto"),
}


@pytest.fixture(scope="module")
def pkg(built_lib):
    import speech_recognition_b200 as p

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return p


@pytest.fixture(scope="module")
def audio():
    rng = np.random.default_rng(11)
    x = np.clip(np.rint(rng.standard_normal((B, L)) * 3000.0), -32768, 32767).astype(np.int16)
    lengths = rng.integers(L // 2, L + 1, B).astype(np.int32)
    lengths[0], lengths[1], lengths[2] = L, 320 + 160 * 130, 320 + 160 * 127
    return x, lengths


def _banded(shape, dtype, poison, dev):
    """A tensor of `shape` in the middle of a poisoned allocation; returns (view, whole, lo, hi)."""
    n = int(np.prod(shape))
    whole = torch.full((n + 2 * BAND,), poison, dtype=torch.int32, device=dev)
    view = whole[BAND:BAND + n].view(dtype).view(shape)
    return view, whole, n


def _bands_intact(whole, n, poison):
    lo = whole[:BAND]
    hi = whole[BAND + n:]
    return bool((lo == poison).all().item()) and bool((hi == poison).all().item())


def _pcm_at_end(x, lengths, tail_value, dtype, dev):
    """The batch as the LAST bytes before a band of `tail_value`; samples behind lengths[b] = tail_value."""
    xx = x.copy()
    for b in range(B):
        xx[b, lengths[b]:] = tail_value
    band = np.full(BAND, tail_value, np.int16)
    flat = np.concatenate([xx.reshape(-1), band])
    t = torch.from_numpy(flat).to(dev)
    if dtype == torch.float32:
        t = t.to(torch.float32) / 32768.0
    return t, t[:B * L].view(B, L)


@pytest.mark.parametrize("name", list(CASES))
def test_no_write_outside_the_result_and_no_read_outside_the_utterance(pkg, audio, name):
    ftype, kw, kernel = CASES[name]
    x, lengths = audio
    dev = torch.device("cuda:0")
    fe = pkg.FeatureFrontend(ftype, **kw)
    if kernel == "tc":
        assert fe.tc_supported
    t_max = fe.num_frames(L)
    len_d = torch.from_numpy(lengths).to(dev)
    results = []
    for tail in (0, 12345):
        out, whole, n = _banded((B, t_max, fe.frequency_dim, fe.feature_dim), torch.float32, POISON_F, dev)
        nf, nf_whole, nn = _banded((B,), torch.int32, POISON_I, dev)
        keep, pcm = _pcm_at_end(x, lengths, tail, torch.int16, dev)
        got, got_nf, st = fe(pcm, lengths=len_d, t_max=t_max, seed=43, first_utterance_id=7, out=out, n_frames=nf,
                             kernel=kernel, return_status=True)
        torch.cuda.synchronize()
        assert int(st.item()) == 0, "status word %#x" % int(st.item())
        assert _bands_intact(whole, n, POISON_F), "%s wrote outside the feature tensor" % name
        assert _bands_intact(nf_whole, nn, POISON_I), "%s wrote outside n_frames" % name
        assert not torch.isnan(got).any(), "poison left inside the result: a part of it was not written"
        results.append((got.clone(), got_nf.clone()))
        del keep
    assert torch.equal(results[0][0].view(torch.int32), results[1][0].view(torch.int32)), \
        "%s: samples behind lengths[b] (or behind the buffer) reached a result" % name
    assert torch.equal(results[0][1], results[1][1])
    want_nf = np.maximum(0, 1 + (lengths.astype(np.int64) - fe.frame_length) // fe.frame_step)
    assert np.array_equal(results[0][1].cpu().numpy(), want_nf)


def test_float32_entry_and_ragged_offsets_stay_inside(pkg, audio):
    """The float32 entry (conversion kernel + gated pair of kernels) and the ragged (offsets) entry."""
    x, lengths = audio
    dev = torch.device("cuda:0")
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    t_max = fe.num_frames(L)
    len_d = torch.from_numpy(lengths).to(dev)
    want, _ = fe(torch.from_numpy(x).to(dev), lengths=len_d, t_max=t_max, kernel="tc")
    for tail in (0, 12345):
        out, whole, n = _banded((B, t_max, 80, 1), torch.float32, POISON_F, dev)
        keep, pcm = _pcm_at_end(x, lengths, tail, torch.float32, dev)
        got, _ = fe(pcm, lengths=len_d, t_max=t_max, out=out)
        torch.cuda.synchronize()
        assert _bands_intact(whole, n, POISON_F)
        assert torch.equal(got, want)
    # packed utterances, unaligned starts: the last one ends exactly at the end of the buffer
    offs = np.zeros(B, np.int64)
    offs[1:] = np.cumsum(lengths[:-1].astype(np.int64))
    flat = np.concatenate([x[b, :lengths[b]] for b in range(B)] + [np.full(BAND, 12345, np.int16)])
    flat_d = torch.from_numpy(flat).to(dev)
    total = int(offs[-1] + lengths[-1])
    out, whole, n = _banded((B, t_max, 80, 1), torch.float32, POISON_F, dev)
    got, _ = fe(flat_d[:total], lengths=len_d, offsets=torch.from_numpy(offs).to(dev), t_max=t_max, out=out, kernel="tc")
    torch.cuda.synchronize()
    assert _bands_intact(whole, n, POISON_F)
    assert torch.equal(got, want)


def test_rows_either_side_stay_inside(pkg, audio):
    """ingest (N1) and the padding mask (N3) with guard bands around their outputs."""
    from speech_recognition_b200 import batching, ingest

    x, lengths = audio
    dev = torch.device("cuda:0")
    files = [x[b, :lengths[b]].tobytes() for b in range(16)]
    got, glen = ingest.ingest_file_images(files, "pcm", device=dev)
    assert got.shape[0] == 16
    for b in range(16):
        assert torch.equal(got[b, :lengths[b]].cpu(), torch.from_numpy(x[b, :lengths[b]]))
        assert not got[b, lengths[b]:].any()
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    feats, _ = fe(torch.from_numpy(x[:16]).to(dev), lengths=torch.from_numpy(lengths[:16]).to(dev))
    m = batching.audio_mask(feats, [11, 11], [2, 1])
    assert m.dtype == torch.bool or m.dtype == torch.uint8


def test_time_warp_stays_inside(pkg, audio):
    """b200mel_time_warp (N4): guard bands around its output, padding rows written as +0.0."""
    import ctypes

    from speech_recognition_b200 import _lib

    dev = torch.device("cuda:0")
    Bw, T, v = 6, 300, 80
    nf = torch.tensor([300, 299, 150, 81, 2, 0], dtype=torch.int32, device=dev)
    feat = torch.randn((Bw, T, v, 1), device=dev)
    out, whole, n = _banded((Bw, T, v, 1), torch.float32, POISON_F, dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = _lib.load().b200mel_time_warp(ctypes.c_void_p(feat.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                                       ctypes.c_void_p(nf.data_ptr()), Bw, T, v, 1, 40, None, 43, 0,
                                       ctypes.c_void_p(st.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert rc == 0, _lib.last_error()
    assert _bands_intact(whole, n, POISON_F)
    assert not torch.isnan(out).any()
    for b in range(Bw):
        assert not out[b, int(nf[b]):].any()
