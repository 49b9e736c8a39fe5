"""b200mel_forward_mfcc_noalloc (include/b200mel.h): the MFCC path with its log-mel scratch in a static
device array, for callers that may not allocate.  No counterpart in the reference,
so the checker is the existing entry point, b200mel_forward, on the same inputs:
both run the same two kernels, the outputs must agree within the MFCC tolerance the parity tests use
(they are in fact identical).  Two calls in a row: the second, on a SMALLER batch with other audio,
finds the first call's rows in the shared array and must not be affected by them.

This file is collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

from .helpers import LIBRI

pytestmark = pytest.mark.gpu

MFCC_TOL = 1e-3 * np.sqrt(160.0)    # tests/test_parity_fullsize_gpu.py: what a 1e-3 log-mel error allows
SA = dict(enable=True, F=7, m_F=2, T=20, p=1.0, m_T=2)
NOALLOC_BYTES = 16 << 20            # B200MEL_NOALLOC_SCRATCH_BYTES


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else None)


def _noalloc(pkg, fe, pcm, lengths, t_max, seed, first_id):
    """The new entry point through ctypes, on torch's current stream; no allocation inside the call."""
    from speech_recognition_b200 import _lib

    lib = _lib.load()
    batch, stride = pcm.shape
    out = torch.empty((batch, t_max, fe.frequency_dim, fe.feature_dim), dtype=torch.float32, device=pcm.device)
    n_frames = torch.empty(batch, dtype=torch.int32, device=pcm.device)
    status = torch.zeros(1, dtype=torch.int32, device=pcm.device)
    assert out.data_ptr() % 32 == 0
    rc = lib.b200mel_forward_mfcc_noalloc(fe._handle, _ptr(pcm), None, _ptr(lengths), batch, stride, _ptr(out), t_max,
                                          _ptr(n_frames), None, seed, first_id, _ptr(status),
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, out, n_frames, status


def _batch(rng, batch, n, ragged, tiny_last):
    x = np.clip(np.rint(rng.standard_normal((batch, n)) * 3000.0), -32768, 32767).astype(np.int16)
    lengths = np.full(batch, n, np.int32)
    if ragged:
        lengths = rng.integers(n // 3, n + 1, batch).astype(np.int32)
        lengths[0] = n
        # shorter than one frame: every row is padding (with masks: 31 frames - a time-mask draw on an empty
        # utterance is the reference's error case, covered in tests/test_parity_gpu.py)
        lengths[-1] = 100 if tiny_last else 320 + 160 * 30
    return torch.from_numpy(x).cuda(), torch.from_numpy(lengths).cuda()


@pytest.mark.parametrize("delta,masks", [(True, True), (False, False)])
def test_noalloc_entry_matches_forward_twice(built_lib, delta, masks):
    import speech_recognition_b200 as pkg
    from speech_recognition_b200 import _lib

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    fe = pkg.FeatureFrontend("mfcc", **LIBRI, num_mfcc=40, use_delta_accelerate=delta, spec_augment=SA if masks else None)
    assert fe.tc_supported
    rng = np.random.default_rng(77)
    # call 1: 24 utterances x 6 s (ragged); call 2: 5 utterances x 2.5 s of other audio, same array
    for call, (batch, n, ragged) in enumerate([(24, 96000, True), (5, 40000, False)]):
        pcm, lengths = _batch(rng, batch, n, ragged, tiny_last=not masks)
        t_max = fe.num_frames(n)
        seed, first_id = 43, 1000 * (call + 1)
        want, want_nf, want_status = fe(pcm, lengths=lengths, t_max=t_max, seed=seed, first_utterance_id=first_id,
                                        kernel="tc", return_status=True)
        rc, got, nf, status = _noalloc(pkg, fe, pcm, lengths, t_max, seed, first_id)
        assert rc == 0, _lib.load().b200mel_last_error().decode()
        torch.cuda.synchronize()
        assert int(status.item()) == int(want_status.item()) == 0
        assert torch.equal(nf, want_nf)
        g, w = got.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(g).all()
        np.testing.assert_allclose(g, w, rtol=0.0, atol=MFCC_TOL, err_msg="call %d" % (call + 1))
        assert np.array_equal(g == 0.0, w == 0.0), "call %d: masked / padding zeros differ" % (call + 1)
        assert w[0].any(), "the comparison is not vacuous"


def test_noalloc_entry_rejects_what_does_not_fit(built_lib):
    import speech_recognition_b200 as pkg
    from speech_recognition_b200 import _lib

    lib = _lib.load()
    fe = pkg.FeatureFrontend("mfcc", **LIBRI, num_mfcc=40)
    max_rows = NOALLOC_BYTES // 4 // 80
    t_max = 1499
    batch = max_rows // t_max + 1                           # one utterance too many
    pcm = torch.zeros((batch, 240000), dtype=torch.int16, device="cuda")
    lengths = torch.full((batch,), 240000, dtype=torch.int32, device="cuda")
    rc, _, _, _ = _noalloc(pkg, fe, pcm, lengths, t_max, 0, 0)
    assert rc != 0
    msg = lib.b200mel_last_error().decode()
    assert "exceeds the static scratch" in msg and str(batch * t_max) in msg
    rc, out, nf, _ = _noalloc(pkg, fe, pcm[:batch - 1], lengths[:batch - 1], t_max, 0, 0)   # the largest batch that fits
    assert rc == 0, lib.b200mel_last_error().decode()
    torch.cuda.synchronize()
    assert (nf.cpu().numpy() == t_max).all() and np.isfinite(out[-1].cpu().numpy()).all()
    # a log-mel handle has no scratch stage: refused with a clear message
    fe2 = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    rc, _, _, _ = _noalloc(pkg, fe2, pcm[:2], lengths[:2], t_max, 0, 0)
    assert rc != 0 and "MFCC" in lib.b200mel_last_error().decode()
