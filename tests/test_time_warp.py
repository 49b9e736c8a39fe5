"""SpecAugment time warp (data.py:275-280).

CPU: properties of the oracle's restatement of tfa.image.sparse_image_warp (PARITY UNPINNED: tensorflow-addons
is not vendored by the reference and not installable here; oracle/time_warp_oracle.py states what it follows).
GPU: b200mel_time_warp against that oracle, and the staged front-end (feature -> warp -> masks -> deltas).
Tolerance: the flow is a sum of 13 terms of size ~1e2..1e3 that cancel to a few frames, evaluated in float32 on
the device (as in TF) and in float64 by the oracle: |delta flow| <= 2e-3 frames, so
|delta out| <= 2e-3 * max |x[t + 1] - x[t]| + 1e-5 |x|; the tests use smooth features.
"""
import numpy as np
import pytest

from oracle import logmel_oracle as O
from oracle import time_warp_oracle as TW

from .helpers import LIBRI, noise_int16


def smooth_features(T, v, C=1, seed=0):
    rng = np.random.default_rng(seed)
    t = np.arange(T)[:, None, None] / 37.0
    f = np.arange(v)[None, :, None] / 11.0
    ph = rng.uniform(0, 6.28, (1, 1, C))
    return (np.sin(t + ph) * np.cos(f) + 0.3 * np.sin(0.31 * t * f + ph) - 2.0).astype(np.float32)


def test_oracle_identity_boundaries_and_control_point():
    x = smooth_features(300, 80, 2)
    assert np.array_equal(TW.sparse_image_warp_time(x, 100, 100, np.float32), x)      # zero flow: exact copy
    flow = TW.time_flow(300, 80, 100, 120)
    assert abs(flow[120, 40] - 20.0) < 1e-9                       # the control flow itself
    b = TW.boundary_points(300, 80)
    assert b.shape == (12, 2)                                     # num_boundary_points = 3 -> 4 x 4 grid, 12 on the edge
    for y, x_ in b:
        if float(y).is_integer() and float(x_).is_integer():
            assert abs(flow[int(y), int(x_)]) < 1e-9              # zero flow at the boundary points
    q = np.arange(300)[:, None] - flow
    assert (np.diff(q, axis=0) > 0).all()                         # frames keep their order
    y = TW.sparse_image_warp_time(x, 100, 120)
    # the pixel at the destination shows what was at the source
    assert np.allclose(y[120, 40], x[100, 40], atol=1e-6)


def test_oracle_draws_follow_the_reference_ranges():
    for utt in range(50):
        src, dst = TW.warp_draw_points(400, 40, seed=43, utterance_id=utt)
        assert 40 <= src < 360 and -40 <= dst - src < 40          # uniform(W, T - W), src + uniform(-W, W)
    assert TW.warp_draw_points(400, 40, raw=[5, 7]) == (40 + 5, 45 - 40 + 7)
    with pytest.raises(ValueError):                               # tf.random.uniform: maxval <= minval
        TW.warp_draw_points(80, 40)
    # the mask stream (stream 0) does not move when the warp draws (stream 1) are taken
    assert TW.warp_raw_draw(43, 7, 0) != O.specaug_raw_draw(43, 7, 0)


def test_oracle_float32_follows_float64():
    x = smooth_features(500, 80)
    a = TW.sparse_image_warp_time(x, 200, 170, np.float64)
    b = TW.sparse_image_warp_time(x, 200, 170, np.float32)
    assert np.abs(a - b).max() < 1e-3


# ------------------------------------------------------------------------------------------------------------
gpu = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe_mod(built_lib):
    import torch

    import speech_recognition_b200 as p

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return p


def _tol(x):
    return 2e-3 * np.abs(np.diff(x, axis=0)).max() + 1e-5 * np.abs(x).max()


@gpu
@pytest.mark.parametrize("C", [1, 3])
def test_time_warp_matches_the_oracle(fe_mod, C):
    import torch

    from speech_recognition_b200 import frontend as F_

    T, v, W = 640, 80, 40
    lens = [640, 400, 81, 2, 300, 80]                             # 81 = the shortest the draw accepts; 2 and 80 are refused
    feats = np.zeros((len(lens), T, v, C), np.float32)
    for b, n in enumerate(lens):
        feats[b, :n] = smooth_features(n, v, C, seed=b)
    out, st = F_.time_warp(torch.from_numpy(feats).cuda(), lens, W, seed=43, first_utterance_id=11)
    out = out.cpu().numpy()
    assert int(st.item()) == 1                                    # utterances 3 and 5: T - W <= W, TF raises
    for b, n in enumerate(lens):
        assert not out[b, n:].any()
        if n - W <= W:
            assert np.array_equal(out[b, :n], feats[b, :n])       # copied unchanged
            continue
        src, dst = TW.warp_draw_points(n, W, 43, 11 + b)
        want = TW.sparse_image_warp_time(feats[b, :n], src, dst)
        assert np.abs(out[b, :n] - want).max() <= _tol(feats[b, :n]), (b, n, src, dst)
    # explicit control points (e.g. recorded TF draws) through warp_table
    tbl = torch.tensor([[100, 130]] * len(lens), dtype=torch.int32)
    out2, _ = F_.time_warp(torch.from_numpy(feats[:2]).cuda(), lens[:2], 0, warp_table=tbl[:2])
    want = TW.sparse_image_warp_time(feats[1, :400], 100, 130)
    assert np.abs(out2[1, :400].cpu().numpy() - want).max() <= _tol(feats[1, :400])
    # src == dst: zero flow, bit-exact copy
    out3, _ = F_.time_warp(torch.from_numpy(feats[:1]).cuda(), lens[:1], 0, warp_table=torch.tensor([[77, 77]], dtype=torch.int32))
    assert np.array_equal(out3[0].cpu().numpy(), feats[0])


@gpu
def test_frontend_with_time_warp_runs_the_reference_stage_order(fe_mod):
    """feature -> warp -> masks -> deltas (run/train.py:88-116 with W set): the staged front-end against the
    same stages done by hand with the oracle's warp and mask draws."""
    import torch

    sa = dict(enable=True, W=20, F=15, m_F=2, T=30, p=1.0, m_T=2)
    fe = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True, spec_augment=sa)
    plain = fe_mod.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    lens = [48000, 30000, 16000]
    pcm = np.zeros((3, 48000), np.int16)
    for i, n in enumerate(lens):
        pcm[i, :n] = noise_int16(n, 20 + i)
    got, nf, st = fe(torch.from_numpy(pcm).cuda(), lengths=lens, seed=43, first_utterance_id=5, return_status=True)
    assert int(st.item()) == 0 and got.shape == (3, 299, 80, 3)
    base, nf0 = plain(torch.from_numpy(pcm).cuda(), lengths=lens)
    base = base.cpu().numpy()
    got = got.cpu().numpy()
    for b in range(3):
        n = int(nf[b])
        src, dst = TW.warp_draw_points(n, 20, 43, 5 + b)
        x = TW.sparse_image_warp_time(base[b, :n], src, dst).astype(np.float32)
        table = O.specaug_draw_masks(n, 80, 15, 2, 30, 1.0, 2, seed=43, utterance_id=5 + b)
        x = O.spec_augment_apply(x, table[:2], table[2:])
        want = O.delta_accelerate(x)
        tol = 3 * (2e-3 * np.abs(np.diff(base[b, :n], axis=0)).max() + 1e-4)    # deltas: differences of warped rows
        assert np.abs(got[b, :n] - want).max() <= tol, b
        zero_w, zero_g = want[..., 0] == 0, got[b, :n, :, 0] == 0
        assert np.array_equal(zero_w, zero_g)                     # the masks sit where the oracle's draws put them
        assert not got[b, n:].any()


@gpu
def test_spec_augment_factory_with_W(fe_mod):
    import torch

    x = torch.from_numpy(smooth_features(300, 80))
    fn = fe_mod.spec_augment(80, W=30, F=None, m_F=None, T=None, p=None, m_T=None)
    y = fn(x.cuda(), utterance_id=3, seed=9).cpu().numpy()
    src, dst = TW.warp_draw_points(300, 30, 9, 3)
    want = TW.sparse_image_warp_time(x.numpy(), src, dst)
    assert np.abs(y - want).max() <= _tol(x.numpy())
    with pytest.raises(ValueError):                               # TF raises when num_time - W <= W
        fe_mod.spec_augment(80, W=30)(x[:50].cuda(), utterance_id=0, seed=0)
