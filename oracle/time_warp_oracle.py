"""TEST INFRASTRUCTURE - CPU restatement of SpecAugment's time warp.

Reference: speech_recognition/data.py:275-280

    src_loc_time = tf.random.uniform((), W, num_time - W, tf.int32)
    dst_loc_time = src_loc_time + tf.random.uniform((), -W, W, tf.int32)
    src_loc = tf.cast(tf.reshape([src_loc_time, v // 2], [1, 1, 2]), tf.float32)
    dst_loc = tf.cast(tf.reshape([dst_loc_time, v // 2], [1, 1, 2]), tf.float32)
    audio, _ = tfa.image.sparse_image_warp(audio, src_loc, dst_loc, num_boundary_points=3)

`tensorflow-addons` is an un-vendored, unpinned dependency of the reference (requirements.txt) and is not
installable here: PARITY UNPINNED.  What follows restates the published algorithm of
tfa.image.sparse_image_warp (tensorflow_addons/image/sparse_image_warp.py, interpolate_spline.py,
dense_image_warp.py, v0.11-v0.23: unchanged in that range) as the reference calls it:

* image = the [T, v, C] feature tensor (height = time, width = feature bins), one batch entry;
* num_boundary_points = 3 -> boundary_points_per_edge = 2 -> the 12 boundary points of a 4 x 4 grid
  (linspace(0, T - 1, 4) x linspace(0, v - 1, 4)) are added as zero-flow control points, 13 in all
  (the reference's docstring speaks of 8; the code passes 3);
* flows = dest - source at the control points; interpolate_spline(train_points = dest points, train_values =
  flows, query = every pixel (y, x), order = 2, regularization 0): polyharmonic spline with
  phi(r) = 0.5 r log(max(r, 1e-10)) of the SQUARED distance r, solved as [[A, B], [B^T, 0]] [w; v] = [f; 0],
  B = [c, 1];
* dense_image_warp: output[y, x] = image sampled at (y, x) - flow(y, x), bilinear, floor clipped to
  [0, size - 2], alpha clipped to [0, 1].

The frequency component of every control flow is zero, so the x-flow is identically zero: the warp moves
frames along time only, by an amount that depends on (t, bin).

Draws: the framework's counter-based stream (oracle/logmel_oracle.py: specaug_raw_draw) with stream id 1 in
the second counter word - the mask draws (stream 0) are the same with or without the warp.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np

from .logmel_oracle import philox4x32_10

EPSILON = 1e-10


def warp_raw_draw(seed: int, utterance_id: int, draw_index: int) -> int:
    seed &= 0xFFFFFFFFFFFFFFFF
    utterance_id &= 0xFFFFFFFFFFFFFFFF
    return philox4x32_10((draw_index, 1, utterance_id & 0xFFFFFFFF, utterance_id >> 32),
                         (seed & 0xFFFFFFFF, seed >> 32))[0]


def warp_draw_points(num_time: int, W: int, seed: int = 0, utterance_id: int = 0,
                     raw: Optional[Sequence[int]] = None) -> Tuple[int, int]:
    """(src_loc_time, dst_loc_time) of data.py:276-277; raises like tf.random.uniform when maxval <= minval."""
    if num_time - W <= W:
        raise ValueError("Need minval < maxval, got %d >= %d" % (W, num_time - W))
    u0 = int(raw[0]) if raw is not None else warp_raw_draw(seed, utterance_id, 0)
    u1 = int(raw[1]) if raw is not None else warp_raw_draw(seed, utterance_id, 1)
    src = W + u0 % (num_time - 2 * W)
    dst = src + (-W + u1 % (2 * W))
    return src, dst


def boundary_points(height: int, width: int, per_edge: int = 2) -> np.ndarray:
    ys, xs = np.meshgrid(np.linspace(0.0, height - 1.0, per_edge + 2), np.linspace(0.0, width - 1.0, per_edge + 2),
                         indexing="ij")
    is_b = (xs == 0) | (xs == width - 1) | (ys == 0) | (ys == height - 1)
    return np.stack([ys[is_b], xs[is_b]], axis=-1)


def _phi(r):
    return 0.5 * r * np.log(np.maximum(r, EPSILON))


def control_points(num_time: int, v: int, src_t: int, dst_t: int):
    b = boundary_points(num_time, v)
    src = np.concatenate([[[float(src_t), float(v // 2)]], b])
    dst = np.concatenate([[[float(dst_t), float(v // 2)]], b])
    return dst, dst - src          # train points, flows


def solve_spline(c: np.ndarray, f: np.ndarray):
    n, d = c.shape
    diff = c[:, None, :] - c[None, :, :]
    a = _phi((diff * diff).sum(-1))
    bm = np.concatenate([c, np.ones((n, 1))], 1)
    lhs = np.block([[a, bm], [bm.T, np.zeros((d + 1, d + 1))]])
    rhs = np.concatenate([f, np.zeros((d + 1, f.shape[1]))], 0)
    wv = np.linalg.solve(lhs, rhs)
    return wv[:n], wv[n:]


def spline_flow(q: np.ndarray, c: np.ndarray, w: np.ndarray, v: np.ndarray) -> np.ndarray:
    diff = q[:, None, :] - c[None, :, :]
    return _phi((diff * diff).sum(-1)) @ w + np.concatenate([q, np.ones((q.shape[0], 1))], 1) @ v


def time_flow(num_time: int, v: int, src_t: int, dst_t: int) -> np.ndarray:
    """[T, v] float64: the time component of the dense flow (the frequency component is zero)."""
    c, f = control_points(num_time, v, src_t, dst_t)
    w, lin = solve_spline(c, f)
    ys, xs = np.meshgrid(np.arange(num_time, dtype=np.float64), np.arange(v, dtype=np.float64), indexing="ij")
    q = np.stack([ys.ravel(), xs.ravel()], -1)
    fl = spline_flow(q, c, w, lin)
    assert np.abs(fl[:, 1]).max() < 1e-6, "frequency flow must vanish"
    return fl[:, 0].reshape(num_time, v)


def sparse_image_warp_time(feat: np.ndarray, src_t: int, dst_t: int, dtype=np.float64) -> np.ndarray:
    """feat [T, v, C] -> warped [T, v, C] (the `audio` of data.py:280)."""
    x = np.asarray(feat, dtype=dtype)
    T, v, _ = x.shape
    if T < 2 or v < 2:
        raise ValueError("the bilinear sampler needs at least 2 x 2 pixels")
    flow = time_flow(T, v, src_t, dst_t).astype(dtype)
    qy = np.arange(T, dtype=dtype)[:, None] - flow
    fy = np.clip(np.floor(qy), 0, T - 2)
    ay = np.clip(qy - fy, 0.0, 1.0).astype(dtype)
    fy = fy.astype(np.int64)
    qx = np.broadcast_to(np.arange(v, dtype=dtype)[None, :], (T, v))
    fx = np.clip(np.floor(qx), 0, v - 2)
    ax = np.clip(qx - fx, 0.0, 1.0).astype(dtype)
    fx = fx.astype(np.int64)
    tl, tr = x[fy, fx], x[fy, fx + 1]
    bl, br = x[fy + 1, fx], x[fy + 1, fx + 1]
    top = ax[..., None] * (tr - tl) + tl
    bot = ax[..., None] * (br - bl) + bl
    return (ay[..., None] * (bot - top) + top).astype(dtype)


def time_warp(feat: np.ndarray, W: int, seed: int = 0, utterance_id: int = 0, raw=None, dtype=np.float64) -> np.ndarray:
    src, dst = warp_draw_points(feat.shape[0], W, seed, utterance_id, raw)
    return sparse_image_warp_time(feat, src, dst, dtype)
