"""Batch contract between the front-end and the models.

The reference applies a *max-over policy* to single examples (``filter_example`` / ``slice_example``,
data.py:331-354, selected at run/train.py:119-130), pads to a fixed shape with zeros
(``padded_batch``, run/train.py:182-197), and each model then recovers the padding mask by scanning the
whole batch for non-zero frames and reducing it by its convolution strides (``_audio_mask``,
models/las.py:205-217, models/deepspeech2.py:68-78).  Here the policy is applied to the batch on the
device: ``slice`` is the front-end's ``t_max`` clip, ``filter`` a keep-mask over ``n_frames`` / token
lengths, and the mask is one streaming kernel over the padded batch.
"""
import ctypes
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def filter_example(max_audio_length: int, max_token_length: int):
    """Batched ``filter_example`` (data.py:331-342): ``_wrapper(n_frames, token_lengths)`` -> bool keep-mask
    ``n_frames <= max_audio_length and token_lengths <= max_token_length``, on the inputs' device."""

    def _wrapper(n_frames: torch.Tensor, token_lengths: torch.Tensor) -> torch.Tensor:
        return torch.logical_and(n_frames <= max_audio_length, token_lengths <= max_token_length)

    return _wrapper


def slice_example(max_audio_length: int, max_token_length: int):
    """Batched ``slice_example`` (data.py:345-354): ``_wrapper(features [B, T, F, C], tokens [B, N])`` ->
    ``(features[:, :max_audio_length], tokens[:, :max_token_length])``.  With ``FeatureFrontend`` pass
    ``t_max=min(t_max, max_audio_length)`` instead and the frames are never computed."""

    def _wrapper(features: torch.Tensor, tokens: torch.Tensor):
        return features[:, :max_audio_length], tokens[:, :max_token_length]

    return _wrapper


def audio_mask_length(t_max: int, time_kernel_sizes: Sequence[int], time_strides: Sequence[int]) -> Tuple[int, int]:
    """(reduced length, stride product) of ``_audio_mask`` for ``t_max`` padded frames: per conv layer
    ``t = (t - (kernel - stride)) // stride`` (models/las.py:209-213, models/deepspeech2.py:71-73)."""
    k = np.ascontiguousarray(time_kernel_sizes, dtype=np.int32)
    s = np.ascontiguousarray(time_strides, dtype=np.int32)
    if k.shape != s.shape or k.ndim != 1:
        raise ValueError("time_kernel_sizes and time_strides must be equally long lists")
    reduced, group = ctypes.c_int64(), ctypes.c_int32()
    _lib.check(_lib.load().b200mel_audio_mask_length(int(t_max), k.ctypes.data_as(ctypes.c_void_p),
                                                     s.ctypes.data_as(ctypes.c_void_p), int(k.size),
                                                     ctypes.byref(reduced), ctypes.byref(group)))
    return reduced.value, group.value


def audio_mask(features: torch.Tensor, time_kernel_sizes: Sequence[int], time_strides: Sequence[int]) -> torch.Tensor:
    """``_audio_mask`` (models/las.py:205-217 with ``[k, k]`` / ``[s, s]``; models/deepspeech2.py:68-78 with the
    time components of its layers) on a padded batch ``[B, Tmax, F, C]`` float32 CUDA: bool ``[B, reduced]``,
    True where any frame of the stride group holds a non-zero value."""
    if not features.is_cuda:
        raise RuntimeError("audio_mask runs on CUDA tensors only: there is no CPU path")
    if features.dtype != torch.float32 or features.dim() < 2:
        raise ValueError("features must be float32 [B, Tmax, ...]")
    features = features.contiguous()
    B, T = features.shape[:2]
    row = int(np.prod(features.shape[2:])) if features.dim() > 2 else 1
    reduced, group = audio_mask_length(T, time_kernel_sizes, time_strides)
    mask = torch.empty((B, reduced), dtype=torch.uint8, device=features.device)
    stream = torch.cuda.current_stream(features.device).cuda_stream
    _lib.check(_lib.load().b200mel_audio_mask(_ptr(features), B, T, row, group, reduced, _ptr(mask),
                                              ctypes.c_void_p(stream)))
    return mask.view(torch.bool)  # 0 / 1 bytes: a view, no second kernel


def assemble_batch(frontend, pcm, lengths, token_lengths: Optional[torch.Tensor] = None,
                   max_audio_length: Optional[int] = None, max_token_length: Optional[int] = None,
                   max_over_policy: Optional[str] = None, audio_pad_length: Optional[int] = None,
                   mask_layers: Optional[Tuple[Sequence[int], Sequence[int]]] = None, **forward_kwargs):
    """Front-end + max-over policy + padding + model mask in one call.

    ``max_over_policy`` ∈ {None, "slice", "filter"} (run/train.py:119-130).  ``audio_pad_length`` fixes
    ``Tmax`` like ``padded_batch`` with ``get_batching_shape`` (run/train.py:182-197).  Returns a dict with
    ``features`` [B, Tmax, F, C], ``n_frames`` (clipped to Tmax under "slice"), ``keep`` (bool [B], all True
    unless "filter") and, if ``mask_layers = (time_kernel_sizes, time_strides)``, ``mask`` [B, reduced]."""
    if max_over_policy not in (None, "slice", "filter"):
        raise ValueError("max_over_policy must be None, 'slice' or 'filter'")
    t_max = forward_kwargs.pop("t_max", None)
    if audio_pad_length is not None:
        t_max = int(audio_pad_length)
    if max_over_policy == "slice" and max_audio_length is not None:
        if t_max is None:
            host = lengths.cpu().numpy() if isinstance(lengths, torch.Tensor) else np.asarray(lengths)
            t_max = frontend.num_frames(int(host.max())) if host.size else 0
        t_max = min(int(t_max), int(max_audio_length))
    feats, n_frames = frontend(pcm, lengths=lengths, t_max=t_max, **forward_kwargs)
    if audio_pad_length is not None and feats.shape[1] < int(audio_pad_length):
        # "slice" below the padded length: the reference slices each example (data.py:345-354) and
        # padded_batch then pads to audio_pad_length (run/train.py:182-197) - zeros behind the slice
        feats = torch.nn.functional.pad(feats, (0, 0, 0, 0, 0, int(audio_pad_length) - feats.shape[1]))
    out = {"features": feats, "n_frames": n_frames}
    keep = torch.ones_like(n_frames, dtype=torch.bool)
    if max_over_policy == "filter":
        tl = token_lengths if token_lengths is not None else torch.zeros_like(n_frames)
        keep = filter_example(max_audio_length if max_audio_length is not None else 2 ** 31 - 1,
                              max_token_length if max_token_length is not None else 2 ** 31 - 1)(
            n_frames, tl.to(n_frames.device))
    elif max_over_policy == "slice":
        out["n_frames"] = torch.clamp(n_frames, max=int(t_max))
    out["keep"] = keep
    if mask_layers is not None:
        out["mask"] = audio_mask(feats, mask_layers[0], mask_layers[1])
    return out
