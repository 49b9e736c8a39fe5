"""FeatureFrontend: the batched, fused form of the reference's feature maps.

The reference maps a per-example callable over a tf.data pipeline and pads afterwards
(run/train.py:88-116, 192-197).  This class does the same work for a whole batch in one kernel
launch on a B200: feature (data.py:122-241) -> SpecAugment masks (data.py:282-301) -> delta
(data.py:310-328) -> zero padding to Tmax, writing only [B, Tmax, F, C] to HBM.

torch is used for device memory and streams only; the arithmetic is in lib/libb200mel.so.
"""
import ctypes
from typing import Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _lib


def _sa_get(sa, key):
    if sa is None:
        return None
    if isinstance(sa, dict):
        return sa.get(key)
    return getattr(sa, key, None)


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


class FeatureFrontend:
    """One immutable plan (window, twiddles, float32 mel filterbank in TF's op order, DCT matrix)
    on one CUDA device.  Parameter names are the YAML keys of ``resources/configs/*_config.yml``
    (configs/data_config.py:23-54)."""

    def __init__(self, audio_feature_type: str, sample_rate: int, frame_length: int, frame_step: int,
                 fft_length: Optional[int] = None, num_mel_bins: Optional[int] = None,
                 num_mfcc: Optional[int] = None, lower_edge_hertz: Optional[float] = None,
                 upper_edge_hertz: Optional[float] = None, epsilon: float = 1e-12,
                 use_delta_accelerate: bool = False, spec_augment=None,
                 device: Union[None, int, str, torch.device] = None):
        if audio_feature_type not in _lib.FEATURE_TYPES:
            raise ValueError("audio_feature_type must be one of %s" % sorted(_lib.FEATURE_TYPES))
        if not torch.cuda.is_available():
            raise RuntimeError("speech_recognition_b200 needs a CUDA device: there is no CPU fallback")
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("FeatureFrontend runs on CUDA devices only (got %s)" % dev)
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        sa_on = bool(_sa_get(spec_augment, "enable")) if spec_augment is not None else False
        # SpecAugment time warping (data.py:275-280; W is null in every shipped config): the warp sits between
        # the feature and the masks, so the stages run one after the other - feature (the fused kernel of an
        # inner front-end without masks and deltas), b200mel_time_warp, masks, deltas - instead of in one launch
        self._warp_W = int(_sa_get(spec_augment, "W") or 0) if sa_on else 0
        self._inner = None
        if self._warp_W:
            self._inner = FeatureFrontend(audio_feature_type, sample_rate, frame_length, frame_step, fft_length,
                                          num_mel_bins, num_mfcc, lower_edge_hertz, upper_edge_hertz, epsilon,
                                          use_delta_accelerate=False, spec_augment=None, device=dev)
            self._handle = None
            self._sa = {k: _sa_get(spec_augment, k) for k in ("F", "m_F", "T", "p", "m_T")}
            self._delta = bool(use_delta_accelerate)
            self.frequency_dim, self.fft_length = self._inner.frequency_dim, self._inner.fft_length
            self.feature_dim = 3 if use_delta_accelerate else 1
            self.audio_feature_type = audio_feature_type
            self.frame_length, self.frame_step = int(frame_length), int(frame_step)
            self.spec_augment_enabled = True
            self.num_freq_masks = int(self._sa["m_F"]) if all([self._sa["F"], self._sa["m_F"]]) else 0
            self.num_time_masks = int(self._sa["m_T"]) if all([self._sa["T"], self._sa["p"], self._sa["m_T"]]) else 0
            return
        cfg = _lib.B200MelConfig()
        cfg.struct_size = ctypes.sizeof(_lib.B200MelConfig)
        cfg.audio_feature_type = _lib.FEATURE_TYPES[audio_feature_type]
        cfg.sample_rate = int(sample_rate)
        cfg.frame_length = int(frame_length)
        cfg.frame_step = int(frame_step)
        cfg.fft_length = int(fft_length) if fft_length else 0
        cfg.num_mel_bins = int(num_mel_bins or 0)
        cfg.num_mfcc = int(num_mfcc or 0)
        cfg.lower_edge_hertz = float(lower_edge_hertz or 0.0)
        cfg.upper_edge_hertz = float(upper_edge_hertz or 0.0)
        cfg.epsilon = float(epsilon)
        cfg.use_delta_accelerate = 1 if use_delta_accelerate else 0
        cfg.spec_augment_enable = 1 if sa_on else 0
        cfg.sa_F = int(_sa_get(spec_augment, "F") or 0) if sa_on else 0
        cfg.sa_m_F = int(_sa_get(spec_augment, "m_F") or 0) if sa_on else 0
        cfg.sa_T = int(_sa_get(spec_augment, "T") or 0) if sa_on else 0
        cfg.sa_m_T = int(_sa_get(spec_augment, "m_T") or 0) if sa_on else 0
        cfg.sa_p = float(_sa_get(spec_augment, "p") or 0.0) if sa_on else 0.0
        self._cfg = cfg
        self._handle = ctypes.c_void_p()
        lib = _lib.load()
        self._forward = lib.b200mel_forward
        _lib.check(lib.b200mel_create(ctypes.byref(cfg), dev.index, ctypes.byref(self._handle)))
        f, c, n = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        _lib.check(lib.b200mel_feature_shape(self._handle, ctypes.byref(f), ctypes.byref(c), ctypes.byref(n)))
        self.frequency_dim, self.feature_dim, self.fft_length = f.value, c.value, n.value
        self.audio_feature_type = audio_feature_type
        self.frame_length, self.frame_step = int(frame_length), int(frame_step)
        self.spec_augment_enabled = sa_on
        self.num_freq_masks = cfg.sa_m_F if (cfg.sa_F and cfg.sa_m_F) else 0
        self.num_time_masks = cfg.sa_m_T if (cfg.sa_T and cfg.sa_p and cfg.sa_m_T) else 0

    # ------------------------------------------------------------------------------------------
    @classmethod
    def from_config(cls, config, training: bool = False, device=None) -> "FeatureFrontend":
        """Build the fused front-end a ``DataConfig`` describes.  SpecAugment is applied only when
        ``training`` (run/train.py:99-110 applies it to the train split only) and enabled."""
        sa = config.spec_augment if (training and config.spec_augment.enable) else None
        return cls(config.audio_feature_type, config.sample_rate, config.frame_length, config.frame_step,
                   config.fft_length, config.num_mel_bins, config.num_mfcc, config.lower_edge_hertz,
                   config.upper_edge_hertz, use_delta_accelerate=config.use_delta_accelerate,
                   spec_augment=sa, device=device)

    def __del__(self):
        try:   # (at interpreter shutdown the module globals may already be gone)
            h = getattr(self, "_handle", None)
            if h is not None and h.value:
                self._handle = None
                _lib.load().b200mel_destroy(h)
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------
    def num_frames(self, n_samples: int) -> int:
        if self._inner is not None:
            return self._inner.num_frames(n_samples)
        out = ctypes.c_int64()
        _lib.check(_lib.load().b200mel_num_frames(self._handle, int(n_samples), ctypes.byref(out)))
        return out.value

    @property
    def tc_supported(self) -> bool:
        if self._inner is not None:
            return self._inner.tc_supported
        return bool(_lib.load().b200mel_tc_supported(self._handle))

    @property
    def launch_count(self) -> int:
        if self._inner is not None:
            return self._inner.launch_count
        return int(_lib.load().b200mel_launch_count(self._handle))

    def mel_weights(self) -> np.ndarray:
        if self._inner is not None:
            return self._inner.mel_weights()
        k = self.fft_length // 2 + 1
        out = np.zeros((k, self._cfg.num_mel_bins), dtype=np.float32)
        _lib.check(_lib.load().b200mel_mel_weights(self._handle, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    # ------------------------------------------------------------------------------------------
    def __call__(self, pcm: torch.Tensor, lengths=None, offsets=None, t_max: Optional[int] = None,
                 seed: int = 0, first_utterance_id: int = 0, mask_table: Optional[torch.Tensor] = None,
                 out: Optional[torch.Tensor] = None, kernel: str = "auto", return_status: bool = False,
                 validate: bool = False, n_frames: Optional[torch.Tensor] = None
                 ) -> Tuple[torch.Tensor, torch.Tensor]:
        """pcm: CUDA int16 or float32; dense [B, L] (``lengths`` optional) or ragged 1-D with
        ``offsets`` [B] (int64, element offsets) and ``lengths`` [B].  Returns
        (features [B, t_max, F, C] float32, n_frames [B] int32), both on the device; padding
        frames are exactly +0.0.  Asynchronous on the current torch stream (usual torch rules: tensors
        allocated on another stream are the caller's to synchronise).  ``out`` / ``n_frames`` may be
        passed to reuse buffers (a steady-state caller then allocates nothing per step).

        Preconditions.  Host-side ``lengths`` / ``offsets`` are range-checked here.  Device-side ones
        are NOT looked at (that would be a stream synchronisation per call): the caller guarantees
        ``0 <= lengths[b]`` and ``offsets[b] + lengths[b] <= pcm.numel()`` (dense: ``lengths[b] <= L``),
        or passes ``validate=True`` to have them checked on the device (one synchronisation).
        SpecAugment draws the reference would raise on (``tf.random.uniform`` with maxval <= minval:
        ``F`` above the feature width, an utterance with fewer frames than a drawn ``t``) leave that
        mask empty and set bit 0 of the device status word: pass ``return_status=True`` and read it
        to get the reference's error behaviour; ``forward_host`` does this itself and raises."""
        if self._inner is not None:
            return self._call_with_time_warp(pcm, lengths, offsets, t_max, seed, first_utterance_id, mask_table, out,
                                             kernel, return_status, validate, n_frames)
        if not pcm.is_cuda or pcm.device != self.device:
            raise RuntimeError("pcm must live on %s (got %s): there is no CPU path" % (self.device, pcm.device))
        if pcm.dtype == torch.int16:
            dt = _lib.PCM_I16
        elif pcm.dtype == torch.float32:
            dt = _lib.PCM_F32
        else:
            raise ValueError("pcm must be int16 or float32, got %s" % pcm.dtype)
        pcm = pcm.contiguous()
        host_lengths = None
        if offsets is None:
            if pcm.dim() == 1:
                pcm = pcm[None]
            if pcm.dim() != 2:
                raise ValueError("dense pcm must be [B, L]")
            B, L = pcm.shape
            row_stride = L
            if lengths is None:
                host_lengths = np.full(B, L, dtype=np.int64)
        else:
            if pcm.dim() != 1:
                raise ValueError("ragged pcm must be 1-D")
            if lengths is None:
                raise ValueError("ragged input needs lengths")
            B, L, row_stride = len(offsets), int(pcm.numel()), 0
        if lengths is not None and not (isinstance(lengths, torch.Tensor) and lengths.is_cuda):
            host_lengths = np.asarray(lengths.cpu() if isinstance(lengths, torch.Tensor) else lengths, dtype=np.int64)
        if host_lengths is not None:
            if host_lengths.shape != (B,):
                raise ValueError("lengths must have shape [B]")
            if host_lengths.size and (host_lengths.min() < 0 or (offsets is None and host_lengths.max() > L)):
                raise ValueError("lengths outside [0, L]")
            d_lengths = torch.from_numpy(host_lengths.astype(np.int32)).to(self.device, non_blocking=True)
        else:
            d_lengths = lengths.to(torch.int32).contiguous()
            if validate and B:
                lo, hi = int(d_lengths.min()), int(d_lengths.max())
                if lo < 0 or (offsets is None and hi > L):
                    raise ValueError("lengths outside [0, L]")
        d_offsets = None
        if offsets is not None:
            if isinstance(offsets, torch.Tensor) and offsets.is_cuda:
                d_offsets = offsets.to(torch.int64).contiguous()
                if validate and B:
                    end = d_offsets + d_lengths.to(torch.int64)
                    if int(d_offsets.min()) < 0 or int(end.max()) > L:
                        raise ValueError("offsets/lengths outside the pcm buffer")
            else:
                ho = np.asarray(offsets.cpu() if isinstance(offsets, torch.Tensor) else offsets, dtype=np.int64)
                if host_lengths is not None and ho.size and (ho.min() < 0 or (ho + host_lengths).max() > L):
                    raise ValueError("offsets/lengths outside the pcm buffer")
                d_offsets = torch.from_numpy(ho).to(self.device, non_blocking=True)
        if t_max is None:
            if host_lengths is not None:
                t_max = self.num_frames(int(host_lengths.max())) if B else 0
            elif offsets is None:
                t_max = self.num_frames(L)
            else:
                raise ValueError("t_max is required when lengths live on the device and pcm is ragged")
        t_max = int(t_max)
        F, C = self.frequency_dim, self.feature_dim
        if out is None:
            out = torch.empty((B, t_max, F, C), dtype=torch.float32, device=self.device)
        else:
            if out.shape != (B, t_max, F, C) or out.dtype != torch.float32 or not out.is_contiguous() \
                    or out.device != self.device:
                raise ValueError("out must be a contiguous float32 [B, t_max, F, C] tensor on %s" % self.device)
        if n_frames is None:
            n_frames = torch.empty((B,), dtype=torch.int32, device=self.device)
        elif n_frames.shape != (B,) or n_frames.dtype != torch.int32 or not n_frames.is_contiguous() \
                or n_frames.device != self.device:
            raise ValueError("n_frames must be a contiguous int32 [B] tensor on %s" % self.device)
        status = torch.zeros((1,), dtype=torch.int32, device=self.device) if return_status else None
        if mask_table is not None:
            n = 2 * (self.num_freq_masks + self.num_time_masks)
            mask_table = mask_table.to(device=self.device, dtype=torch.int32).contiguous()
            if mask_table.shape != (B, n):
                raise ValueError("mask_table must be int32 [B, %d]" % n)
        if kernel not in _lib.KERNELS:
            raise ValueError("kernel must be one of %s" % sorted(_lib.KERNELS))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        # (temporaries made above - device copies of host lengths / offsets - were allocated on this same
        #  stream: the caching allocator reuses them in stream order, no record_stream needed)
        _lib.check(self._forward(
            self._handle, _ptr(pcm), dt, _ptr(d_offsets), _ptr(d_lengths), B, row_stride, _ptr(out), t_max,
            _ptr(n_frames), _ptr(mask_table), seed & 0xFFFFFFFFFFFFFFFF, first_utterance_id & 0xFFFFFFFFFFFFFFFF,
            _ptr(status), _lib.KERNELS[kernel], ctypes.c_void_p(stream)))
        if return_status:
            return out, n_frames, status
        return out, n_frames

    def _call_with_time_warp(self, pcm, lengths, offsets, t_max, seed, first_utterance_id, mask_table, out, kernel,
                             return_status, validate, n_frames):
        """The reference's stage order with W set (run/train.py:88-116, data.py:275-301): feature ->
        time warp -> masks -> deltas, one launch each.  The mask draws need the frame counts on the host
        (one synchronisation) - this path is for configs that turn W on; none of the shipped ones does."""
        feats, nf, st = self._inner(pcm, lengths=lengths, offsets=offsets, t_max=t_max, kernel=kernel,
                                    validate=validate, n_frames=n_frames, return_status=True)
        feats, st_w = time_warp(feats, nf, self._warp_W, seed, first_utterance_id)
        st |= st_w
        sa = self._sa
        if self.num_freq_masks + self.num_time_masks > 0:
            if mask_table is None:
                mask_table, _, _, st_m = specaug_draw(nf.cpu().numpy(), self.frequency_dim, sa["F"], sa["m_F"], sa["T"],
                                                      sa["p"], sa["m_T"], seed, first_utterance_id, self.device)
                st |= st_m
            feats = specaug_apply(feats, mask_table.to(device=self.device, dtype=torch.int32), self.num_freq_masks,
                                  self.num_time_masks)
        if self._delta:
            feats = delta(feats, nf)
        if out is not None:
            out.copy_(feats)
            feats = out
        if return_status:
            return feats, nf, st
        return feats, nf

    # ------------------------------------------------------------------------------------------
    def bind(self, pcm: torch.Tensor, lengths: torch.Tensor, out: torch.Tensor, n_frames: torch.Tensor,
             offsets: Optional[torch.Tensor] = None, seed: int = 0, first_utterance_id: int = 0,
             mask_table: Optional[torch.Tensor] = None, status: Optional[torch.Tensor] = None,
             kernel: str = "auto", stream: Optional[torch.cuda.Stream] = None):
        """Steady-state form of ``__call__`` for a loop that reuses its buffers (a training input pipeline
        does): every argument is a device tensor, checked ONCE here; the returned zero-argument callable
        only makes the C call (``b200mel_forward`` with the same pointers), on ``stream`` (default: the
        torch stream current at bind time).  Host cost per step: the ctypes call + the launch (~5 us)
        instead of ~17 us of argument handling - that matters when the kernel is 75 us and the timed
        region starts on an idle GPU.  The callable keeps the tensors alive; results land in ``out`` /
        ``n_frames`` (and ``status``, OR-ed, when given) exactly as with ``__call__``."""
        if self._inner is not None:
            raise NotImplementedError("bind() is the single-launch form; with time warping (W) call the front-end")
        if kernel not in _lib.KERNELS:
            raise ValueError("kernel must be one of %s" % sorted(_lib.KERNELS))
        for name, t, dt in (("pcm", pcm, None), ("lengths", lengths, torch.int32), ("out", out, torch.float32),
                            ("n_frames", n_frames, torch.int32), ("offsets", offsets, torch.int64),
                            ("mask_table", mask_table, torch.int32), ("status", status, torch.int32)):
            if t is None:
                if name in ("offsets", "mask_table", "status"):
                    continue
                raise ValueError("%s is required" % name)
            if not t.is_cuda or t.device != self.device or not t.is_contiguous():
                raise ValueError("%s must be a contiguous tensor on %s" % (name, self.device))
            if dt is not None and t.dtype != dt:
                raise ValueError("%s must be %s" % (name, dt))
        if pcm.dtype == torch.int16:
            dt = _lib.PCM_I16
        elif pcm.dtype == torch.float32:
            dt = _lib.PCM_F32
        else:
            raise ValueError("pcm must be int16 or float32, got %s" % pcm.dtype)
        if offsets is None:
            if pcm.dim() != 2:
                raise ValueError("dense pcm must be [B, L]")
            B, row_stride = int(pcm.shape[0]), int(pcm.shape[1])
        else:
            if pcm.dim() != 1:
                raise ValueError("ragged pcm must be 1-D")
            B, row_stride = int(offsets.shape[0]), 0
        F, C = self.frequency_dim, self.feature_dim
        if out.dim() != 4 or out.shape[0] != B or out.shape[2] != F or out.shape[3] != C:
            raise ValueError("out must be [B, t_max, %d, %d]" % (F, C))
        t_max = int(out.shape[1])
        if lengths.shape != (B,) or n_frames.shape != (B,):
            raise ValueError("lengths and n_frames must have shape [B]")
        if mask_table is not None and mask_table.shape != (B, 2 * (self.num_freq_masks + self.num_time_masks)):
            raise ValueError("mask_table must be int32 [B, %d]" % (2 * (self.num_freq_masks + self.num_time_masks)))
        st = stream if stream is not None else torch.cuda.current_stream(self.device)
        args = (self._handle, _ptr(pcm), dt, _ptr(offsets), _ptr(lengths), B, row_stride, _ptr(out), t_max,
                _ptr(n_frames), _ptr(mask_table), seed & 0xFFFFFFFFFFFFFFFF, first_utterance_id & 0xFFFFFFFFFFFFFFFF,
                _ptr(status), _lib.KERNELS[kernel], ctypes.c_void_p(st.cuda_stream))
        keep = (pcm, lengths, out, n_frames, offsets, mask_table, status, st)
        fwd, check = self._forward, _lib.check

        def bound_step(_args=args, _keep=keep):
            rc = fwd(*_args)
            if rc:
                check(rc)

        return bound_step

    # ------------------------------------------------------------------------------------------
    def forward_host(self, pcm, lengths=None, offsets=None, t_max: Optional[int] = None, seed: int = 0,
                     first_utterance_id: int = 0, out=None, kernel: str = "auto"):
        """End-to-end call with HOST buffers (numpy arrays or CPU torch tensors; pinned memory gives
        full PCIe speed): H2D, kernel and D2H are pipelined inside the library and the call returns
        after the result is in ``out``.  Returns (features, n_frames) as host arrays of the same kind
        as ``out`` (numpy by default)."""
        res = self.forward_host_submit(pcm, lengths, offsets, t_max, seed, first_utterance_id, out, kernel, slot=0)
        self.forward_host_wait(0)
        return res

    def forward_host_wait(self, slot: int = 0):
        """Block until the step submitted on ``slot`` has delivered its host results."""
        try:
            _lib.check(_lib.load().b200mel_forward_host_wait(self._handle, int(slot)))
        finally:   # (the results are delivered even when the step's status word makes the call raise)
            getattr(self, "_host_pending", {}).pop(int(slot), None)

    def forward_host_submit(self, pcm, lengths=None, offsets=None, t_max: Optional[int] = None, seed: int = 0,
                            first_utterance_id: int = 0, out=None, kernel: str = "auto", slot: int = 0):
        """``forward_host`` without the final wait: enqueue the step on staging ``slot`` (0 or 1) and return
        ``(features, n_frames)`` - host arrays that are complete only after ``forward_host_wait(slot)``.
        Submitting step i+1 on the other slot before waiting for step i overlaps its uploads with step i's
        downloads (both directions of the link busy), like the reference's ``prefetch`` (run/train.py:194).
        Pinned buffers are needed for the calls to be asynchronous."""
        if self._inner is not None:
            raise NotImplementedError("the host-buffer call is the single-launch form; with time warping (W) call the "
                                      "front-end on device tensors")
        pcm_np = pcm.numpy() if isinstance(pcm, torch.Tensor) else np.asarray(pcm)
        if pcm_np.dtype == np.int16:
            dt = _lib.PCM_I16
        elif pcm_np.dtype == np.float32:
            dt = _lib.PCM_F32
        else:
            raise ValueError("pcm must be int16 or float32")
        if not pcm_np.flags["C_CONTIGUOUS"]:
            pcm_np = np.ascontiguousarray(pcm_np)
        if offsets is None:
            if pcm_np.ndim == 1:
                pcm_np = pcm_np[None]
            B, L = pcm_np.shape
            row_stride = L
            lens = np.full(B, L, dtype=np.int32) if lengths is None else np.ascontiguousarray(lengths, dtype=np.int32)
            offs = None
        else:
            B, L, row_stride = len(offsets), pcm_np.size, 0
            lens = np.ascontiguousarray(lengths, dtype=np.int32)
            offs = np.ascontiguousarray(offsets, dtype=np.int64)
        if lens.shape != (B,):
            raise ValueError("lengths must have shape [B]")
        if B and (lens.min() < 0 or (offs is None and lens.max() > L)):
            raise ValueError("lengths outside [0, L]")
        if t_max is None:
            t_max = self.num_frames(int(lens.max())) if B else 0
        F, C = self.frequency_dim, self.feature_dim
        if out is None:
            out = np.empty((B, t_max, F, C), dtype=np.float32)
        out_np = out.numpy() if isinstance(out, torch.Tensor) else out
        if out_np.shape != (B, t_max, F, C) or out_np.dtype != np.float32 or not out_np.flags["C_CONTIGUOUS"]:
            raise ValueError("out must be a C-contiguous float32 [B, t_max, F, C] array")
        n_frames = np.empty((B,), dtype=np.int32)
        vp = ctypes.c_void_p
        _lib.check(_lib.load().b200mel_forward_host_submit(
            self._handle, pcm_np.ctypes.data_as(vp), dt, offs.ctypes.data_as(vp) if offs is not None else None,
            lens.ctypes.data_as(vp), B, row_stride, pcm_np.size, out_np.ctypes.data_as(vp), int(t_max),
            n_frames.ctypes.data_as(vp), seed & 0xFFFFFFFFFFFFFFFF, first_utterance_id & 0xFFFFFFFFFFFFFFFF,
            _lib.KERNELS[kernel], int(slot)))
        # the library reads / writes these buffers until the wait: keep them alive
        if not hasattr(self, "_host_pending"):
            self._host_pending = {}
        self._host_pending[int(slot)] = (pcm, pcm_np, lens, offs, out, out_np, n_frames)
        return out, n_frames


# ----------------------------------------------------------------------------------------------
# stand-alone stages (device tensors)
# ----------------------------------------------------------------------------------------------

def specaug_draw(num_time: Sequence[int], v: int, F, m_F, T, p, m_T, seed: int, first_utterance_id: int,
                 device) -> Tuple[torch.Tensor, int, int, torch.Tensor]:
    """Draw the mask tables of a batch on the device.  Returns (table int32 [B, 2(m_F+m_T)],
    n_freq_masks, n_time_masks, status int32[1])."""
    nf = int(m_F) if all([F, m_F]) else 0
    nt = int(m_T) if all([T, p, m_T]) else 0
    nt_dev = torch.as_tensor(np.asarray(num_time, dtype=np.int32)).to(device)
    B = nt_dev.numel()
    table = torch.zeros((B, 2 * (nf + nt)), dtype=torch.int32, device=device)
    status = torch.zeros((1,), dtype=torch.int32, device=device)
    stream = torch.cuda.current_stream(device).cuda_stream
    _lib.check(_lib.load().b200mel_specaug_draw(
        _ptr(nt_dev), B, int(v), int(F or 0), int(m_F or 0), int(T or 0), float(p or 0.0), int(m_T or 0),
        seed & 0xFFFFFFFFFFFFFFFF, first_utterance_id & 0xFFFFFFFFFFFFFFFF, _ptr(table), _ptr(status),
        ctypes.c_void_p(stream)))
    return table, nf, nt, status


def specaug_apply(feat: torch.Tensor, table: torch.Tensor, n_freq: int, n_time: int) -> torch.Tensor:
    """feat [B, T, v, C] float32 CUDA -> masked copy."""
    feat = feat.contiguous()
    B, T, v, C = feat.shape
    out = torch.empty_like(feat)
    stream = torch.cuda.current_stream(feat.device).cuda_stream
    _lib.check(_lib.load().b200mel_specaug_apply(_ptr(feat), _ptr(out), B, T, v, C, _ptr(table.contiguous()),
                                                 n_freq, n_time, ctypes.c_void_p(stream)))
    return out


def time_warp(feat: torch.Tensor, n_frames, W: int, seed: int = 0, first_utterance_id: int = 0,
              warp_table: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """SpecAugment time warp (data.py:275-280) of feat [B, T, v, C] float32 CUDA, per utterance over its own
    n_frames[b] rows.  Returns (warped copy, status int32[1]: bit 0 = an utterance too short for the
    reference's draw was copied unchanged).  ``warp_table`` int32 [B, 2] = (src_t, dst_t) replaces the draws."""
    feat = feat.contiguous()
    B, T, v, C = feat.shape
    nf = torch.as_tensor(n_frames).to(device=feat.device, dtype=torch.int32).contiguous()
    if nf.shape != (B,):
        raise ValueError("n_frames must have shape [B]")
    if warp_table is not None:
        warp_table = warp_table.to(device=feat.device, dtype=torch.int32).contiguous()
        if warp_table.shape != (B, 2):
            raise ValueError("warp_table must be int32 [B, 2]")
    out = torch.empty_like(feat)
    status = torch.zeros((1,), dtype=torch.int32, device=feat.device)
    stream = torch.cuda.current_stream(feat.device).cuda_stream
    _lib.check(_lib.load().b200mel_time_warp(_ptr(feat), _ptr(out), _ptr(nf), B, T, v, C, int(W or 0), _ptr(warp_table),
                                             seed & 0xFFFFFFFFFFFFFFFF, first_utterance_id & 0xFFFFFFFFFFFFFFFF,
                                             _ptr(status), ctypes.c_void_p(stream)))
    return out, status


def delta(feat: torch.Tensor, n_frames: Optional[torch.Tensor] = None) -> torch.Tensor:
    """feat [B, T, v, 1] float32 CUDA -> [B, T, v, 3]."""
    feat = feat.contiguous()
    B, T, v, C = feat.shape
    if C != 1:
        raise ValueError("delta_accelerate expects a single channel, got %d" % C)
    out = torch.empty((B, T, v, 3), dtype=torch.float32, device=feat.device)
    if n_frames is not None:
        n_frames = n_frames.to(device=feat.device, dtype=torch.int32).contiguous()
    stream = torch.cuda.current_stream(feat.device).cuda_stream
    _lib.check(_lib.load().b200mel_delta(_ptr(feat), _ptr(out), _ptr(n_frames), B, T, v, ctypes.c_void_p(stream)))
    return out
