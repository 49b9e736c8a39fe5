"""A small invocation of EVERY kernel of libb200mel.so through the public API (= the C ABI), for
compute-sanitizer (scripts/gpu_sanitize.sh).  Sizes are tiny (the tools slow kernels down 10-100x) but
every case covers more than one tile per CTA and a ragged tail.  Each case also checks the status word
and, for the tensor-core kernel, agreement with the generic kernel, so a plain run is a test as well."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import speech_recognition_b200 as pkg                                     # noqa: E402
from speech_recognition_b200 import batching, ingest                      # noqa: E402

LIBRI = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
             lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
SA = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
SA_MFCC = dict(enable=True, F=7, m_F=2, T=40, p=1.0, m_T=2)
dev = torch.device("cuda:0")
rng = np.random.default_rng(5)
# enough tiles for every CTA of the persistent kernel to loop: 200 utterances x 3.2 s = 800 tiles / 148 SMs
B, L = 200, 51200
x = np.clip(np.rint(rng.standard_normal((B, L)) * 3000.0), -32768, 32767).astype(np.int16)
lengths = rng.integers(L // 2, L + 1, B).astype(np.int32)   # (>= 159 frames: every SpecAugment draw below is valid)
lengths[0], lengths[1] = L, 320 + 160 * 120
pcm = torch.from_numpy(x).to(dev)
len_d = torch.from_numpy(lengths).to(dev)


def ok(name):
    torch.cuda.synchronize()
    print("ok:", name, flush=True)


def pair(name, ftype, tol, **kw):
    """tensor-core kernel against the generic kernel on the same batch."""
    fe = pkg.FeatureFrontend(ftype, **kw)
    a, nfa, sa_ = fe(pcm, lengths=len_d, seed=43, first_utterance_id=7, kernel="tc", return_status=True)
    b, nfb, sb_ = fe(pcm, lengths=len_d, seed=43, first_utterance_id=7, kernel="generic", return_status=True)
    assert int(sa_.item()) == 0 and int(sb_.item()) == 0, (name, int(sa_.item()), int(sb_.item()))
    assert torch.equal(nfa, nfb)
    err = (a - b).abs().max().item()
    assert err <= tol, (name, err)
    ok("%s (tc vs generic max abs err %.2e)" % (name, err))
    return fe


fe = pair("log-mel C=1", "log-mel-spectrogram", 2e-2, **LIBRI)
pair("log-mel C=1 + masks", "log-mel-spectrogram", 2e-2, **LIBRI, spec_augment=SA)
pair("log-mel C=3", "log-mel-spectrogram", 4e-2, **LIBRI, use_delta_accelerate=True)
pair("log-mel C=3 + masks", "log-mel-spectrogram", 4e-2, **LIBRI, use_delta_accelerate=True, spec_augment=SA)
pair("mfcc C=1", "mfcc", 5e-2, **LIBRI, num_mfcc=40)
fe_m = pair("mfcc C=3 + masks", "mfcc", 1e-1, **LIBRI, num_mfcc=40, use_delta_accelerate=True, spec_augment=SA_MFCC)
pair("mfcc-13 (scalar second stage)", "mfcc", 5e-2, **LIBRI, num_mfcc=13)

# float32 entry: exact int16 / 32768 (conversion + tensor-core kernel) and inexact (generic kernel takes over)
xf = pcm.to(torch.float32) / 32768.0
a, _ = fe(xf, lengths=len_d)
b, _ = fe(pcm, lengths=len_d, kernel="tc")
assert torch.equal(a, b)
ok("float32 entry, exact samples")
xf2 = xf.clone()
xf2[3, 1000] += 1e-6
a, _ = fe(xf2, lengths=len_d)
assert (a - b).abs().max().item() < 2e-2
ok("float32 entry, inexact sample -> generic kernel")

# ragged (packed) input through offsets
offs = np.zeros(B, np.int64)
offs[1:] = np.cumsum(lengths[:-1].astype(np.int64))
flat = torch.from_numpy(np.concatenate([x[i, :lengths[i]] for i in range(B)])).to(dev)
t_max = fe.num_frames(L)
a, _ = fe(flat, lengths=len_d, offsets=torch.from_numpy(offs).to(dev), t_max=t_max, kernel="tc")
assert torch.equal(a, b)
ok("ragged offsets")

# other feature types of the generic kernel
sp = pkg.FeatureFrontend("spectrogram", sample_rate=16000, frame_length=400, frame_step=160, fft_length=512)
sp(pcm[:16], lengths=len_d[:16])
ok("spectrogram 400/160/512")
lm2 = pkg.FeatureFrontend("log-mel-spectrogram", sample_rate=16000, frame_length=400, frame_step=160, fft_length=512,
                          num_mel_bins=40, lower_edge_hertz=20.0, upper_edge_hertz=8000.0, use_delta_accelerate=True,
                          spec_augment=dict(enable=True, F=10, m_F=1, T=30, p=0.5, m_T=1))
lm2(pcm[:16], lengths=len_d[:16], seed=1)
ok("generic log-mel 400/160/512 C=3 + masks")

# host-buffer call (staging, chunked copies)
out_h, nf_h = fe.forward_host(x[:32], lengths=lengths[:32])
assert np.array_equal(out_h, b[:32].cpu().numpy())
ok("forward_host")

# no-alloc entry
import ctypes                                                             # noqa: E402
from speech_recognition_b200 import _lib                                  # noqa: E402

lib = _lib.load()
out = torch.empty((24, t_max, 40, 3), dtype=torch.float32, device=dev)
nf = torch.empty(24, dtype=torch.int32, device=dev)
st = torch.zeros(1, dtype=torch.int32, device=dev)
want, _ = fe_m(pcm[:24], lengths=len_d[:24], seed=43, first_utterance_id=7, kernel="tc")
rc = lib.b200mel_forward_mfcc_noalloc(fe_m._handle, ctypes.c_void_p(pcm.data_ptr()), None, ctypes.c_void_p(len_d.data_ptr()),
                                      24, L, ctypes.c_void_p(out.data_ptr()), t_max, ctypes.c_void_p(nf.data_ptr()), None,
                                      43, 7, ctypes.c_void_p(st.data_ptr()),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
assert rc == 0, lib.b200mel_last_error()
assert torch.equal(out, want)
ok("b200mel_forward_mfcc_noalloc")

# stand-alone stages and the rows either side of the path
from speech_recognition_b200 import frontend as F_                        # noqa: E402

tbl, n_f, n_t, st2 = F_.specaug_draw(nf.cpu().numpy(), 80, 27, 2, 100, 1.0, 2, 43, 7, dev)
assert int(st2.item()) == 0
masked = F_.specaug_apply(b[:24], tbl, n_f, n_t)
ok("specaug_draw / specaug_apply")
d3 = F_.delta(b[:24], nf)
assert d3.shape[-1] == 3
ok("delta")
m = batching.audio_mask(b[:8], [11, 11], [2, 1])
ok("audio_mask")
pcm_files = [x[i, :lengths[i]].tobytes() for i in range(2, 10)]
got, glen = ingest.ingest_file_images(pcm_files, "pcm", device=dev)
assert torch.equal(got[0, :lengths[2]].cpu(), torch.from_numpy(x[2, :lengths[2]]))
ok("ingest_pcm16")
print("all cases passed", flush=True)
