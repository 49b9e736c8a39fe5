"""Device-time of the rows either side of the path: ingest of 256 x 15 s WAV payloads, the
audio mask over the padded batch, and the chain ingest -> front-end -> mask.  CUDA events, ring of
buffers larger than L2."""
import json
import sys
import os

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import speech_recognition_b200 as pkg  # noqa: E402
from speech_recognition_b200 import _lib, batching  # noqa: E402
import ctypes  # noqa: E402


def timeit(fn, n=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3  # us


def main():
    B, L = 256, 240000
    dev = torch.device("cuda:0")
    ring = 5
    rng = np.random.default_rng(0)
    res = {}
    for hdr in (44, 48, 45):
        per = (hdr + 2 * L + 15) // 16 * 16
        bufs = [torch.from_numpy(rng.integers(0, 256, B * per, dtype=np.uint8)).to(dev) for _ in range(ring)]
        off = torch.arange(B, dtype=torch.int64, device=dev) * per + hdr
        ln = torch.full((B,), 2 * L, dtype=torch.int64, device=dev)
        outs = [torch.empty((B, L), dtype=torch.int16, device=dev) for _ in range(ring)]
        lens = torch.empty(B, dtype=torch.int32, device=dev)
        lib = _lib.load()
        i = [0]

        def run(dtype=_lib.PCM_I16):
            k = i[0] % ring
            i[0] += 1
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(lib.b200mel_ingest_pcm16(ctypes.c_void_p(bufs[k].data_ptr()), ctypes.c_void_p(off.data_ptr()),
                                                ctypes.c_void_p(ln.data_ptr()), None, B,
                                                ctypes.c_void_p(outs[k].data_ptr()), dtype, None, L,
                                                ctypes.c_void_p(lens.data_ptr()), ctypes.c_void_p(st)))
        us = timeit(run)
        res["ingest_i16_hdr%d" % hdr] = {"us": us, "GB/s": 2 * B * L * 2 / us / 1e3}
        # check
        got = outs[(i[0] - 1) % ring][3, :100].cpu().numpy()
        src = bufs[(i[0] - 1) % ring][3 * per + hdr: 3 * per + hdr + 200].cpu().numpy().view("<i2")
        assert np.array_equal(got, src), hdr
        del bufs, outs
    fe = pkg.FeatureFrontend("log-mel-spectrogram", 16000, 320, 160, 320, 80, None, 80.0, 7600.0)
    pcms = [torch.from_numpy(rng.integers(-3000, 3000, (B, L), dtype=np.int16)).to(dev) for _ in range(ring)]
    feats = [torch.empty((B, 1499, 80, 1), dtype=torch.float32, device=dev) for _ in range(ring)]
    j = [0]

    def fwd():
        k = j[0] % ring
        j[0] += 1
        fe(pcms[k], out=feats[k])
    res["frontend"] = {"us": timeit(fwd)}

    reduced, group = batching.audio_mask_length(1499, [3, 3], [2, 2])
    mout = torch.empty((B, reduced), dtype=torch.uint8, device=dev)

    def mask():
        k = j[0] % ring
        j[0] += 1
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.b200mel_audio_mask(ctypes.c_void_p(feats[k].data_ptr()), B, 1499, 80, group, reduced,
                                          ctypes.c_void_p(mout.data_ptr()), ctypes.c_void_p(st)))
    us = timeit(mask)
    res["audio_mask_C1"] = {"us": us, "GB/s": B * reduced * group * 80 * 4 / us / 1e3}
    feats3 = [torch.randn((B, 1499, 80, 3), dtype=torch.float32, device=dev) for _ in range(3)]

    def mask3():
        k = j[0] % 3
        j[0] += 1
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.b200mel_audio_mask(ctypes.c_void_p(feats3[k].data_ptr()), B, 1499, 240, group, reduced,
                                          ctypes.c_void_p(mout.data_ptr()), ctypes.c_void_p(st)))
    us = timeit(mask3)
    res["audio_mask_C3"] = {"us": us, "GB/s": B * reduced * group * 240 * 4 / us / 1e3}
    del feats3

    def both():
        k = j[0] % ring
        j[0] += 1
        fe(pcms[k], out=feats[k])
        batching.audio_mask(feats[k], [3, 3], [2, 2])
    res["frontend+mask"] = {"us": timeit(both)}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
