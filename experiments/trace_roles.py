"""Developer tool: run the TC kernel once on BASELINE config 2 with a -DTC_TRACE variant and print the
role timeline of CTA 0 (cycles relative to the first stamp).  B200MEL_LIB must point at the variant."""
import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_recognition_b200 import FeatureFrontend, _lib
cfg = dict(sample_rate=16000, frame_length=320, frame_step=160, fft_length=320, num_mel_bins=80,
           lower_edge_hertz=80.0, upper_edge_hertz=7600.0)
SA = dict(enable=True, W=None, F=27, m_F=2, T=100, p=1.0, m_T=2) if os.environ.get("TRACE_SA") else None
fe = FeatureFrontend("log-mel-spectrogram", **cfg, spec_augment=SA, device="cuda:0")
x = torch.randint(-3000, 3000, (256, 240000), dtype=torch.int16, device="cuda")
for _ in range(3):
    out, nf = fe(x, kernel="tc")
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_longlong * (5 * 32 * 8))()
lib.b200mel_debug_trace.argtypes = [ctypes.c_void_p]
assert lib.b200mel_debug_trace(buf) == 0
t = np.array(buf[:]).reshape(5, 32, 8)
t0 = t[t > 0].min()
names = ["P ", "M ", "EA", "EB", "EC"]
ev = {0: ["start", "w0", "w1", "w2", "w3", "arr_w4", "done", "arr_w7"], 1: ["top", "afull0", "pre_accE", "accE_ok", "committed", "afull1_ok"],
      2: ["top", "accfull", "released", "loopend", "", "done", "", ""], 3: ["top", "accfull", "released", "loopend", "phase1", "done", "bar_ab", ""],
      4: ["top", "accfull", "released", "loopend", "phase1", "done", "bar_bc", "stored"]}
print("kernel: top=%d setup_done=%d dep_wait_done=%d teardown=%d" % tuple(t[1, 31, e] - t0 for e in range(4)))
print("list (epilogue thread 0): enter=%d ranges_known=%d written=%d list_full_seen=%d" % tuple(t[2, 30, e] - t0 for e in range(4)))
t[1, 31, :] = 0
t[2, 30, :] = 0
for n in list(range(0, 2)) + list(range(9, 12)) + list(range(20, 21)):
    print("tile", n)
    for r in range(5):
        print("  ", names[r], " ".join("%s=%d" % (ev[r][e], t[r, n, e] - t0) for e in range(len(ev[r])) if ev[r][e] and t[r, n, e] > 0))
