"""The tensor-core kernel's generated epilogue, checked on the CPU.

tests/host/tc_epilogue_host.cc compiles the SAME generated code the kernel runs
(csrc/tc_tables_generated.inc: accumulator column order, per-group mel FFMAs, chain hand-over A -> B -> C,
float4 staging) and the same packed radix-5 arithmetic (csrc/tc_epilogue_math.cuh, host emulation of the
f32x2 instructions) with g++, feeds it accumulator columns computed in double precision, and this test
compares the 80 staged values per frame with the numpy oracle (log-mel of data.py:145-189).  What is NOT
covered here is the GPU plumbing (MMA operand layout, TMEM, barriers): the -m gpu parity tests do that."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import logmel_oracle as O

from .helpers import LIBRI, assert_logmel_close, coloured_int16, noise_int16, tone_int16

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "tc_epilogue_host.cc")


@pytest.fixture(scope="module")
def harness(tmp_path_factory, built_lib):   # built_lib: the generated tables are refreshed by the build
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("g++ not available")
    exe = str(tmp_path_factory.mktemp("tc_host") / "tc_epilogue_host")
    subprocess.run([gxx, "-O1", "-std=c++17", "-o", exe, SRC], check=True)
    return exe


def run(harness, frames, tmp_path, mask=None):
    fin, fout = str(tmp_path / "frames.i16"), str(tmp_path / "out.f32")
    np.ascontiguousarray(frames, np.int16).tofile(fin)
    cmd = [harness, fin, fout] + ([str(mask[0]), str(mask[1])] if mask else [])
    subprocess.run(cmd, check=True)
    return np.fromfile(fout, np.float32).reshape(-1, 80)


def frames_of(pcm):
    n = (len(pcm) - 320) // 160 + 1
    return np.stack([pcm[160 * t:160 * t + 320] for t in range(n)])


def test_generated_epilogue_matches_the_oracle(harness, tmp_path):
    signals = [noise_int16(16000, 0, 0.1), noise_int16(8000, 1, 1e-3), coloured_int16(16000, 2),
               tone_int16(8000, 1000.0), tone_int16(8000, 1234.5),
               np.where((np.arange(8000) // 37) % 2 == 0, 32767, -32768).astype(np.int16), np.zeros(1000, np.int16)]
    for i, pcm in enumerate(signals):
        got = run(harness, frames_of(pcm), tmp_path)
        assert not np.isnan(got).any(), "a staged value was never written"
        want = O.make_log_mel_spectrogram(pcm.astype(np.float64) / 32768.0, **LIBRI, dtype=np.float64)[..., 0]
        assert got.shape == want.shape
        assert_logmel_close(got[..., None], want[..., None], what="signal %d" % i)
    silence = run(harness, np.zeros((3, 320), np.int16), tmp_path)
    assert np.all(silence == np.float32(np.log(np.float32(1e-12))))     # the reference's golden value (logf on the host)


@pytest.mark.parametrize("start,width", [(0, 5), (30, 27), (51, 26), (79, 1), (10, 0)])
def test_generated_epilogue_masks(harness, tmp_path, start, width):
    """SpecAugment frequency masks in the finishing macros: every bin carries its literal index, in all
    three groups and on forwarded values (zero set and sign of the zeros as audio *= mask gives them)."""
    pcm = noise_int16(4000, 5, 0.1)
    base = run(harness, frames_of(pcm), tmp_path)
    got = run(harness, frames_of(pcm), tmp_path, mask=(start, width))
    keep = np.ones(80, bool)
    keep[start:start + width] = False
    assert np.array_equal(got[:, keep], base[:, keep])
    assert not got[:, ~keep].any()
    assert np.array_equal(np.signbit(got[:, ~keep]), np.signbit(base[:, ~keep]))
