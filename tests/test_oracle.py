"""Pin the oracle: against the reference's golden fixture, its shape contract, independent
implementations (torch.stft, scipy dct) and analytic known answers.  CPU only."""
import os

import numpy as np
import pytest

from oracle import logmel_oracle as O

from .helpers import GOLDEN_DIR, LIBRI, REF_LOGMEL_GRID, REF_MFCC_GRID, REF_SPECTROGRAM_GRID, noise_int16

golden = np.load(os.path.join(GOLDEN_DIR, "ref_wav_dataset_tfrecord.npz"))


def test_golden_tfrecord_exact():
    # reference tests/test_data.py:53-57 compares with exact equality
    x = golden["wav_pcm_int16"].astype(np.float32) / 32768.0
    assert x.shape == (66150,)
    for dt in (np.float32, np.float64):
        y = O.make_log_mel_spectrogram(x, 16000, 320, 160, 320, 80, 80.0, 7600.0, dtype=dt)
        for i in range(2):
            want = golden["feature_%d" % i]
            assert want.shape == (412, 80, 1)
            assert np.array_equal(y.astype(np.float32), want)
    assert golden["tokens_0"].shape == (22,) and golden["tokens_1"].shape == (9,)


@pytest.mark.parametrize("frame_length,frame_step,fft_length", REF_SPECTROGRAM_GRID)
def test_spectrogram_shape_contract(frame_length, frame_step, fft_length):
    # reference tests/test_data.py:60-78 (note: its `fft_length = frame_length` fallback for None)
    L = 66150
    y = O.make_spectrogram(np.zeros(L, np.float32), frame_length, frame_step, fft_length, np.float32)
    n = fft_length if fft_length is not None else frame_length
    assert y.shape == ((L - frame_length + frame_step) // frame_step, n // 2 + 1, 1)


@pytest.mark.parametrize("sr,fl,fs,n,m,lo,hi", REF_LOGMEL_GRID)
def test_logmel_shape_contract(sr, fl, fs, n, m, lo, hi):
    L = 66150
    y = O.make_log_mel_spectrogram(np.zeros(L, np.float32), sr, fl, fs, n, m, lo, hi, dtype=np.float32)
    assert y.shape == ((L - fl + fs) // fs, m, 1)


@pytest.mark.parametrize("sr,fl,fs,n,m,c,lo,hi", REF_MFCC_GRID)
def test_mfcc_shape_contract(sr, fl, fs, n, m, c, lo, hi):
    L = 66150
    y = O.make_mfcc(np.zeros(L, np.float32), sr, fl, fs, n, m, c, lo, hi, dtype=np.float32)
    assert y.shape == ((L - fl + fs) // fs, c, 1)


def test_frame_count_edges():
    for L, T in [(0, 0), (319, 0), (320, 1), (479, 1), (480, 2), (66150, 412), (240000, 1499)]:
        assert O.frame_count(L, 320, 160) == T


def test_stft_matches_torch():
    """torch.stft as the independent implementation for EVERY case, not only fl == n: TF left-aligns the
    window inside fft_length (zero-padding at the end) - emulated by a length-n window whose tail is
    zero - and crops to the first n windowed samples when fl > n (a length-n head of the window)."""
    torch = pytest.importorskip("torch")
    x = noise_int16(16000, 3).astype(np.float64) / 32768.0
    for fl, fs, n in [(320, 160, 320), (128, 64, 256), (400, 160, 512), (512, 512, 256), (127, 50, 128),
                      (1024, 1024, 1024), (128, 80, 128)]:
        S = O.stft(x, fl, fs, n, np.float64)
        if fl % 2 == 1:  # TF's odd-length rule (divide by Lw - 1): no torch equivalent, pinned in test_hann_window_rules
            win = torch.tensor(O.hann_window(fl, np.float64))
        else:
            win = torch.hann_window(fl, periodic=True, dtype=torch.float64)
        full = torch.zeros(n, dtype=torch.float64)
        full[:min(fl, n)] = win[:min(fl, n)]
        # (fl < n: the last TF frames reach n - fl samples past the signal - under the window's zero tail)
        xp = np.concatenate([x, np.zeros(max(n - fl, 0))])
        St = torch.stft(torch.tensor(xp), n, fs, n, window=full, center=False, return_complex=True).T.numpy()
        T = O.frame_count(len(x), fl, fs)        # TF counts frames with frame_length, torch with n_fft
        assert S.shape == (T, n // 2 + 1) and St.shape[0] >= T
        assert np.abs(S - St[:T]).max() < 1e-12, (fl, fs, n)


@pytest.mark.parametrize("m,n,sr,lo,hi", [(80, 320, 16000, 80.0, 7600.0), (80, 1024, 22050, 10.0, 10000.0),
                                          (123, 256, 16000, 12.0, 88.0), (40, 512, 16000, 0.0, 8000.0)])
def test_mel_matrix_matches_kaldi_banks(m, n, sr, lo, hi):
    """Independent filterbank: torchaudio.compliance.kaldi.get_mel_banks is linear-in-mel HTK
    (1127 ln(1 + f/700)) like tf.signal.linear_to_mel_weight_matrix; it returns [m, n/2] (no Nyquist
    row).  TF zeroes the DC row by construction; Kaldi's DC weight is zero whenever lo > 0."""
    kaldi = pytest.importorskip("torchaudio.compliance.kaldi")
    banks, _ = kaldi.get_mel_banks(m, n, float(sr), lo, hi, 100.0, -500.0, 1.0)
    K = banks.numpy().T.astype(np.float64)                     # [n/2, m]
    W = O.linear_to_mel_weight_matrix(m, n // 2 + 1, sr, lo, hi, dtype=np.float64)
    first = 1 if lo == 0.0 else 0                              # (lo = 0: Kaldi keeps a DC weight, TF forces 0)
    # torchaudio computes in float32: 7.8e-6 on the shipped set, up to 2.5e-5 where the filters are a
    # fraction of a mel wide (123 filters between 12 and 88 Hz) and the rounding of the edges is amplified
    assert np.abs(W[first:n // 2] - K[first:]).max() < (1e-5 if m == 80 and n == 320 else 5e-5)
    assert np.array_equal(W[first:n // 2] > 0, K[first:] > 0) or np.abs(W[first:n // 2] - K[first:])[(W[first:n // 2] > 0) != (K[first:] > 0)].max() < 5e-5
    if hi <= sr / 2 * (1 - 2.0 / n):
        assert np.all(W[n // 2] == 0)                          # Nyquist row
    W32 = O.linear_to_mel_weight_matrix(m, n // 2 + 1, sr, lo, hi)
    assert np.abs(W32[first:n // 2] - K[first:]).max() < 6e-5


def test_impulse_known_answer():
    # unit impulse at n0 inside frame 0: |X[k]| = w[n0] for every k
    x = np.zeros(320)
    x[77] = 1.0
    S = O.stft(x, 320, 160, 320, np.float64)
    w = O.hann_window(320, np.float64)
    assert np.allclose(np.abs(S[0]), w[77], atol=1e-14)


def test_hann_window_rules():
    w = O.hann_window(320, np.float64)
    assert w[0] == 0.0 and abs(w[160] - 1.0) < 1e-15 and abs(w[1] - w[319]) < 1e-15
    w_odd = O.hann_window(5, np.float64)   # odd: divides by Lw - 1 (symmetric)
    assert np.allclose(w_odd, 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(5) / 4))
    assert np.array_equal(O.hann_window(1), np.ones(1))


def test_mel_matrix_properties():
    W = O.linear_to_mel_weight_matrix(80, 161, 16000, 80.0, 7600.0)
    assert W.dtype == np.float32 and W.shape == (161, 80)
    assert (W > 0).sum() == 295
    nz_rows = np.nonzero(W.sum(1))[0]
    assert nz_rows[0] == 2 and nz_rows[-1] == 151
    assert (W > 0).sum(1).max() <= 2 and (W > 0).sum(0).max() <= 9
    assert np.all(W[0] == 0)                         # DC row zeroed
    W64 = O.linear_to_mel_weight_matrix(80, 161, 16000, 80.0, 7600.0, dtype=np.float64)
    assert np.abs(W - W64).max() < 2e-5
    # each column is a triangle: contiguous support, single peak
    for m in range(80):
        nz = np.nonzero(W[:, m])[0]
        assert np.array_equal(nz, np.arange(nz[0], nz[-1] + 1))


def test_mel_argument_validation():
    with pytest.raises(ValueError):
        O.linear_to_mel_weight_matrix(80, 161, 16000, 80.0, 9000.0)
    with pytest.raises(ValueError):
        O.linear_to_mel_weight_matrix(80, 161, 16000, 7600.0, 80.0)
    with pytest.raises(ValueError):
        O.linear_to_mel_weight_matrix(0, 161, 16000, 80.0, 7600.0)
    with pytest.raises(ValueError):
        O.linear_to_mel_weight_matrix(80, 161, 16000, -1.0, 7600.0)


def test_dct_matches_scipy():
    scipy_fft = pytest.importorskip("scipy.fft")
    x = np.random.default_rng(0).normal(size=(7, 80))
    want = scipy_fft.dct(x, type=2, norm=None, axis=-1) / np.sqrt(2 * 80)
    got = x @ O.dct2_matrix(80, 40)
    assert np.abs(got - want[:, :40]).max() < 1e-12


def test_f32_vs_f64_budget():
    x = noise_int16(240000, 0).astype(np.float64) / 32768.0
    y64 = O.make_log_mel_spectrogram(x, **LIBRI, dtype=np.float64)
    y32 = O.make_log_mel_spectrogram(x.astype(np.float32), **LIBRI, dtype=np.float32)
    assert y64.shape == (1499, 80, 1)
    assert np.abs(y32 - y64).max() < 5e-5


def test_delta_on_ramp():
    T, F = 6, 3
    x = (np.arange(T, dtype=np.float32)[:, None, None] * 2.0 + 1.0) * np.ones((1, F, 1), np.float32)
    y = O.delta_accelerate(x)
    assert y.shape == (T, F, 3)
    assert np.array_equal(y[:, 0, 1], [1, 2, 2, 2, 2, 2])     # delta[0] = x[0]
    assert np.array_equal(y[:, 0, 2], [1, 1, 0, 0, 0, 0])     # deltas[0] = delta[0]


@pytest.mark.parametrize("W,F,m_F,T,p,m_T", [(None, 27, 1, 100, 1.0, 1), (None, 15, 2, 70, 0.2, 2)])
def test_spec_augment_invariants(W, F, m_F, T, p, m_T):
    # reference tests/test_data.py:148-163 (without the time-warp argument)
    num_time, num_freq = 234, 80
    data = np.random.default_rng(1).uniform(0.1, 1.0, (num_time, num_freq, 1)).astype(np.float32)
    for utt in range(20):
        aug = O.spec_augment(data, num_freq, W, F, m_F, T, p, m_T, seed=43, utterance_id=utt)
        is_zero = np.all(aug == 0.0, axis=2)
        assert np.count_nonzero(is_zero.all(axis=0)) <= F * m_F
        assert np.count_nonzero(is_zero.all(axis=1)) <= T * m_T
        assert np.count_nonzero(is_zero.all(axis=1)) <= int(np.float32(num_time) * np.float32(p)) or p == 1.0
        assert aug.shape == data.shape


def test_spec_augment_fixed_table_and_negative_zero():
    x = -np.ones((10, 8, 1), np.float32)
    y = O.spec_augment_apply(x, [(2, 3)], [(4, 2), (0, 0)])
    assert np.all(y[:, 2:5] == 0) and np.all(y[4:6] == 0)
    assert np.all(np.signbit(y[:, 2:5]))          # masked negatives are -0.0 (audio *= mask)
    keep = np.ones((10, 8), bool)
    keep[:, 2:5] = False
    keep[4:6] = False
    assert np.all(y[keep] == -1.0)


def test_draw_mapping_and_errors():
    raw = [5, 1000, 99, 7, 250, 3]
    tbl = O.specaug_draw_masks(50, 80, F=27, m_F=1, T=100, p=0.5, m_T=2, raw=raw)
    assert tbl[0] == (1000 % (80 - 5), 5)
    assert tbl[1] == (7 % (50 - 25), 25)           # t = min(99, int(50*0.5)=25)
    assert tbl[2] == (3 % 50, 0)                   # budget exhausted -> t = 0
    with pytest.raises(ValueError):                # tf.random.uniform: maxval <= minval
        O.specaug_draw_masks(10, 80, T=100, p=1.0, m_T=1, raw=[10, 0])


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    assert O.philox4x32_10((0, 0, 0, 0), (0, 0)) == (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)
    assert O.philox4x32_10((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2) == (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)
    assert O.philox4x32_10((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0)) == \
        (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)


# ---- the plain-C restatement (second oracle, also the timed CPU stand-in) -----------------------
def test_c_port_matches_numpy_oracle():
    from oracle.ref_c import RefPlan

    from .helpers import coloured_int16

    pcms = [noise_int16(24000, 1), coloured_int16(16000, 2), noise_int16(100, 3), noise_int16(320, 4)]
    L = max(len(p) for p in pcms)
    batch = np.zeros((len(pcms), L), np.int16)
    for i, p in enumerate(pcms):
        batch[i, :len(p)] = p
    lengths = [len(p) for p in pcms]
    for ftype, kw in [("log-mel-spectrogram", {}), ("mfcc", {}), ("spectrogram", {})]:
        for delta in (False, True):
            plan = RefPlan(ftype, 16000, 320, 160, 320, 80, 40, 80.0, 7600.0, use_delta_accelerate=delta)
            got, nf = plan(batch, lengths)
            want, nf_w = O.batched_features([p.astype(np.float64) / 32768.0 for p in pcms], ftype, 16000, 320, 160,
                                            320, 80, 40, 80.0, 7600.0, use_delta=delta)
            assert np.array_equal(nf, nf_w) and got.shape == want.shape
            tol = 2e-4 if ftype != "mfcc" else 2e-3
            if ftype == "spectrogram":
                tol = 2e-6 * np.abs(want).max()
            assert np.abs(got - want).max() <= tol * (4 if delta else 1), (ftype, delta, np.abs(got - want).max())


@pytest.mark.parametrize("sr,fl,fs,n,m,lo,hi", REF_LOGMEL_GRID + [(16000, 400, 160, 512, 80, 80.0, 7600.0),
                                                                    (16000, 15, 7, 15, 10, 100.0, 7000.0),
                                                                    (16000, 14, 5, 22, 10, 100.0, 7000.0)])
def test_c_port_parameter_grid(sr, fl, fs, n, m, lo, hi):
    from oracle.ref_c import RefPlan

    from .helpers import assert_logmel_close

    pcm = noise_int16(20000, 7)
    got, _ = RefPlan("log-mel-spectrogram", sr, fl, fs, n, m, 0, lo, hi)(pcm[None])
    want = O.make_log_mel_spectrogram(pcm.astype(np.float64) / 32768.0, sr, fl, fs, n, m, lo, hi, dtype=np.float64)
    assert_logmel_close(got[0], want, what=str((sr, fl, fs, n, m)))


def test_c_port_golden_and_ragged():
    from oracle.ref_c import RefPlan

    plan = RefPlan("log-mel-spectrogram", 16000, 320, 160, 320, 80, 40, 80.0, 7600.0)
    got, nf = plan(np.stack([golden["wav_pcm_int16"]] * 2))
    assert np.array_equal(got[0], golden["feature_0"]) and np.array_equal(got[1], golden["feature_1"])
    lengths = [5000, 12000, 700]
    pcms = [noise_int16(n, i) for i, n in enumerate(lengths)]
    offs = np.concatenate([[0], np.cumsum(lengths)[:-1]])
    ragged, _ = plan(np.concatenate(pcms), lengths, offs)
    dense = np.zeros((3, 12000), np.int16)
    for i, p in enumerate(pcms):
        dense[i, :len(p)] = p
    d, _ = plan(dense, lengths)
    assert np.array_equal(ragged, d)
