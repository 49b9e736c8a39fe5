"""Batch contract: max-over policy on the batch and the models' conv-reduced
padding mask (models/las.py:205-217, models/deepspeech2.py:68-78) against the oracle's line-by-line
restatement."""
import numpy as np
import pytest
import torch

from oracle import pipeline_oracle as P

from .helpers import LIBRI, noise_int16


@pytest.fixture(scope="module")
def bat(built_lib):
    from speech_recognition_b200 import batching

    return batching


def _py_len(t, ks, ss):
    g = 1
    for k, s in zip(ks, ss):
        t -= k - s
        t = t // s
        g *= s
    return max(t, 0), g


@pytest.mark.parametrize("layers", [([3, 3], [2, 2]), ([11, 11], [2, 1]), ([41, 21, 21], [2, 2, 1]), ([5], [3]), ([], [])])
def test_mask_length_follows_reference_arithmetic(bat, layers):
    for t in list(range(0, 70)) + [412, 1499, 2048]:
        assert bat.audio_mask_length(t, *layers) == _py_len(t, *layers)
    with pytest.raises(ValueError):
        bat.audio_mask_length(10, [3], [0])
    with pytest.raises(ValueError):
        bat.audio_mask_length(10, [3, 3], [2])


def test_policy_wrappers_cpu(bat):
    keep = bat.filter_example(100, 5)(torch.tensor([99, 100, 101, 50]), torch.tensor([5, 6, 1, 0]))
    assert keep.tolist() == [True, False, False, True]   # data.py:337: both <= max
    f, t = bat.slice_example(3, 2)(torch.zeros(2, 7, 4, 1), torch.zeros(2, 9))
    assert f.shape == (2, 3, 4, 1) and t.shape == (2, 2)   # data.py:350
    with pytest.raises(RuntimeError):
        bat.audio_mask(torch.zeros(1, 8, 4, 1), [3, 3], [2, 2])   # no CPU path


gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("shape", [(3, 64, 80, 1), (2, 1499, 80, 3), (4, 37, 40, 3), (2, 9, 161, 1), (1, 3, 5, 1)])
def test_audio_mask_matches_las(bat, shape):
    rng = np.random.default_rng(shape[1])
    x = np.zeros(shape, np.float32)
    B, T = shape[:2]
    for b in range(B):
        n = int(rng.integers(0, T + 1))
        x[b, :n] = rng.normal(0, 1, (n,) + shape[2:])
        if n > 6:
            x[b, 2:5] = 0.0               # a SpecAugment time mask: frames of zeros inside the utterance
            x[b, 5] = -0.0                # masked negatives are -0.0 (data.py:288,301): still padding
        if n > 12:
            x[b, 9] = 0.0
            x[b, 9, -1, -1] = 1e-30       # one tiny value keeps the frame
            x[b, 11, 0, 0] = np.nan       # NaN != 0 is True
    got = bat.audio_mask(torch.from_numpy(x).cuda(), [3, 3], [2, 2])
    torch.cuda.synchronize()
    want = P.las_audio_mask(x, 3, 2)
    assert got.dtype == torch.bool and got.shape == want.shape
    assert np.array_equal(got.cpu().numpy(), want)


@gpu
@pytest.mark.parametrize("layers", [([11, 11], [2, 1]), ([41, 21, 21], [2, 2, 1]), ([5], [3]), ([], [])])
def test_audio_mask_matches_conv_stack(bat, layers):
    rng = np.random.default_rng(len(layers[0]))
    x = rng.normal(0, 1, (3, 211, 7, 3)).astype(np.float32)
    x[0, 100:] = 0
    x[1, 13:] = 0
    x[2, ::2] = 0
    got = bat.audio_mask(torch.from_numpy(x).cuda(), *layers)
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), P.conv_audio_mask(x, *layers))
    # a non-contiguous / misaligned view goes through the scalar path
    y = torch.from_numpy(x).cuda()[:, 1:, :, :]
    got = bat.audio_mask(y, *layers)
    assert np.array_equal(got.cpu().numpy(), P.conv_audio_mask(x[:, 1:], *layers))


@gpu
def test_assemble_batch_policies(bat):
    import speech_recognition_b200 as pkg

    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True,
                             spec_augment=dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2))
    lens = [48000, 16000, 31999, 8000]
    pcm = np.zeros((4, 48000), np.int16)
    for i, n in enumerate(lens):
        pcm[i, :n] = noise_int16(n, seed=i)
    d = torch.from_numpy(pcm).cuda()
    full = bat.assemble_batch(fe, d, lens, seed=43, mask_layers=([3, 3], [2, 2]))
    assert full["features"].shape == (4, 299, 80, 3) and full["keep"].all()
    x = full["features"].cpu().numpy()
    assert np.array_equal(full["mask"].cpu().numpy(), P.las_audio_mask(x, 3, 2))
    # the mask sees the SpecAugment time masks (zero frames inside the utterance), not only the padding
    frame_real = (x.reshape(4, 299, -1) != 0).any(-1)
    assert (~frame_real[0]).any()
    # slice: frames beyond max_audio_length are never computed; n_frames is clipped
    sl = bat.assemble_batch(fe, d, lens, seed=43, max_over_policy="slice", max_audio_length=150, max_token_length=9)
    assert sl["features"].shape == (4, 150, 80, 3) and sl["n_frames"].tolist() == [150, 99, 150, 49]
    # slice below a larger audio_pad_length: [B, audio_pad_length, ...] with zeros behind the slice
    # (the reference slices per example, data.py:345-354, then padded_batch pads, run/train.py:182-197)
    sp = bat.assemble_batch(fe, d, lens, seed=43, max_over_policy="slice", max_audio_length=150, max_token_length=9,
                            audio_pad_length=320, mask_layers=([3, 3], [2, 2]))
    assert sp["features"].shape == (4, 320, 80, 3) and sp["n_frames"].tolist() == [150, 99, 150, 49]
    assert torch.equal(sp["features"][:, :150], sl["features"]) and not sp["features"][:, 150:].any()
    assert np.array_equal(sp["mask"].cpu().numpy(), P.las_audio_mask(sp["features"].cpu().numpy(), 3, 2))
    # filter: keep-mask over n_frames and token lengths (data.py:337)
    fl = bat.assemble_batch(fe, d, lens, seed=43, max_over_policy="filter", max_audio_length=198, max_token_length=9,
                            token_lengths=torch.tensor([1, 1, 10, 9]), audio_pad_length=320)
    assert fl["features"].shape == (4, 320, 80, 3) and fl["keep"].tolist() == [False, True, False, True]
    assert not fl["features"][:, 299:].any()
