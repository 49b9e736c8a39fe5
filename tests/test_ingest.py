"""PCM / WAV ingest (load_audio_file, data.py:82-119).

CPU: the RIFF walk of libb200mel.so against the reference fixture's header, Python's ``wave`` module and
hand-built images.  GPU: the byte -> sample kernels against the oracle's restatement, bit for bit, and the
ingest -> front-end chain against the front-end fed the decoded samples."""
import io
import os
import struct
import wave

import numpy as np
import pytest
import torch

from oracle import pipeline_oracle as P

from .helpers import GOLDEN_DIR, LIBRI, noise_int16


@pytest.fixture(scope="module")
def ing(built_lib):
    from speech_recognition_b200 import ingest

    return ingest


def wav_image(samples: np.ndarray, rate=16000, extra_chunks=()):
    """samples int16 [n] or [n, C] -> canonical WAV image via the standard library, optionally with
    extra chunks (id, body) spliced in front of the data chunk."""
    samples = np.asarray(samples, dtype="<i2")
    ch = 1 if samples.ndim == 1 else samples.shape[1]
    buf = io.BytesIO()
    with wave.open(buf, "wb") as w:
        w.setnchannels(ch)
        w.setsampwidth(2)
        w.setframerate(rate)
        w.writeframes(samples.tobytes())
    img = buf.getvalue()
    if extra_chunks:
        pos = img.index(b"data")
        extra = b"".join(cid + struct.pack("<I", len(body)) + body + (b"\x00" if len(body) & 1 else b"")
                         for cid, body in extra_chunks)
        img = img[:pos] + extra + img[pos:]
        img = img[:4] + struct.pack("<I", len(img) - 8) + img[8:]
    return img


# ---------------------------------------------------------------------------------------------- CPU
def test_reference_wav_header(ing):
    """tests/data/audio_files/test.wav of the reference: mono, 22 050 Hz, 16 bit, 66 150 samples."""
    with open(os.path.join(GOLDEN_DIR, "ref_test_wav_header.bin"), "rb") as f:
        head = f.read()
    img = head + bytes(132300)
    h = ing.parse_wav_header(img)
    assert (h.format, h.channels, h.sample_rate, h.bits_per_sample) == (1, 1, 22050, 16)
    assert (h.data_offset, h.data_bytes) == (44, 132300)
    ch, rate, bits, payload = P.wav_payload(img)
    assert (ch, rate, bits, len(payload)) == (1, 22050, 16, 132300)


@pytest.mark.parametrize("channels", [1, 2, 6])
@pytest.mark.parametrize("extra", [(), ((b"LIST", b"INFOabc"),), ((b"junk", b""), (b"LIST", b"xy"))])
def test_wav_parse_against_wave_module(ing, channels, extra):
    x = np.arange(30 * channels, dtype=np.int16).reshape(30, channels)
    img = wav_image(x if channels > 1 else x[:, 0], rate=8000, extra_chunks=extra)
    h = ing.parse_wav_header(img)
    with wave.open(io.BytesIO(img), "rb") as w:
        assert (h.channels, h.sample_rate, h.bits_per_sample) == (w.getnchannels(), w.getframerate(), 16)
        assert h.data_bytes == w.getnframes() * channels * 2
    assert img[h.data_offset:h.data_offset + h.data_bytes] == x.astype("<i2").tobytes()
    assert P.wav_payload(img)[3] == x.astype("<i2").tobytes()


def test_wav_parse_errors_and_clipping(ing):
    good = wav_image(np.zeros(10, np.int16))
    for bad in (b"", b"RIFF", b"RIFX" + good[4:], good[:8] + b"AVI " + good[12:], good[:36]):
        with pytest.raises(ValueError):
            ing.parse_wav_header(bad)
    nofmt = good[:12] + good[36:]
    with pytest.raises(ValueError, match="before fmt"):
        ing.parse_wav_header(nofmt)
    # streamed writers leave the data size at 0xFFFFFFFF, or the file is cut short: clip to what is there
    streamed = good[:40] + struct.pack("<I", 0xFFFFFFFF) + good[44:]
    assert ing.parse_wav_header(streamed).data_bytes == 20
    assert ing.parse_wav_header(good[:-6]).data_bytes == 14


def test_invalid_format_raises_like_reference(ing):
    with pytest.raises(ValueError, match="is not valid"):
        ing.load_audio_file(16000, "ogg")
    with pytest.raises(ValueError, match="is not valid"):
        ing.ingest_file_images([b""], "ogg")
    with pytest.raises(NotImplementedError):
        ing.ingest_file_images([b""], "flac")


# ---------------------------------------------------------------------------------------------- GPU
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("dtype", [torch.int16, torch.float32])
def test_ingest_raw_pcm_matches_oracle(ing, dtype):
    """Raw .pcm (data.py:100-105): every length class incl. empty and odd byte counts."""
    rng = np.random.default_rng(3)
    sizes = [0, 1, 2, 3, 15, 16, 17, 31, 640, 641, 4099, 66150 * 2, 66150 * 2 + 1]
    images = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in sizes]
    pcm, lengths = ing.ingest_file_images(images, "pcm", dtype=dtype)
    torch.cuda.synchronize()
    pcm, lengths = pcm.cpu().numpy(), lengths.cpu().numpy()
    assert pcm.shape[1] % 8 == 0
    for i, img in enumerate(images):
        want = P.load_audio_pcm16(img, "pcm")
        assert lengths[i] == want.size == (sizes[i] + 1) // 2
        got = pcm[i, :lengths[i]]
        if dtype == torch.int16:
            assert np.array_equal(got.astype(np.float32) / np.float32(32768.0), want)
        else:
            assert got.tobytes() == want.tobytes()
        assert not pcm[i, lengths[i]:].any(), "dense rows are zero behind the samples"


@gpu
def test_ingest_wav_mono_all_misalignments(ing):
    """The payload may start at any byte offset (extra header chunks of odd / even sizes): the funnel-shift
    path is exact for all four alignments."""
    x = noise_int16(5003, seed=5)
    images = [wav_image(x, extra_chunks=((b"LIST", b"z" * k),)) for k in (1, 2, 3, 4, 5, 6, 7, 8)]
    images.append(wav_image(x))
    pcm, lengths = ing.ingest_file_images(images, "wav")
    assert pcm.dtype == torch.int16
    torch.cuda.synchronize()
    for i in range(len(images)):
        assert int(lengths[i]) == x.size and np.array_equal(pcm[i, :x.size].cpu().numpy(), x)
    # ragged layout, float32 output
    pcm_r, off, len_r = ing.ingest_file_images(images[:3], "wav", dtype=torch.float32, ragged=True)
    torch.cuda.synchronize()
    for i in range(3):
        o = int(off[i])
        assert o % 8 == 0
        assert pcm_r[o:o + int(len_r[i])].cpu().numpy().tobytes() == P.load_audio_pcm16(images[i], "wav").tobytes()


@gpu
@pytest.mark.parametrize("channels", [2, 3, 6])
def test_ingest_wav_channel_mean(ing, channels):
    """reduce_mean over channels (data.py:116), float32 bit for bit; mixed channel counts in one batch."""
    rng = np.random.default_rng(channels)
    a = rng.integers(-32768, 32768, (777, channels)).astype(np.int16)
    b = rng.integers(-32768, 32768, 1001).astype(np.int16)
    images = [wav_image(a), wav_image(b), wav_image(a[:5], extra_chunks=((b"LIST", b"q"),))]
    pcm, lengths = ing.ingest_file_images(images, "wav")
    assert pcm.dtype == torch.float32
    torch.cuda.synchronize()
    for i, img in enumerate(images):
        want = P.load_audio_pcm16(img, "wav")
        assert int(lengths[i]) == want.size
        assert pcm[i, :want.size].cpu().numpy().tobytes() == want.tobytes()
        assert not pcm[i, want.size:].cpu().numpy().any()
    with pytest.raises(ValueError, match="float32"):
        ing.ingest_file_images(images, "wav", dtype=torch.int16)


@gpu
def test_dense_truncation_and_unsupported_wav(ing):
    x = noise_int16(1000, seed=1)
    pcm, lengths = ing.ingest_file_images([wav_image(x), wav_image(x[:10])], "wav", row_stride=640)
    torch.cuda.synchronize()
    assert pcm.shape == (2, 640) and lengths.tolist() == [640, 10]
    assert np.array_equal(pcm[0].cpu().numpy(), x[:640])
    # a damaged file whose data chunk ends inside a sample frame: the partial frame is dropped
    cut = wav_image(x)[:-1]
    pcm, lengths = ing.ingest_file_images([cut], "wav")
    torch.cuda.synchronize()
    assert lengths.tolist() == [999] and np.array_equal(pcm[0, :999].cpu().numpy(), x[:999])
    img8 = bytearray(wav_image(x))
    img8[34:36] = struct.pack("<H", 8)
    with pytest.raises(ValueError, match="16-bit"):
        ing.ingest_file_images([bytes(img8)], "wav")


@gpu
def test_load_audio_file_wrapper_and_frontend_chain(ing, tmp_path):
    """load_audio_file mirror (per example) and the batched chain file bytes -> ingest -> front-end:
    identical to the front-end fed the decoded samples; the reference's fixture (silence) gives its golden."""
    import speech_recognition_b200 as pkg

    xs = [noise_int16(n, seed=n) for n in (16000, 9999, 24001)]
    paths = []
    for i, x in enumerate(xs):
        p = tmp_path / ("u%d.wav" % i)
        p.write_bytes(wav_image(x))
        paths.append(str(p))
    (tmp_path / "k.pcm").write_bytes(xs[1].astype("<i2").tobytes() + b"\x07")
    audio = ing.load_audio_file(16000, "wav")(paths[0])
    assert audio.dtype == torch.float32 and audio.shape == (16000,)
    assert audio.cpu().numpy().tobytes() == P.load_audio_pcm16(wav_image(xs[0]), "wav").tobytes()
    raw = ing.load_audio_file(16000, "pcm")(str(tmp_path / "k.pcm"))
    assert raw.shape == (10000,) and float(raw[-1]) == 7 / 32768.0

    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    pcm, lengths = ing.ingest_files(paths, "wav")
    got, nf = fe(pcm, lengths=lengths, t_max=fe.num_frames(24001))
    L = pcm.shape[1]
    direct = np.zeros((3, L), np.int16)
    for i, x in enumerate(xs):
        direct[i, :x.size] = x
    want, nf2 = fe(torch.from_numpy(direct).cuda(), lengths=[x.size for x in xs])
    torch.cuda.synchronize()
    assert torch.equal(nf, nf2) and torch.equal(got, want)

    with open(os.path.join(GOLDEN_DIR, "ref_test_wav_header.bin"), "rb") as f:
        silence = f.read() + bytes(132300)
    pcm, lengths = ing.ingest_file_images([silence, silence], "wav")
    out, nf = fe(pcm, lengths=lengths)
    g = np.load(os.path.join(GOLDEN_DIR, "ref_wav_dataset_tfrecord.npz"))
    assert nf.tolist() == [412, 412] and np.array_equal(out[0].cpu().numpy(), g["feature_0"])


@gpu
def test_full_size_round_trip_and_mask_property(ing):
    """BASELINE config-2 size (256 x 15 s) through size-independent properties: ingest(bytes) gives back the
    samples bit for bit for every payload alignment; without SpecAugment the models' padding mask of the
    front-end's batch is exactly `any frame of the stride group < n_frames`."""
    import speech_recognition_b200 as pkg
    from speech_recognition_b200 import batching

    B, L = 256, 240000
    rng = np.random.default_rng(11)
    x = rng.integers(-32768, 32768, (B, L), dtype=np.int16)
    lens = rng.integers(160, L + 1, B)
    lens[:4] = [L, 0, 319, 320]
    header = wav_image(np.zeros(0, np.int16))[:40]
    images = []
    for b in range(B):
        body = x[b, :lens[b]].astype("<i2").tobytes()
        pad = b"z" * (b % 5)                                   # odd / even junk chunks shift the payload alignment
        extra = (b"LIST" + struct.pack("<I", len(pad)) + pad + (b"\x00" if len(pad) & 1 else b"")) if pad else b""
        img = header[:36] + extra + b"data" + struct.pack("<I", len(body)) + body
        images.append(img[:4] + struct.pack("<I", len(img) - 8) + img[8:])
    pcm, lengths = ing.ingest_file_images(images, "wav", row_stride=L)
    torch.cuda.synchronize()
    assert lengths.cpu().numpy().tolist() == lens.tolist()
    got = pcm.cpu().numpy()
    valid = np.arange(L)[None, :] < lens[:, None]
    assert np.array_equal(got[valid], x[valid]) and not got[~valid].any()

    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI)
    feats, nf = fe(pcm, lengths=lengths)
    mask = batching.audio_mask(feats, [3, 3], [2, 2]).cpu().numpy()
    nfr = np.minimum(nf.cpu().numpy(), feats.shape[1])
    want = (np.arange(mask.shape[1])[None, :] * 4) < nfr[:, None]
    assert mask.shape == (B, 374) and np.array_equal(mask, want)
