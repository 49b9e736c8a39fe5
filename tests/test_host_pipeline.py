"""Streaming form of the host-buffer call (b200mel_forward_host_submit / _wait): two staging slots in flight
give the same results as the synchronous call, step by step."""
import numpy as np
import pytest
import torch

from .helpers import LIBRI, noise_int16


def test_slot_and_handle_validation(built_lib):
    from speech_recognition_b200 import _lib

    lib = _lib.load()
    assert lib.b200mel_forward_host_wait(None, 0) == _lib.ERR_INVALID_ARGUMENT
    assert lib.b200mel_forward_host_submit(None, None, 0, None, None, 1, 0, 0, None, 0, None, 0, 0, 0, 0) == _lib.ERR_INVALID_ARGUMENT
    assert b"null handle" in lib.b200mel_last_error()


@pytest.mark.gpu
@pytest.mark.parametrize("delta", [False, True])
def test_submit_wait_matches_synchronous_call(built_lib, delta):
    import speech_recognition_b200 as pkg

    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=delta,
                             spec_augment=dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2))
    B, L = 24, 32000
    steps = 5
    ins = [torch.from_numpy(np.stack([noise_int16(L, seed=100 * s + b) for b in range(B)])).pin_memory() for s in range(steps)]
    lens = [np.array([L - 37 * ((b + s) % 7) for b in range(B)], np.int32) for s in range(steps)]
    want = [fe.forward_host(ins[s], lengths=lens[s], seed=7, first_utterance_id=B * s) for s in range(steps)]
    want = [(w[0].copy(), w[1].copy()) for w in want]
    T, C = want[0][0].shape[1], want[0][0].shape[3]
    outs = [torch.empty((B, T, 80, C), dtype=torch.float32).pin_memory() for _ in range(2)]
    got = []
    res = {0: fe.forward_host_submit(ins[0], lengths=lens[0], seed=7, first_utterance_id=0, out=outs[0], slot=0)}
    for s in range(steps):
        if s + 1 < steps:
            res[(s + 1) % 2] = fe.forward_host_submit(ins[s + 1], lengths=lens[s + 1], seed=7,
                                                      first_utterance_id=B * (s + 1), out=outs[(s + 1) % 2], slot=(s + 1) % 2)
        fe.forward_host_wait(s % 2)
        o, nf = res[s % 2]
        got.append((o.numpy().copy(), nf.copy()))
    for s in range(steps):
        assert np.array_equal(got[s][1], want[s][1])
        assert got[s][0].tobytes() == want[s][0].tobytes(), "step %d differs" % s
    # waiting on an idle slot is a no-op; a re-submit on a busy slot waits for it first
    fe.forward_host_wait(0)
    fe.forward_host_wait(1)
    a = fe.forward_host_submit(ins[0], lengths=lens[0], seed=7, out=outs[0], slot=0)
    b = fe.forward_host_submit(ins[1], lengths=lens[1], seed=7, first_utterance_id=B, out=outs[1], slot=0)
    fe.forward_host_wait(0)
    assert outs[1].numpy().tobytes() == want[1][0].tobytes()
    with pytest.raises(ValueError):
        fe.forward_host_submit(ins[0], out=outs[0], slot=2)


@pytest.mark.gpu
def test_bound_step_equals_the_call(built_lib):
    """FeatureFrontend.bind: arguments checked once, the returned callable is the b200mel_forward call alone."""
    import speech_recognition_b200 as pkg

    sa = dict(enable=True, F=27, m_F=2, T=100, p=1.0, m_T=2)
    fe = pkg.FeatureFrontend("log-mel-spectrogram", **LIBRI, use_delta_accelerate=True, spec_augment=sa)
    B, L = 5, 48000
    pcm = torch.from_numpy(np.stack([noise_int16(L, 40 + i) for i in range(B)])).cuda()
    lengths = torch.tensor([48000, 47999, 20000, 320, 100], dtype=torch.int32, device="cuda")
    want, nf, st = fe(pcm, lengths=lengths, seed=7, first_utterance_id=3, return_status=True)
    out = torch.full_like(want, float("nan"))
    nf2 = torch.zeros_like(nf)
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    step = fe.bind(pcm, lengths, out, nf2, seed=7, first_utterance_id=3, status=status)
    step()
    step()                                                        # (a second time: same buffers, same result)
    torch.cuda.synchronize()
    assert torch.equal(out, want) and torch.equal(nf2, nf)
    assert int(status.item()) == int(st.item())                   # (utterance 4 has no frame: its time-mask draw is refused, bit 0)
    with pytest.raises(ValueError):
        fe.bind(pcm, lengths.to(torch.int64), out, nf2)
    with pytest.raises(ValueError):
        fe.bind(pcm, lengths, out[:, :10], nf2)
    with pytest.raises(ValueError):
        fe.bind(pcm.cpu(), lengths, out, nf2)
