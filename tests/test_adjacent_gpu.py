"""GPU parity tests of the steps either side of the generator, through the C ABI:
16-bit PCM output fused into the conv_post epilogue, and the mel front-end kernel."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
from oracle import adjacent_oracle as A
from oracle import hifigan_oracle as O


@pytest.mark.parametrize("v", "123")
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_pcm16_is_the_quantised_fp32_output_bit_exact(v, precision):
    h = HParams.builtin(v)
    g = Generator(h, precision=precision).load_weights(synthetic_weights(h, "B"))
    mel = synthetic_mel(3, 19, h.num_mels)
    audio = g(mel)
    pcm = g.synthesize_pcm16(mel)
    assert pcm.dtype == np.int16 and pcm.shape == audio.shape[:2]
    assert (pcm == A.pcm16(audio[..., 0])).all()               # same kernel, same tanh: integer work is bit-exact
    # and within the path's waveform tolerance of the quantised oracle
    ref = A.pcm16(O.generator_forward(h, synthetic_weights(h, "B"), mel, "torch32")[..., 0])
    tol = 1 + int(32768 * (1e-5 if precision == "fp32" else 2e-3))
    assert np.abs(pcm.astype(np.int32) - ref.astype(np.int32)).max() <= tol
    g.close()


def test_pcm16_saturates_and_device_path_matches_host_path():
    import torch
    h = HParams.builtin("2")
    w = synthetic_weights(h, "B")
    w["conv_post.b"] = np.array([30.0], np.float32)            # tanh -> exactly 1.0f everywhere: must give 32767, not wrap
    g = Generator(h, precision="fp16").load_weights(w)
    mel = synthetic_mel(2, 8, h.num_mels)
    pcm = g.synthesize_pcm16(mel)
    assert (pcm == 32767).all()
    w["conv_post.b"] = np.array([-30.0], np.float32)
    g.load_weights(w)
    assert (g.synthesize_pcm16(mel) == -32768).all()
    w["conv_post.b"] = np.array([0.01], np.float32)
    g.load_weights(w)
    host = g.synthesize_pcm16(mel)
    dev = g.synthesize_pcm16(torch.from_numpy(mel).cuda())
    assert dev.is_cuda and dev.dtype == torch.int16
    assert (dev.cpu().numpy() == host).all()
    assert g.synthesize_pcm16(np.zeros((0, 8, 80), np.float32)).shape == (0, 8 * 256)
    g.close()


def _speechlike(B, N, seed=0):
    rng = np.random.default_rng(seed)
    t = np.arange(N)[None, :] / 22050.0
    f0 = rng.uniform(90, 250, size=(B, 1))
    x = sum(rng.uniform(0.02, 0.3, size=(B, 1)) * np.sin(2 * np.pi * f0 * h * t + rng.uniform(0, 6.28, size=(B, 1))) for h in range(1, 30))
    x = x / np.abs(x).max() * 0.8 + 0.01 * rng.standard_normal((B, N))
    return x.astype(np.float32)


@pytest.mark.parametrize("B,N", [(1, 1024), (3, 1024 + 256 * 6), (2, 22050), (1, 1024 + 255)])
def test_mel_frontend_matches_oracle(B, N):
    from neuraltexttospeech_b200 import STFT
    wav = _speechlike(B, N)
    fe = STFT()
    mel = fe.mel_spectrogram(wav)
    ref = A.mel_spectrogram(wav.astype(np.float64))
    if B == 1:
        ref = ref[0]
    assert mel.shape == ref.shape and mel.dtype == np.float32
    # fp32 FFT vs the fp64 oracle, compared in the log domain the generator consumes
    assert np.abs(mel - ref).max() <= 2e-3
    assert np.abs(mel - ref).mean() <= 2e-5
    fe.close()


def test_mel_frontend_edge_cases_and_device_path():
    import torch
    from neuraltexttospeech_b200 import STFT
    fe = STFT()
    assert fe.mel_spectrogram(np.zeros(1000, np.float32)).shape == (0, 80)          # shorter than one frame
    assert fe.mel_spectrogram(np.zeros((0, 4096), np.float32)).shape == (0, 13, 80)  # empty batch
    sil = fe.mel_spectrogram(np.zeros(4096, np.float32))
    assert np.allclose(sil, np.log(1e-5))                                           # silence sits on the 1e-5 floor
    wav = _speechlike(2, 8192, seed=3)
    host = fe.mel_spectrogram(wav)
    dev = fe.mel_spectrogram(torch.from_numpy(wav).cuda())
    assert dev.is_cuda and (dev.cpu().numpy() == host).all()
    with pytest.raises(ValueError):
        STFT(filter_length=1000)
    # other supported FFT sizes (frame_length == filter_length, the only arguments the reference's own matmul accepts)
    for nfft, fl, st in ((256, 256, 64), (4096, 4096, 1024)):
        f2 = STFT(filter_length=nfft, frame_length=fl, frame_step=st)
        m = f2.mel_spectrogram(wav)
        r = A.mel_spectrogram(wav.astype(np.float64), frame_length=fl, frame_step=st, fft_length=nfft)
        assert m.shape == r.shape and np.abs(m - r).max() <= 2e-3
        f2.close()
    fe.close()


def test_wav_to_mel_to_wav_round_trip_runs_on_device():
    """Front-end output feeds the generator directly (same layout); PCM comes back as int16."""
    import torch
    from neuraltexttospeech_b200 import STFT
    h = HParams.builtin("3")
    g = Generator(h, precision="fp16").load_weights(synthetic_weights(h, "A"))
    fe = STFT()
    wav = torch.from_numpy(_speechlike(2, 1024 + 256 * 15)).cuda()
    mel = fe.mel_spectrogram(wav)
    assert mel.shape == (2, 16, 80)
    audio = g(mel)
    pcm = g.synthesize_pcm16(mel)
    assert audio.shape == (2, 16 * 256, 1) and pcm.shape == (2, 16 * 256)
    ref = O.generator_forward(h, synthetic_weights(h, "A"), mel.cpu().numpy(), "torch32")
    assert np.abs(audio.cpu().numpy() - ref).max() <= 2e-3
    g.close(); fe.close()


def test_mel_frontend_matches_committed_fixture(golden_dir):
    import os
    from neuraltexttospeech_b200 import STFT
    z = np.load(os.path.join(golden_dir, "adjacent_frontend.npz"))
    fe = STFT()
    mel = fe.mel_spectrogram(z["wav"])
    assert mel.shape == z["mel"].shape and np.abs(mel - z["mel"]).max() <= 2e-3
    fe.close()


def test_mel_frontend_rejects_frame_length_other_than_filter_length():
    # the reference's own matmul fails for these constructor arguments (audio_processing_tf.py:55, 85-97)
    from neuraltexttospeech_b200 import STFT
    with pytest.raises(ValueError):
        STFT(filter_length=1024, frame_length=512)
