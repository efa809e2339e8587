"""CPU checks of the oracle for the steps either side of the generator (PCM writer, mel front-end)
and of the host-side WAV framing.  Known-answer tests stand in for the golden vectors the reference does not hold."""
import os
import wave

import numpy as np
import pytest

from neuraltexttospeech_b200.pcm import write_wav, write_wavs
from oracle import adjacent_oracle as A


def test_pcm16_known_answers():
    x = np.array([0.0, 1.0, -1.0, 2.0, -2.0, 0.5, -0.5, 1.0 / 32768, -1.0 / 32768, 0.99999, 1.5 / 32768, -1.5 / 32768], np.float32)
    want = np.array([0, 32767, -32768, 32767, -32768, 16384, -16384, 1, -1, 32767, 1, -1], np.int16)
    assert A.pcm16(x).tolist() == want.tolist()
    # identical to the reference's expression wherever that expression is well defined (|x| < 1)
    rng = np.random.default_rng(0)
    y = rng.uniform(-0.999, 0.999, 10000).astype(np.float32)
    assert (A.pcm16(y) == (np.clip(y, -1, 1) * 32768).astype(np.int16)).all()


def test_wav_round_trip(tmp_path):
    rng = np.random.default_rng(1)
    pcm = rng.integers(-32768, 32767, size=(2, 1000), dtype=np.int16)
    p = str(tmp_path / "a.wav")
    write_wav(p, pcm[0], 22050)
    with wave.open(p, "rb") as f:
        assert (f.getnchannels(), f.getsampwidth(), f.getframerate(), f.getnframes()) == (1, 2, 22050, 1000)
        assert (np.frombuffer(f.readframes(1000), "<i2") == pcm[0]).all()
    paths = [str(tmp_path / f"{i}.wav") for i in range(2)]
    write_wavs(paths, pcm, 22050, lengths=[10, 1000])
    with wave.open(paths[0], "rb") as f:
        assert f.getnframes() == 10
    with pytest.raises(TypeError):
        write_wav(p, pcm[0].astype(np.float32))


def test_mel_weight_matrix_known_properties():
    M = A.mel_weight_matrix()
    assert M.shape == (513, 80)
    assert (M[0] == 0).all()                                   # DC bin zeroed
    assert (M >= 0).all() and M.max() <= 1.0
    # HTK triangles: each has a single peak, supports overlap only with neighbours, nothing above mel_fmax
    f = np.linspace(0, 11025, 513)
    assert (M[f > 8000.0] == 0).all()
    centres = 700.0 * np.expm1(np.linspace(A.hertz_to_mel(0.0), A.hertz_to_mel(8000.0), 82)[1:-1] / 1127.0)
    peak_hz = f[M.argmax(axis=0)]
    assert np.abs(peak_hz - centres).max() <= 11025 / 512     # peak within one bin of the triangle centre
    assert A.hertz_to_mel(1000.0) == pytest.approx(1127.0 * np.log(1 + 1000 / 700.0))
    # interior bins between the first and last centre are covered by exactly two triangles summing to ~1
    inner = (f > centres[0]) & (f < centres[-1])
    assert np.abs(M[inner].sum(axis=1) - 1.0).max() < 1e-9


def test_mel_spectrogram_frames_and_pure_tone():
    assert A.num_frames(1023) == 0 and A.num_frames(1024) == 1 and A.num_frames(1024 + 255) == 1 and A.num_frames(1280) == 2
    assert A.num_frames(220672) == 859                         # no centre padding: 10 s -> 859 frames, not 862
    sr, n = 22050, 1024 + 256 * 7
    k = 64                                                     # bin-centred tone: f = k * sr / 1024
    t = np.arange(n)
    wav = 0.5 * np.cos(2 * np.pi * k * t / 1024)
    mel = A.mel_spectrogram(wav)
    assert mel.shape == (8, 80)
    # periodic Hann of a bin-centred tone: |X[k]| = A*N/4, |X[k+-1]| = A*N/8, other bins ~0
    power = np.zeros(513)
    power[k] = (0.5 * 1024 / 4) ** 2
    power[k - 1] = power[k + 1] = (0.5 * 1024 / 8) ** 2
    want = np.log(np.clip(power @ A.mel_weight_matrix(), 1e-5, None))
    assert np.abs(mel - want[None]).max() < 1e-6
    # silence hits the 1e-5 floor
    assert np.allclose(A.mel_spectrogram(np.zeros(2048)), np.log(1e-5))
    # batch and dtype
    b = A.mel_spectrogram(np.stack([wav, wav]).astype(np.float32), dtype=np.float32)
    assert b.shape == (2, 8, 80) and b.dtype == np.float32 and np.abs(b[0] - mel).max() < 1e-2


def test_oracle_reproduces_committed_adjacent_fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "adjacent_frontend.npz"))
    mel = A.mel_spectrogram(z["wav"].astype(np.float64))
    assert mel.shape == z["mel"].shape == (2, 10, 80)
    assert np.abs(mel - z["mel"]).max() < 1e-5
    assert (A.pcm16(z["wav"]) == z["pcm"]).all()


def test_frame_length_must_equal_filter_length_like_the_reference():
    # audio_processing_tf.py:55: the filterbank has frame_length // 2 + 1 bins, the STFT fft_length // 2 + 1: the reference's
    # matmul raises when they differ; the restatement (and the CUDA front-end, tests/test_adjacent_gpu.py) refuse too
    import pytest
    wav = np.zeros(4096, np.float32)
    with pytest.raises(ValueError):
        A.mel_spectrogram(wav, frame_length=512, fft_length=1024)
    assert A.mel_spectrogram(wav, frame_length=1024, fft_length=1024).shape == (13, 80)
