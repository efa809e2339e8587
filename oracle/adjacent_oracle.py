"""CPU ORACLE for the steps either side of the generator -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may import this file.

  * pcm16()            -- the step AFTER the path: the reference's intended post-vocoder quantisation
                          `(audio.clamp(-1, 1).numpy() * 32768).astype(np.int16)`
                          (Grad-TTS_TF/inference.py:89; same pattern
                          FastSpeech2_TF/FastSpeech2/utils/model.py:82-85 with max_wav_value 32768).
  * mel_spectrogram()  -- the step BEFORE the path: `STFT.mel_spectrogram`
                          (HiFiGAN_TF/common/audio_processing_tf.py:41-104, used data.py:142,210, gan.py:48-52,147).

PARITY UNPINNED at the TensorFlow boundary (tf.signal is not importable here and the reference holds no golden
mels): tf.signal's published semantics are restated below and pinned by known-answer tests instead
(tests/test_adjacent_oracle.py).
  * tf.signal.stft(x, frame_length, frame_step, fft_length): window_fn = hann_window(periodic=True),
    pad_end=False -> frames = 1 + (N - frame_length) // frame_step (0 if N < frame_length); frame f covers
    samples [f*step, f*step + frame_length); rfft of length fft_length -> fft_length//2 + 1 bins.  No centre padding.
  * tf.signal.hann_window(n, periodic=True): 0.5 - 0.5 cos(2 pi i / n).
  * tf.signal.linear_to_mel_weight_matrix(num_mel_bins, num_spectrogram_bins, sample_rate, lower, upper): HTK mel
    scale mel(f) = 1127 ln(1 + f/700); the DC bin is zeroed (bands_to_zero = 1); triangle m has edges
    linspace(mel(lower), mel(upper), num_mel_bins + 2)[m : m + 3]; weight = max(0, min(up-slope, down-slope)); no area
    normalisation (unlike librosa's Slaney filters).
  * reference quirks kept: POWER spectrogram (tf.math.square(tf.abs(.)), audio_processing_tf.py:93-96), then
    dynamic_range_compression = log(clip(x, 1e-5) * 1) (audio_processing_tf.py:27-32).
"""
from __future__ import annotations

import numpy as np


def pcm16(audio):
    """fp32 waveform in [-1, 1] -> int16: clamp, * 32768, truncate toward zero (numpy astype), with the single
    out-of-range value +1.0 * 32768 = 32768 saturated to 32767 (numpy's cast would wrap it to -32768)."""
    a = np.clip(np.asarray(audio, dtype=np.float32), -1.0, 1.0) * np.float32(32768.0)
    return np.minimum(np.trunc(a), 32767.0).astype(np.int16)


def hertz_to_mel(f):
    return 1127.0 * np.log1p(np.asarray(f, dtype=np.float64) / 700.0)


def mel_weight_matrix(num_mel_bins=80, num_spectrogram_bins=513, sample_rate=22050, lower=0.0, upper=8000.0):
    """tf.signal.linear_to_mel_weight_matrix restated: [num_spectrogram_bins, num_mel_bins] float64."""
    nyquist = sample_rate / 2.0
    lin = np.linspace(0.0, nyquist, num_spectrogram_bins)[1:]
    spec_mel = hertz_to_mel(lin)[:, None]
    edges = np.linspace(hertz_to_mel(lower), hertz_to_mel(upper), num_mel_bins + 2)
    lo, ce, up = edges[:-2][None, :], edges[1:-1][None, :], edges[2:][None, :]
    w = np.maximum(0.0, np.minimum((spec_mel - lo) / (ce - lo), (up - spec_mel) / (up - ce)))
    return np.pad(w, [[1, 0], [0, 0]])


def hann_periodic(n):
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


def num_frames(n_samples, frame_length=1024, frame_step=256):
    return 0 if n_samples < frame_length else 1 + (n_samples - frame_length) // frame_step


def mel_spectrogram(wav, frame_length=1024, frame_step=256, fft_length=1024, n_mel_channels=80, sampling_rate=22050,
                    mel_fmin=0.0, mel_fmax=8000.0, dtype=np.float64):
    """wav [B, N] (or [N]) -> log-mel [B, T, n_mel_channels], T = num_frames(N) (audio_processing_tf.py:62-104)."""
    if frame_length != fft_length:
        # the reference builds the filterbank for frame_length // 2 + 1 bins (audio_processing_tf.py:55) but the STFT has
        # fft_length // 2 + 1: its matmul raises for these arguments, and so does this restatement
        raise ValueError(f"frame_length ({frame_length}) != filter_length ({fft_length}): the reference's "
                         "spectrogram @ mel_filterbank has mismatched inner dimensions")
    x = np.asarray(wav, dtype=dtype)
    squeeze = x.ndim == 1
    if squeeze:
        x = x[None]
    T = num_frames(x.shape[1], frame_length, frame_step)
    idx = np.arange(T)[:, None] * frame_step + np.arange(frame_length)[None, :]
    frames = x[:, idx] * hann_periodic(frame_length).astype(dtype)[None, None, :]
    spec = np.fft.rfft(frames, n=fft_length, axis=-1)
    power = (np.abs(spec) ** 2).astype(dtype)
    mel = power @ mel_weight_matrix(n_mel_channels, frame_length // 2 + 1, sampling_rate, mel_fmin, mel_fmax).astype(dtype)   # audio_processing_tf.py:55
    out = np.log(np.clip(mel, 1e-5, None))
    return out[0] if squeeze else out
