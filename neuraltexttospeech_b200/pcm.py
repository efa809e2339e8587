"""WAV framing for the 16-bit PCM the generator emits (`Generator.synthesize_pcm16`).

The reference intends this step right after the vocoder (Grad-TTS_TF/inference.py:89-92:
`audio = (vocoder.forward(y_dec)...clamp(-1, 1).numpy() * 32768).astype(np.int16)` then
`write(path, 22050, audio)`; same pattern FastSpeech2_TF/FastSpeech2/utils/model.py:74-92).  The
quantisation itself runs on the GPU inside the conv_post epilogue (csrc/kernels.cuh `pcm16_of`);
this module only frames the int16 samples as RIFF/WAVE -- there is no arithmetic here.
"""
from __future__ import annotations

import wave
from typing import Optional, Sequence

import numpy as np


def write_wav(path: str, pcm: np.ndarray, sampling_rate: int = 22050) -> None:
    """One utterance of int16 mono PCM -> RIFF/WAVE file."""
    pcm = np.asarray(pcm)
    if pcm.dtype != np.int16:
        raise TypeError(f"pcm must be int16 (Generator.synthesize_pcm16 output), got {pcm.dtype}")
    pcm = pcm.reshape(-1)
    with wave.open(path, "wb") as f:
        f.setnchannels(1)
        f.setsampwidth(2)
        f.setframerate(int(sampling_rate))
        f.writeframes(pcm.astype("<i2", copy=False).tobytes())


def write_wavs(paths: Sequence[str], pcm: np.ndarray, sampling_rate: int = 22050,
               lengths: Optional[Sequence[int]] = None) -> None:
    """A batch [B, L] of PCM -> one file per utterance, optionally trimmed to `lengths` samples
    (the `wavs[i][:lengths[i]]` trim of FastSpeech2_TF/FastSpeech2/utils/model.py:88-90)."""
    pcm = np.asarray(pcm)
    if pcm.ndim != 2 or len(paths) != pcm.shape[0]:
        raise ValueError("pcm must be [B, L] with one path per utterance")
    for i, p in enumerate(paths):
        n = pcm.shape[1] if lengths is None else int(lengths[i])
        write_wav(p, pcm[i, :n], sampling_rate)
