"""Drop-in `STFT` for the reference's HiFiGAN_TF/common/audio_processing_tf.py:41-104 -- the mel front-end that
feeds the generator.

Same constructor arguments and the same `mel_spectrogram(x)` method as the reference class; the arithmetic
(periodic-Hann STFT without centre padding -> POWER spectrogram -> HTK mel filterbank -> log(clip(., 1e-5)))
runs in one hand-written CUDA kernel (csrc/mel_frontend.cu) reached through the C ABI.  No TensorFlow, no CPU path.

  x of shape (wav_length,) or (1, wav_length)  ->  (T, n_mel_channels)      like the reference
  x of shape (B, wav_length), B > 1            ->  (B, T, n_mel_channels)   batched extension = generator input
with T = 1 + (wav_length - frame_length) // frame_step.  numpy in -> numpy out; torch CUDA in -> torch CUDA out.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

try:
    import torch
except Exception:                       # pragma: no cover
    torch = None


class STFT:
    def __init__(self, filter_length=1024, frame_length=1024, frame_step=256, n_mel_channels=80,
                 sampling_rate=22050, mel_fmin=0.0, mel_fmax=8000.0, device: int = 0):
        self.frame_length = int(frame_length)
        self.frame_step = int(frame_step)
        self.fft_length = int(filter_length)
        self.sampling_rate = int(sampling_rate)
        self.n_mel_channels = int(n_mel_channels)
        self.mel_fmin = float(mel_fmin)
        self.mel_fmax = float(mel_fmax)
        self.device = int(device)
        self._lib = _lib.load()
        cfg = _lib.HfgMelConfig(self.fft_length, self.frame_length, self.frame_step, self.n_mel_channels,
                                self.sampling_rate, self.mel_fmin, self.mel_fmax)
        self._handle = C.c_void_p()
        _lib.check(self._lib.hfg_mel_create(C.byref(cfg), self.device, C.byref(self._handle)), None, "hfg_mel_create", mel=True)

    def num_frames(self, wav_length: int) -> int:
        n = C.c_int()
        _lib.check(self._lib.hfg_mel_num_frames(self._handle, int(wav_length), C.byref(n)), self._handle, mel=True)
        return int(n.value)

    def mel_spectrogram(self, x, check_shape=True):
        is_torch = torch is not None and isinstance(x, torch.Tensor)
        shape = tuple(x.shape)
        if len(shape) == 1:
            batched, B, N = False, 1, shape[0]
        elif len(shape) == 2:
            # reference audio_processing_tf.py:80-83 squeezes a leading axis of 1
            batched, B, N = shape[0] != 1, shape[0], shape[1]
        else:
            raise ValueError(f"Invalid signal shape {shape}, expected (wav_length,) or (B, wav_length)")
        T = self.num_frames(N)
        if is_torch and x.is_cuda:
            wav = x.float().contiguous()
            if wav.device.index != self.device:
                raise ValueError(f"signal is on cuda:{wav.device.index}, front-end on cuda:{self.device}")
            mel = torch.empty((B, T, self.n_mel_channels), dtype=torch.float32, device=wav.device)
            if B and T:
                stream = torch.cuda.current_stream(wav.device).cuda_stream
                _lib.check(self._lib.hfg_mel_forward(self._handle, C.c_void_p(wav.data_ptr()), B, N, C.c_void_p(mel.data_ptr()),
                                                     C.c_void_p(stream)), self._handle, "hfg_mel_forward", mel=True)
            return mel if batched else mel[0]
        wav = np.ascontiguousarray(x.detach().numpy() if is_torch else np.asarray(x), dtype=np.float32).reshape(B, N)
        mel = np.empty((B, T, self.n_mel_channels), dtype=np.float32)
        if B and T:
            _lib.check(self._lib.hfg_mel_forward_host(self._handle, wav.ctypes.data_as(C.c_void_p), B, N,
                                                      mel.ctypes.data_as(C.c_void_p)), self._handle, "hfg_mel_forward_host", mel=True)
        out = mel if batched else mel[0]
        return torch.from_numpy(out) if is_torch else out

    def close(self) -> None:
        if getattr(self, "_handle", None) is not None and self._handle:
            self._lib.hfg_mel_destroy(self._handle)
            self._handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
