"""Drop-in `Generator` for reference HiFiGAN_TF/model.py:115-173.

Same constructor argument (`hparams`), same `call(mel) -> audio` / `__call__(mel, training=...)`
signature and channels-last layout ([B, T, num_mels] -> [B, T*prod(upsample_rates), 1], fp32,
reference HiFiGAN_TF/train.py:70-74).  All arithmetic runs in hand-written sm_100a CUDA kernels
reached through the C ABI of libhifigan_b200.so (ctypes); no TensorFlow, no Triton, no CPU path.

Differences to the Keras model that are on purpose:
  * forward only: `training=True` raises NotImplementedError (`.trainable_variables`, `.save`
    are training/serialisation surface, out of scope -- SURVEY.md section 8b);
  * weights are given explicitly (`load_weights`) under the names of SURVEY.md section 8b, in the
    Keras layouts; weight norm (tfa WeightNormalization, data_init=False) is folded once at
    load instead of on every call;
  * `precision`: "fp32" (FFMA validation mode), "fp16" (tcgen05 tensor cores, default) or "bf16".
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Mapping, Optional, Union

import numpy as np

from . import _lib
from .hparams import coerce_hparams
from .synthetic import synthetic_weights
from .topology import enumerate_layers, total_upsample

try:                                    # torch is plumbing only (device buffers / streams)
    import torch
except Exception:                       # pragma: no cover
    torch = None


def _config_from_hparams(h, compat: str) -> _lib.HfgConfig:
    if compat not in ("reference", "upstream"):
        raise ValueError("compat must be 'reference' or 'upstream'")
    cfg = _lib.HfgConfig()
    cfg.num_mels = int(h.num_mels)
    cfg.upsample_initial_channel = int(h.upsample_initial_channel)
    # reference model.py:126: ResBlock1 if hparams.resblock == '1' else ResBlock2
    cfg.resblock = 1 if h.resblock == "1" else 2
    rates, ksz = list(h.upsample_rates), list(h.upsample_kernel_sizes)
    if len(rates) != len(ksz):
        raise ValueError("upsample_rates and upsample_kernel_sizes differ in length")
    if len(rates) > _lib.HFG_MAX_UPSAMPLES:
        raise ValueError("too many upsample stages")
    cfg.num_upsamples = len(rates)
    for i, (u, k) in enumerate(zip(rates, ksz)):
        cfg.upsample_rates[i] = int(u)
        cfg.upsample_kernel_sizes[i] = int(k)
    rks, rds = list(h.resblock_kernel_sizes), list(h.resblock_dilation_sizes)
    if len(rks) != len(rds):
        raise ValueError("resblock_kernel_sizes and resblock_dilation_sizes differ in length")
    if len(rks) > _lib.HFG_MAX_KERNELS:
        raise ValueError("too many resblock kernels")
    cfg.num_kernels = len(rks)
    for j, (k, ds) in enumerate(zip(rks, rds)):
        ds = list(ds)
        # ResBlock1 consumes dilation[0..2], ResBlock2 dilation[0..1] (model.py:21-66, 84-101)
        need = 3 if cfg.resblock == 1 else 2
        if len(ds) < need:
            raise ValueError(f"resblock_dilation_sizes[{j}] needs {need} entries")
        ds = ds[:need]
        cfg.resblock_kernel_sizes[j] = int(k)
        cfg.num_dilations[j] = len(ds)
        for m, d in enumerate(ds):
            cfg.resblock_dilation_sizes[j][m] = int(d)
    cfg.compat = 0 if compat == "reference" else 1
    return cfg


class _PinnedBlock:
    """One page-locked host allocation (hfg_host_alloc) exposed through the array interface.  A numpy array built on
    it keeps it alive through `.base`; when the last view is dropped the block goes back to its pool, so a loop of
    `audio = generator(mel)` calls reuses the same few pinned buffers instead of allocating (and page-faulting) a fresh
    pageable array every call."""

    def __init__(self, pool, ptr: int, nbytes: int, shape, typestr: str):
        self._pool, self._ptr, self._nbytes = pool, ptr, nbytes
        self.__array_interface__ = {"version": 3, "data": (ptr, False), "shape": tuple(shape), "typestr": typestr}

    def __del__(self):
        try:
            self._pool.release(self._ptr, self._nbytes)
        except Exception:
            pass


class _PinnedPool:
    """Free lists of pinned host buffers by size (at most `keep` cached per size)."""

    def __init__(self, lib, keep: int = 3):
        self._lib, self._keep, self._free, self._closed = lib, keep, {}, False

    def array(self, shape, dtype) -> np.ndarray:
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        if nbytes == 0 or self._closed:
            return np.empty(shape, dtype)
        lst = self._free.get(nbytes)
        if lst:
            ptr = lst.pop()
        else:
            p = C.c_void_p()
            _lib.check(self._lib.hfg_host_alloc(C.byref(p), nbytes), None, "hfg_host_alloc")
            ptr = int(p.value)
        return np.asarray(_PinnedBlock(self, ptr, nbytes, shape, dtype.str))

    def release(self, ptr: int, nbytes: int) -> None:
        lst = self._free.setdefault(nbytes, [])
        if self._closed or len(lst) >= self._keep:
            self._lib.hfg_host_free(C.c_void_p(ptr))
        else:
            lst.append(ptr)

    def close(self) -> None:
        self._closed = True
        for lst in self._free.values():
            for ptr in lst:
                self._lib.hfg_host_free(C.c_void_p(ptr))
        self._free = {}


class Generator:
    def __init__(self, hparams, precision: str = "fp16", device: int = 0, compat: str = "reference"):
        self.hparams = coerce_hparams(hparams)
        if precision not in _lib.PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_lib.PRECISIONS)}")
        self.precision = precision
        self.device = int(device)
        self.compat = compat
        self.num_kernels = len(self.hparams.resblock_kernel_sizes)
        self.num_upsamples = len(self.hparams.upsample_rates)
        self._layers = enumerate_layers(self.hparams)
        self._total_up = total_upsample(self.hparams)
        self._lib = _lib.load()
        self._handle = C.c_void_p()
        cfg = _config_from_hparams(self.hparams, compat)
        _lib.check(self._lib.hfg_create(C.byref(cfg), self.device, C.byref(self._handle)), None, "hfg_create")
        self._weights_loaded = False
        self._ws = None           # torch workspace for the device-pointer path
        self._pinned = _PinnedPool(self._lib)    # outputs of the host path (numpy in -> numpy out)

    # ------------------------------------------------------------------ weights
    def load_weights(self, weights: Union[str, Mapping[str, np.ndarray]]) -> "Generator":
        """weights: mapping (or .npz path) with '<layer>.v', '<layer>.g', '<layer>.b' per layer."""
        if isinstance(weights, str):
            weights = dict(np.load(weights))
        for L in self._layers:
            for suffix, shape in (("v", L.v_shape), ("g", L.g_shape), ("b", L.b_shape)):
                key = f"{L.name}.{suffix}"
                if key not in weights:
                    raise ValueError(f"missing weight {key!r}")
                arr = np.ascontiguousarray(np.asarray(weights[key], dtype=np.float32))
                if tuple(arr.shape) != tuple(shape):
                    raise ValueError(f"{key}: shape {tuple(arr.shape)} != expected {tuple(shape)}")
                shp = (C.c_int64 * arr.ndim)(*arr.shape)
                _lib.check(self._lib.hfg_set_weight(self._handle, key.encode(), arr.ctypes.data_as(C.c_void_p),
                                                    shp, arr.ndim), self._handle, key)
        _lib.check(self._lib.hfg_finalize_weights(self._handle, _lib.PRECISIONS[self.precision]),
                   self._handle, "hfg_finalize_weights")
        self._weights_loaded = True
        return self

    def init_random(self, seed: int = 1234, variant: str = "A") -> Dict[str, np.ndarray]:
        """Glorot-uniform v, g = ||v||, b = 0 (variant 'A': what Keras/tfa produce at random init,
        reference model.py:26-29 defaults) or perturbed g / non-zero b (variant 'B')."""
        w = synthetic_weights(self.hparams, variant, seed)
        self.load_weights(w)
        return w

    def set_option(self, key: str, value: int) -> None:
        _lib.check(self._lib.hfg_set_option(self._handle, key.encode(), int(value)), self._handle, key)
        if key in ("force_ffma", "auto_fold"):
            self._weights_loaded = False

    # ------------------------------------------------------------------ forward
    def __call__(self, x, training: bool = False):
        if training:
            raise NotImplementedError("this Generator is forward-only (inference); training=True is not supported")
        return self.call(x)

    def call(self, x):
        """mel [B, T, num_mels] fp32 -> audio [B, T*prod(rates), 1] fp32 (reference model.py:155-173).
        numpy in -> numpy out (host path, includes H2D/D2H); torch CUDA tensor in -> torch CUDA out
        (device path, asynchronous on the current stream); torch CPU tensor in -> torch CPU out."""
        if not self._weights_loaded:
            raise RuntimeError("weights not loaded: call load_weights() or init_random() first")
        if torch is not None and isinstance(x, torch.Tensor):
            if x.is_cuda:
                return self._call_device(x)
            return torch.from_numpy(self._call_host(x.detach().numpy()))
        if hasattr(x, "__cuda_array_interface__"):      # cupy / numba / any raw device buffer
            return self._call_device(self._from_cuda_array_interface(x))
        return self._call_host(np.asarray(x))

    def _from_cuda_array_interface(self, x):
        """Zero-copy torch view of a foreign device array (fp32, C-contiguous); the result of the call is a torch
        CUDA tensor, which exposes `__cuda_array_interface__` itself."""
        if torch is None:
            raise RuntimeError("device-pointer inputs need torch for the output / workspace allocation")
        cai = x.__cuda_array_interface__
        if cai.get("typestr") not in ("<f4", "=f4", "|f4") or cai.get("strides") is not None:
            raise ValueError("device mel must be float32 and C-contiguous")
        return torch.as_tensor(x, device=torch.device("cuda", self.device))

    def _check_shape(self, shape):
        if len(shape) != 3 or int(shape[2]) != int(self.hparams.num_mels):
            raise ValueError(f"mel must be [B, T, {self.hparams.num_mels}] (channels-last), got {tuple(shape)}")

    def _call_host(self, mel: np.ndarray, out: Optional[np.ndarray] = None) -> np.ndarray:
        self._check_shape(mel.shape)
        mel = np.ascontiguousarray(mel, dtype=np.float32)
        B, T = int(mel.shape[0]), int(mel.shape[1])
        if out is None:
            # the result is an ordinary numpy array owned by the caller; its memory is page-locked and recycled through
            # the generator's pool when the array is dropped, so the device->host copy runs asynchronously at link speed
            out = self._pinned.array((B, T * self._total_up, 1), np.float32)
        _lib.check(self._lib.hfg_forward_host(self._handle, mel.ctypes.data_as(C.c_void_p), B, T,
                                              out.ctypes.data_as(C.c_void_p)), self._handle, "hfg_forward_host")
        return out

    # ------------------------------------------------------------------ ragged batches
    def call_ragged(self, mels, pcm16: bool = False):
        """Utterances of different lengths: `mels` is a sequence of [T_i, num_mels] arrays (numpy, or torch CUDA
        tensors on this generator's device).  Equal-length utterances ride one dense batch (longest first); nothing is
        padded, so every waveform is bit-identical to the one `call(mel[None])` gives (utterances are independent,
        reference model.py:155-173).  Returns a list, in input order, of [T_i*prod(rates), 1] fp32 waveforms (or
        [T_i*prod(rates)] int16 PCM with pcm16=True), of the same container kind as the input."""
        from .sharding import group_by_length
        mels = list(mels)
        for m in mels:
            if len(m.shape) != 2 or int(m.shape[1]) != int(self.hparams.num_mels):
                raise ValueError(f"each mel must be [T, {self.hparams.num_mels}] (channels-last), got {tuple(m.shape)}")
        on_device = bool(mels) and torch is not None and isinstance(mels[0], torch.Tensor) and mels[0].is_cuda
        out = [None] * len(mels)
        for group in group_by_length([int(m.shape[0]) for m in mels], range(len(mels))):
            if on_device:
                batch = torch.stack([mels[i].float() for i in group])
            else:
                batch = np.stack([np.asarray(mels[i].detach().numpy() if torch is not None and isinstance(mels[i], torch.Tensor)
                                             else mels[i], dtype=np.float32) for i in group])
            wav = self.synthesize_pcm16(batch) if pcm16 else self.call(batch)
            for j, i in enumerate(group):
                out[i] = wav[j]
        return out

    # ------------------------------------------------------------------ 16-bit PCM output
    def synthesize_pcm16(self, x):
        """mel [B, T, num_mels] fp32 -> int16 PCM [B, T*prod(rates)]: `(audio.clamp(-1, 1) * 32768).astype(int16)`
        (the step the reference intends after the vocoder, Grad-TTS_TF/inference.py:89), fused into the conv_post
        epilogue on the GPU; +1.0 saturates to 32767.  numpy in -> numpy out, torch CUDA in -> torch CUDA out."""
        if not self._weights_loaded:
            raise RuntimeError("weights not loaded: call load_weights() or init_random() first")
        if torch is not None and isinstance(x, torch.Tensor) and x.is_cuda:
            mel = x.float().contiguous()
            self._check_shape(mel.shape)
            if mel.device.index != self.device:
                raise ValueError(f"mel is on cuda:{mel.device.index}, generator on cuda:{self.device}")
            B, T = int(mel.shape[0]), int(mel.shape[1])
            out = torch.empty((B, T * self._total_up), dtype=torch.int16, device=mel.device)
            if B == 0 or T == 0:
                return out
            ws = self._workspace(B, T, mel.device)
            stream = torch.cuda.current_stream(mel.device).cuda_stream
            _lib.check(self._lib.hfg_forward_pcm16(self._handle, C.c_void_p(mel.data_ptr()), B, T, C.c_void_p(out.data_ptr()),
                                                   C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(stream)),
                       self._handle, "hfg_forward_pcm16")
            return out
        as_torch = torch is not None and isinstance(x, torch.Tensor)
        mel = x.detach().numpy() if as_torch else np.asarray(x)
        self._check_shape(mel.shape)
        mel = np.ascontiguousarray(mel, dtype=np.float32)
        B, T = int(mel.shape[0]), int(mel.shape[1])
        out = self._pinned.array((B, T * self._total_up), np.int16)
        _lib.check(self._lib.hfg_forward_host_pcm16(self._handle, mel.ctypes.data_as(C.c_void_p), B, T,
                                                    out.ctypes.data_as(C.c_void_p)), self._handle, "hfg_forward_host_pcm16")
        return torch.from_numpy(out) if as_torch else out

    def forward_host_pcm16_ptr(self, mel_ptr: int, B: int, T: int, pcm_ptr: int) -> None:
        """Raw host-pointer entry of the PCM path (pinned buffers)."""
        _lib.check(self._lib.hfg_forward_host_pcm16(self._handle, C.c_void_p(mel_ptr), B, T, C.c_void_p(pcm_ptr)),
                   self._handle, "hfg_forward_host_pcm16")

    def _workspace(self, B: int, T: int, device):
        need = self.workspace_bytes(B, T)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=device)
        return self._ws

    def forward_host_ptr(self, mel_ptr: int, B: int, T: int, audio_ptr: int) -> None:
        """Raw host-pointer entry (pinned buffers from hfg_host_alloc / torch pin_memory)."""
        _lib.check(self._lib.hfg_forward_host(self._handle, C.c_void_p(mel_ptr), B, T, C.c_void_p(audio_ptr)),
                   self._handle, "hfg_forward_host")

    def workspace_bytes(self, B: int, T: int) -> int:
        n = C.c_size_t()
        _lib.check(self._lib.hfg_workspace_bytes(self._handle, B, T, C.byref(n)), self._handle)
        return int(n.value)

    def _call_device(self, mel, out=None):
        self._check_shape(mel.shape)
        if mel.dtype != torch.float32:
            mel = mel.float()
        mel = mel.contiguous()
        if mel.device.index != self.device:
            raise ValueError(f"mel is on cuda:{mel.device.index}, generator on cuda:{self.device}")
        B, T = int(mel.shape[0]), int(mel.shape[1])
        if out is None:
            out = torch.empty((B, T * self._total_up, 1), dtype=torch.float32, device=mel.device)
        if B == 0 or T == 0:
            return out
        self._workspace(B, T, mel.device)
        stream = torch.cuda.current_stream(mel.device).cuda_stream
        _lib.check(self._lib.hfg_forward(self._handle, C.c_void_p(mel.data_ptr()), B, T, C.c_void_p(out.data_ptr()),
                                         C.c_void_p(self._ws.data_ptr()), self._ws.numel(), C.c_void_p(stream)),
                   self._handle, "hfg_forward")
        return out

    # ------------------------------------------------------------------ introspection
    def last_launch_count(self) -> int:
        n = C.c_longlong()
        _lib.check(self._lib.hfg_last_launch_count(self._handle, C.byref(n)), self._handle)
        return int(n.value)

    def layer_engines(self) -> Dict[str, str]:
        """layer name -> "tc" (persistent tcgen05 kernel) | "ffma" (direct kernel) | "post"."""
        buf = C.create_string_buffer(1 << 14)
        _lib.check(self._lib.hfg_layer_engines(self._handle, buf, len(buf)), self._handle)
        return dict(line.split(":") for line in buf.value.decode().strip().split("\n"))

    def auto_micro_batch(self, B: int, T: int, host_path: bool = False) -> int:
        n = C.c_int()
        _lib.check(self._lib.hfg_auto_micro_batch(self._handle, int(B), int(T), int(host_path), C.byref(n)), self._handle)
        return int(n.value)

    def layer_folds(self) -> Dict[str, tuple]:
        """layer name -> (fold_in, fold_out, kernel-view taps) chosen for the tcgen05 kernel."""
        buf = C.create_string_buffer(1 << 14)
        _lib.check(self._lib.hfg_layer_folds(self._handle, buf, len(buf)), self._handle)
        return {p[0]: tuple(int(v) for v in p[1:]) for p in (line.split(":") for line in buf.value.decode().strip().split("\n"))}

    def stage_times_ms(self):
        buf = (C.c_float * 16)()
        n = C.c_int(16)
        _lib.check(self._lib.hfg_stage_times_ms(self._handle, buf, C.byref(n)), self._handle)
        return [float(buf[i]) for i in range(n.value)]

    def stage_launches(self):
        """Kernels launched per measurement stage ([conv_pre, stage 0.., conv_post]) by the last forward."""
        buf = (C.c_int * 16)()
        n = C.c_int(16)
        _lib.check(self._lib.hfg_stage_launches(self._handle, buf, C.byref(n)), self._handle)
        return [int(buf[i]) for i in range(n.value)]

    def summary(self, print_fn=print) -> None:
        """Layer table (the Keras `.summary()` surface touched by reference gan.py:192)."""
        print_fn(f"Generator(resblock={self.hparams.resblock!r}, precision={self.precision}, compat={self.compat})")
        total = 0
        for L in self._layers:
            n = L.kernel * L.c_in * L.c_out + L.c_out
            total += n
            print_fn(f"  {L.name:28s} {L.kind:5s} k={L.kernel:<2d} d={L.dilation:<2d} s={L.stride:<2d} "
                     f"{L.c_in:>4d}->{L.c_out:<4d} pad={L.padding:6s} params={n}")
        print_fn(f"  total effective params: {total}")

    def close(self) -> None:
        if getattr(self, "_handle", None) is not None and self._handle:
            self._lib.hfg_destroy(self._handle)
            self._handle = C.c_void_p()
        self._ws = None
        if getattr(self, "_pinned", None) is not None:
            self._pinned.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
