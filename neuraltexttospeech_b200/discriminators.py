"""Host-side mirror of the reference's discriminators for the FORWARD pass:
`MultiPeriodDiscriminator` (HiFiGAN_TF/model.py:246-274) and `MultiScaleDiscriminator` (model.py:349-389), same call
convention -- `y_d_rs, y_d_gs, fmap_rs, fmap_gs = mpd((y, y_hat))` -- over the hfg_disc_* C ABI (csrc/disc.cu).
No backward pass: these are for scoring / feature extraction, not for training.  There is no CPU fallback."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import numpy as np

from . import _lib

MPD_PERIODS = (2, 3, 5, 7, 11)          # model.py:250-256


class _Discriminator:
    KIND = 0
    N_SUB = 0
    N_LAYERS = 0                         # convs + conv_post per sub-discriminator

    def __init__(self, compat: str = "reference", device: int = 0):
        if compat not in ("reference", "upstream"):
            raise ValueError("compat must be 'reference' or 'upstream'")
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.device = int(device)
        _lib.check(self._lib.hfg_disc_create(self.KIND, 0 if compat == "reference" else 1, self.device, C.byref(self._h)),
                   None, "hfg_disc_create", disc=True)
        n = C.c_int()
        self._check(self._lib.hfg_disc_num_maps(self._h, C.byref(n)), "hfg_disc_num_maps")
        self.num_maps = n.value

    def _check(self, rc, what):
        _lib.check(rc, self._h, what, disc=True)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hfg_disc_destroy(h)

    def layer_names(self) -> List[str]:
        out = []
        for i in range(self.N_SUB):
            out += [f"discriminators.{i}.convs.{l}" for l in range(self.N_LAYERS - 1)] + [f"discriminators.{i}.conv_post"]
        return out

    def load_weights(self, weights: Dict[str, np.ndarray]):
        """`weights`: "<layer>.v/.g/.b" (tfa WeightNormalization's variables) or "<layer>.w/.b" (a kernel used as stored,
        e.g. the SpectralNormalization-wrapped first DiscriminatorS at inference) for every name of layer_names()."""
        for name in self.layer_names():
            parts = ("w", "b") if f"{name}.w" in weights else ("v", "g", "b")
            for part in parts:
                key = f"{name}.{part}"
                if key not in weights:
                    raise KeyError(f"missing weight {key}")
                a = np.ascontiguousarray(weights[key], dtype=np.float32)
                shape = (C.c_int64 * a.ndim)(*a.shape)
                self._check(self._lib.hfg_disc_set_weight(self._h, key.encode(), a.ctypes.data_as(C.c_void_p), shape, a.ndim), key)
        self._check(self._lib.hfg_disc_finalize_weights(self._h), "hfg_disc_finalize_weights")
        return self

    def layout(self, B: int, T: int):
        offs = (C.c_longlong * self.num_maps)()
        dims = (C.c_longlong * (4 * self.num_maps))()
        total = C.c_longlong()
        self._check(self._lib.hfg_disc_layout(self._h, int(B), int(T), offs, dims, C.byref(total)), "hfg_disc_layout")
        return list(offs), [tuple(dims[4 * i:4 * i + 4]) for i in range(self.num_maps)], total.value

    def _views(self, buf, B, T):
        """Split the flat output into the reference's per-sub-discriminator (score, fmap list); DiscriminatorP maps come
        back as [B, len, period, C] views (model.py:235), DiscriminatorS maps as [B, len, C]."""
        offs, dims, _ = self.layout(B, T)
        scores, fmaps = [], []
        for i in range(self.N_SUB):
            fm = []
            for l in range(self.N_LAYERS):
                o, (b, p, n, c) = offs[i * self.N_LAYERS + l], dims[i * self.N_LAYERS + l]
                m = buf[o:o + b * p * n * c].reshape(b, p, n, c)
                fm.append(m[:, 0] if self.KIND == 1 else (m.transpose(0, 2, 1, 3) if isinstance(m, np.ndarray) else m.permute(0, 2, 1, 3)))
            fmaps.append(fm)
            scores.append(fm[-1].reshape(B, -1))
        return scores, fmaps

    def forward(self, y):
        """One waveform batch: numpy / torch CUDA [B, T] or [B, T, 1] -> (scores, fmaps) lists over the sub-discriminators."""
        is_np = isinstance(y, np.ndarray)
        if y.ndim == 3:
            if y.shape[2] != 1:
                raise ValueError("expected [B, T, 1] audio")
            y = y[:, :, 0]
        if y.ndim != 2:
            raise ValueError("expected [B, T] or [B, T, 1] audio")
        B, T = (int(v) for v in y.shape)
        _, _, total = self.layout(B, T)
        if is_np:
            a = np.ascontiguousarray(y, dtype=np.float32)
            out = np.empty(total, np.float32)
            self._check(self._lib.hfg_disc_forward_host(self._h, a.ctypes.data_as(C.c_void_p), B, T, out.ctypes.data_as(C.c_void_p)),
                        "hfg_disc_forward_host")
            return self._views(out, B, T)
        import torch
        if not y.is_cuda or y.device.index != self.device:
            raise ValueError(f"expected a CUDA tensor on device {self.device}")
        a = y.contiguous().float()
        out = torch.empty(total, dtype=torch.float32, device=a.device)
        self._check(self._lib.hfg_disc_forward(self._h, C.c_void_p(a.data_ptr()), B, T, C.c_void_p(out.data_ptr()),
                                               C.c_void_p(torch.cuda.current_stream(a.device).cuda_stream)), "hfg_disc_forward")
        return self._views(out, B, T)

    def call(self, x):
        """The reference's signature (model.py:259-274 / 367-389): x = (y, y_hat) -> y_d_rs, y_d_gs, fmap_rs, fmap_gs.
        Both inputs go through ONE forward over the concatenated batch (no op couples utterances)."""
        y, y_hat = x
        if tuple(y.shape) != tuple(y_hat.shape):
            raise ValueError("y and y_hat must have the same shape")
        B = int(y.shape[0])
        both = np.concatenate([y, y_hat], axis=0) if isinstance(y, np.ndarray) else __import__("torch").cat([y, y_hat], dim=0)
        scores, fmaps = self.forward(both)
        return ([s[:B] for s in scores], [s[B:] for s in scores],
                [[f[:B] for f in fm] for fm in fmaps], [[f[B:] for f in fm] for fm in fmaps])

    __call__ = call

    def pair_losses(self, x):
        """x = (y, y_hat) -> the reference's loss terms for this discriminator (HiFiGAN_TF/losses.py:8-43, as gan.py:58-92
        combines them), reduced on the GPU from one forward over the concatenated batch: dict with feature_loss,
        discriminator_loss, generator_loss (floats) and r_losses, g_losses, gen_losses (one float per sub-discriminator)."""
        import torch
        y, y_hat = x
        if tuple(y.shape) != tuple(y_hat.shape):
            raise ValueError("y and y_hat must have the same shape")
        to_t = lambda a: (torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda(self.device) if isinstance(a, np.ndarray) else a)
        both = torch.cat([to_t(y), to_t(y_hat)], dim=0)
        if both.ndim == 3:
            both = both[:, :, 0]
        both = both.contiguous().float()
        B2, T = (int(v) for v in both.shape)
        _, _, total = self.layout(B2, T)
        maps = torch.empty(total, dtype=torch.float32, device=both.device)
        stream = C.c_void_p(torch.cuda.current_stream(both.device).cuda_stream)
        self._check(self._lib.hfg_disc_forward(self._h, C.c_void_p(both.data_ptr()), B2, T, C.c_void_p(maps.data_ptr()), stream), "hfg_disc_forward")
        out = (C.c_float * (3 + 3 * self.N_SUB))()
        self._check(self._lib.hfg_disc_pair_losses(self._h, C.c_void_p(maps.data_ptr()), B2 // 2, T, out, stream), "hfg_disc_pair_losses")
        n = self.N_SUB
        return {"feature_loss": out[0], "discriminator_loss": out[1], "generator_loss": out[2],
                "r_losses": list(out[3:3 + n]), "g_losses": list(out[3 + n:3 + 2 * n]), "gen_losses": list(out[3 + 2 * n:3 + 3 * n])}


class MultiPeriodDiscriminator(_Discriminator):
    KIND, N_SUB, N_LAYERS = 0, 5, 6


class MultiScaleDiscriminator(_Discriminator):
    KIND, N_SUB, N_LAYERS = 1, 3, 8


# (c_in, c_out, kernel, groups) per layer, in forward order (model.py:186-216 / 284-332)
_P_LAYERS = ((1, 32, 5, 1), (32, 128, 5, 1), (128, 512, 5, 1), (512, 1024, 5, 1), (1024, 1024, 5, 1), (1024, 1, 3, 1))
_S_LAYERS = ((1, 128, 15, 1), (128, 128, 41, 4), (128, 256, 41, 16), (256, 512, 41, 16), (512, 1024, 41, 16), (1024, 1024, 41, 16),
             (1024, 1024, 5, 1), (1024, 1, 3, 1))


def disc_flops(model: _Discriminator, B: int, T: int):
    """Algorithmic work of one forward over [B, T]: (FLOPs = 2 * MACs of every conv, bytes of all feature maps written)."""
    _, dims, total = model.layout(B, T)
    layers = _P_LAYERS if model.KIND == 0 else _S_LAYERS
    flops = 0
    for i, (b, p, n, c) in enumerate(dims):
        cin, cout, k, groups = layers[i % model.N_LAYERS]
        flops += 2 * b * p * n * cout * k * (cin // groups)
    return flops, total * 4
