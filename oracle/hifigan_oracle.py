"""CPU ORACLE for the HiFi-GAN generator forward pass -- TEST INFRASTRUCTURE ONLY.

This file is a checker, never a product path: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import it.

It restates, operation by operation, what the reference's TF2/Keras generator
computes:
  * Generator.call                      reference HiFiGAN_TF/model.py:155-173
  * ResBlock1.call / ResBlock2.call     reference HiFiGAN_TF/model.py:71-80, 106-112
  * get_padding -> "causal"/"same"      reference HiFiGAN_TF/utils.py:29-37
  * LRELU_SLOPE = 0.1, LeakyReLU() 0.3  reference HiFiGAN_TF/model.py:14, 152

PARITY UNPINNED at the TensorFlow boundary: the arithmetic of the reference lives
in third-party libraries that are not vendored under /root/reference and cannot be
installed here (tensorflow / keras ~2.7, tensorflow-addons ~0.15; versions are not
pinned by the reference).  Their published semantics are restated below:
  * keras.layers.Conv1D: NWC input, kernel [k, Cin, Cout], cross-correlation
    y[t] = b + sum_j x[t*1 + j*d - pad_left] W[j]; "same" pads d(k-1) split
    left = total//2; "causal" pads d(k-1) on the left only.
  * keras.layers.Conv1DTranspose(padding="same", strides=u): kernel [k, Cout, Cin],
    full[i*u + j] += x[i] W[j]; output = full[pad_left : pad_left + L*u] with
    pad_left = (k-u)//2  (gradient of a SAME strided conv).
  * tfa.layers.WeightNormalization(data_init=False): kernel =
    v * rsqrt(max(sum_{all axes but last} v^2, 1e-12)) * g, g indexed by the
    kernel's LAST axis (Cout for Conv1D, Cin for Conv1DTranspose).
What IS pinned: with the reference quirks switched off (compat="upstream") this
restatement is checked against the importable upstream PyTorch generator that the
reference vendors as a submodule (FastSpeech2_TF/FastSpeech2/hifigan/models.py:
112-165) -- see tests/golden/make_golden.py and tests/test_oracle.py.

Two back ends share one code path:
  * numpy (float64 or float32) -- ground truth
  * torch CPU float32 (F.conv1d) -- the timed CPU baseline
"""
from __future__ import annotations

import numpy as np

LRELU_SLOPE = 0.1
FINAL_SLOPE_REFERENCE = 0.3   # keras LeakyReLU() default (reference quirk Q2)
FINAL_SLOPE_UPSTREAM = 0.01   # torch F.leaky_relu default (upstream jik876/hifi-gan)
L2_EPS = 1e-12


# ----------------------------------------------------------------------------
# padding rules
# ----------------------------------------------------------------------------
def padding_mode(kernel_size, dilation, compat="reference"):
    """reference HiFiGAN_TF/utils.py:29-37 (int pad -> keras mode string)."""
    if compat == "upstream":
        return "same"
    pad = int((kernel_size * dilation - dilation) / 2)
    if pad == 0:
        return "valid"
    if pad == 1:
        return "causal"
    return "same"


def pads_for(kernel_size, dilation, mode):
    span = dilation * (kernel_size - 1)
    if mode == "valid":
        return 0, 0
    if mode == "causal":
        return span, 0
    return span // 2, span - span // 2


# ----------------------------------------------------------------------------
# weight norm fold (tfa.layers.WeightNormalization, data_init=False)
# ----------------------------------------------------------------------------
def fold_weight_norm(v, g, dtype=np.float64):
    v = np.asarray(v, dtype=dtype)
    g = np.asarray(g, dtype=dtype)
    sq = np.sum(v * v, axis=(0, 1), keepdims=True)
    return v / np.sqrt(np.maximum(sq, dtype(L2_EPS))) * g.reshape(1, 1, -1)


# ----------------------------------------------------------------------------
# numpy primitives (channels-last)
# ----------------------------------------------------------------------------
def lrelu(x, slope):
    return np.where(x >= 0, x, x * x.dtype.type(slope))


def conv1d_ntc(x, w, b, dilation, mode):
    """x [B,L,Cin], w [k,Cin,Cout] -> [B,L',Cout]."""
    k = w.shape[0]
    pl, pr = pads_for(k, dilation, mode)
    B, L, _ = x.shape
    xp = np.pad(x, ((0, 0), (pl, pr), (0, 0)))
    Lo = L + pl + pr - dilation * (k - 1)
    y = np.zeros((B, Lo, w.shape[2]), dtype=x.dtype)
    for j in range(k):
        y += xp[:, j * dilation: j * dilation + Lo, :] @ w[j]
    return y + b.reshape(1, 1, -1)


def conv1d_transpose_ntc(x, w, b, stride):
    """x [B,L,Cin], w [k,Cout,Cin] (keras), 'same' -> [B,L*stride,Cout]."""
    k = w.shape[0]
    B, L, _ = x.shape
    full = np.zeros((B, (L - 1) * stride + k, w.shape[1]), dtype=x.dtype)
    for j in range(k):
        full[:, j: j + (L - 1) * stride + 1: stride, :] += x @ w[j].T
    pl = (k - stride) // 2
    return full[:, pl: pl + L * stride, :] + b.reshape(1, 1, -1)


# ----------------------------------------------------------------------------
# torch primitives (float32 CPU) -- same maths through F.conv1d
# ----------------------------------------------------------------------------
class _TorchOps:
    def __init__(self):
        import torch
        import torch.nn.functional as F
        self.torch, self.F = torch, F

    def prep(self, x):
        return self.torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).permute(0, 2, 1).contiguous()

    def done(self, x):
        return x.permute(0, 2, 1).contiguous().numpy()

    def weight(self, w, kind):
        t = self.torch.from_numpy(np.ascontiguousarray(w, dtype=np.float32))
        if kind == "conv":      # [k,Cin,Cout] -> torch [Cout,Cin,k]
            return t.permute(2, 1, 0).contiguous()
        return t.permute(2, 1, 0).contiguous()   # [k,Cout,Cin] -> torch convT [Cin,Cout,k]

    def lrelu(self, x, slope):
        return self.F.leaky_relu(x, slope)

    def conv(self, x, w, b, dilation, mode):
        k = w.shape[2]
        pl, pr = pads_for(k, dilation, mode)
        return self.F.conv1d(self.F.pad(x, (pl, pr)), w, b, dilation=dilation)

    def convT(self, x, w, b, stride):
        k = w.shape[2]
        L = x.shape[2]
        full = self.F.conv_transpose1d(x, w, None, stride=stride)
        pl = (k - stride) // 2
        return full[:, :, pl: pl + L * stride] + b.view(1, -1, 1)

    def tanh(self, x):
        return self.torch.tanh(x)


class _NumpyOps:
    def __init__(self, dtype):
        self.dtype = dtype

    def prep(self, x):
        return np.asarray(x, dtype=self.dtype)

    def done(self, x):
        return x

    def weight(self, w, kind):
        return np.asarray(w, dtype=self.dtype)

    def lrelu(self, x, slope):
        return lrelu(x, slope)

    def conv(self, x, w, b, dilation, mode):
        return conv1d_ntc(x, w, b, dilation, mode)

    def convT(self, x, w, b, stride):
        return conv1d_transpose_ntc(x, w, b, stride)

    def tanh(self, x):
        return np.tanh(x)


# ----------------------------------------------------------------------------
# bf16-storage emulation (torch only) -- used by tests/ to budget the error of
# the tensor-core path before it ever runs on a GPU; NOT a product path.
# ----------------------------------------------------------------------------
class _TorchBf16Ops(_TorchOps):
    """Same maths as _TorchOps, but conv operands (activations after the
    pre-activation, and weights) are rounded to bf16 and every tensor that the
    CUDA path stores between kernels is rounded through `store()`.  Accumulation
    stays fp32, like tcgen05 kind::f16 with an f32 accumulator."""

    def __init__(self, trunk="bf16", operand="bf16"):
        super().__init__()
        self.trunk = trunk
        self.operand = operand

    def _r(self, t, kind=None):
        kind = kind or self.operand
        if kind == "tf32":      # round-to-nearest-even to a 10-bit mantissa
            i = t.contiguous().view(self.torch.int32)
            i = (i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF
            return i.view(self.torch.float32)
        dt = self.torch.bfloat16 if kind == "bf16" else self.torch.float16
        return t.to(dt).to(self.torch.float32)

    def weight(self, w, kind):
        return self._r(super().weight(w, kind))

    def conv(self, x, w, b, dilation, mode):
        return super().conv(self._r(x), w, b, dilation, mode)

    def convT(self, x, w, b, stride):
        return super().convT(self._r(x), w, b, stride)

    def store(self, x):
        return x if self.trunk == "f32" else self._r(x, self.trunk)


def _make_ops(backend):
    if backend == "numpy64":
        return _NumpyOps(np.float64)
    if backend == "numpy32":
        return _NumpyOps(np.float32)
    if backend == "torch32":
        return _TorchOps()
    if backend.startswith("torch_emul:"):       # e.g. "torch_emul:fp16:f32" = operand:trunk
        _, operand, trunk = backend.split(":")
        return _TorchBf16Ops(trunk, operand)
    raise ValueError(f"unknown oracle backend {backend!r}")


# ----------------------------------------------------------------------------
# topology walk (reference HiFiGAN_TF/model.py:116-152)
# ----------------------------------------------------------------------------
def _get(h, name):
    return h[name] if isinstance(h, dict) else getattr(h, name)


def used_dilations(h):
    """The dilation entries the reference's resblocks actually consume: ResBlock1 builds convs from dilation[0..2]
    (reference HiFiGAN_TF/model.py:21-66), ResBlock2 from dilation[0..1] (model.py:84-101); longer lists are ignored."""
    need = 3 if _get(h, "resblock") == "1" else 2
    return [[int(d) for d in list(ds)[:need]] for ds in _get(h, "resblock_dilation_sizes")]


def layer_names(h):
    """Weight-bearing layers in construction/call order -> (name, kind, k, dil, stride, cin, cout)."""
    c0 = int(_get(h, "upsample_initial_channel"))
    num_mels = int(h.get("num_mels", 80)) if isinstance(h, dict) else int(getattr(h, "num_mels", 80))
    rks = list(_get(h, "resblock_kernel_sizes"))
    rds = used_dilations(h)
    rb1 = _get(h, "resblock") == "1"
    out = [("conv_pre", "conv", 7, 1, 1, num_mels, c0)]
    cprev = c0
    for i, (u, ku) in enumerate(zip(_get(h, "upsample_rates"), _get(h, "upsample_kernel_sizes"))):
        c = c0 // (2 ** (i + 1))
        out.append((f"ups.{i}", "convT", int(ku), 1, int(u), cprev, c))
        for j, (k, dils) in enumerate(zip(rks, rds)):
            n = i * len(rks) + j
            for m, d in enumerate(dils):
                if rb1:
                    out.append((f"resblocks.{n}.convs1.{m}", "conv", int(k), int(d), 1, c, c))
                    out.append((f"resblocks.{n}.convs2.{m}", "conv", int(k), 1, 1, c, c))
                else:
                    out.append((f"resblocks.{n}.convs.{m}", "conv", int(k), int(d), 1, c, c))
        cprev = c
    out.append(("conv_post", "conv", 7, 1, 1, cprev, 1))
    return out


def fold_all(h, weights, dtype=np.float64):
    """{name.v,name.g,name.b} -> {name: (kernel, bias)} with weight norm folded."""
    folded = {}
    for (name, kind, k, d, s, cin, cout) in layer_names(h):
        v, g, b = weights[f"{name}.v"], weights[f"{name}.g"], weights[f"{name}.b"]
        exp_v = (k, cin, cout) if kind == "conv" else (k, cout, cin)
        if tuple(v.shape) != exp_v:
            raise ValueError(f"{name}.v has shape {tuple(v.shape)}, expected {exp_v}")
        folded[name] = (fold_weight_norm(v, g, dtype), np.asarray(b, dtype=dtype))
    return folded


# ----------------------------------------------------------------------------
# Generator.call (reference HiFiGAN_TF/model.py:155-173)
# ----------------------------------------------------------------------------
def generator_forward(h, weights, mel, backend="numpy64", compat="reference", taps=None):
    """mel [B,T,num_mels] -> audio [B, T*prod(rates), 1] (channels-last, like the reference).

    compat="reference": causal (k=3,d=1) convs + final slope 0.3 (reference quirks Q1/Q2).
    compat="upstream" : symmetric padding everywhere + final slope 0.01 (jik876/hifi-gan),
                        used only to pin this restatement against the vendored PyTorch model.
    taps: optional dict; if given, intermediate tensors are recorded (numpy NTC) under
          "conv_pre", "ups.i", "stage.i".
    """
    ops = _make_ops(backend)
    fold_dtype = np.float64 if backend == "numpy64" else np.float32
    folded = fold_all(h, weights, fold_dtype)
    store = getattr(ops, "store", lambda t: t)

    def W(name, kind):
        w, b = folded[name]
        wt = ops.weight(w, kind)
        if backend.startswith("torch"):
            bt = ops.torch.from_numpy(np.ascontiguousarray(b, dtype=np.float32))
        else:
            bt = np.asarray(b, dtype=ops.dtype)
        return wt, bt

    rks = list(_get(h, "resblock_kernel_sizes"))
    rds = used_dilations(h)
    rates = list(_get(h, "upsample_rates"))
    rb1 = _get(h, "resblock") == "1"
    nk = len(rks)
    final_slope = FINAL_SLOPE_REFERENCE if compat == "reference" else FINAL_SLOPE_UPSTREAM

    def record(key, t):
        if taps is not None:
            taps[key] = np.array(ops.done(t))

    x = ops.prep(mel)
    w, b = W("conv_pre", "conv")
    x = store(ops.conv(x, w, b, 1, "same"))                       # model.py:156
    record("conv_pre", x)
    for i, u in enumerate(rates):                                 # model.py:157
        x = ops.lrelu(x, LRELU_SLOPE)                             # model.py:159
        w, b = W(f"ups.{i}", "convT")
        x = store(ops.convT(x, w, b, int(u)))                     # model.py:160
        record(f"ups.{i}", x)
        xs = None
        for j in range(nk):                                       # model.py:162-166
            n = i * nk + j
            k = int(rks[j])
            r = x
            for m, d in enumerate(rds[j]):
                d = int(d)
                if rb1:                                           # ResBlock1.call model.py:71-80
                    w1, b1 = W(f"resblocks.{n}.convs1.{m}", "conv")
                    w2, b2 = W(f"resblocks.{n}.convs2.{m}", "conv")
                    xt = ops.lrelu(r, LRELU_SLOPE)
                    xt = store(ops.conv(xt, w1, b1, d, padding_mode(k, d, compat)))
                    xt = ops.lrelu(xt, LRELU_SLOPE)
                    xt = ops.conv(xt, w2, b2, 1, padding_mode(k, 1, compat))
                    r = store(xt + r)
                else:                                             # ResBlock2.call model.py:106-112
                    w1, b1 = W(f"resblocks.{n}.convs.{m}", "conv")
                    xt = ops.lrelu(r, LRELU_SLOPE)
                    xt = ops.conv(xt, w1, b1, d, padding_mode(k, d, compat))
                    r = store(xt + r)
            xs = r if xs is None else xs + r
        x = store(xs / nk)                                        # model.py:167
        record(f"stage.{i}", x)
    x = ops.lrelu(x, final_slope)                                 # model.py:169
    w, b = W("conv_post", "conv")
    x = ops.conv(x, w, b, 1, "same")                              # model.py:170
    x = ops.tanh(x)                                               # model.py:171
    return np.array(ops.done(x))
