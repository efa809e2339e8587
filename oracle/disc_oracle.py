"""CPU restatement of the reference's HiFi-GAN DISCRIMINATOR forward passes -- TEST INFRASTRUCTURE ONLY.

SURVEY.md section 8f next-4: `DiscriminatorP`, `MultiPeriodDiscriminator`, `DiscriminatorS`, `MultiScaleDiscriminator`
(reference HiFiGAN_TF/model.py:176-243, 246-274, 277-346, 349-389).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg may import this file; the product path (neuraltexttospeech_b200/discriminators.py -> hfg_disc_* in
csrc/disc.cu) never does.

PARITY PINNING.  The reference's arithmetic lives in TensorFlow / Keras / tensorflow-addons, which no box here carries,
and the reference holds no golden vectors for these models: in compat="reference" this oracle is "parity unpinned" at the
TensorFlow boundary, exactly like oracle/hifigan_oracle.py.  What IS pinned (tests/golden/make_golden_disc.py ->
tests/golden/disc_upstream_{mpd,msd}.npz, tests/test_disc_oracle.py): compat="upstream" reproduces the torch
discriminators the reference vendors (FastPitch_TF/DeepLearningExamples/PyTorch/SpeechSynthesis/HiFi-GAN/hifigan/
models.py:273-409) run in the build container on the same weights -- layer list, grouped / strided conv placement,
period reshape, reflect padding, weight-norm fold axes, pooling windows.  The differences between the two modes are the
Keras semantics of the reference file, restated below from the TensorFlow documentation:
  * padding="same" with a stride (model.py:188-210, 292-321): out = ceil(in / s), total = max((out-1)*s + k - in, 0),
    left = total // 2, right = total - left (TensorFlow's SAME rule; torch pads (k-1)/2 on both sides);
  * the first DiscriminatorS conv is padding="valid" (model.py:286-289; upstream pads 7);
  * AveragePooling1D(4, 2, "same") (model.py:357-364) divides every window by its number of IN-RANGE samples (TensorFlow
    average pooling never counts padding); NVIDIA's torch copy uses AvgPool1d(4, 2, padding=1), which divides by 4;
  * SpectralNormalization (first DiscriminatorS, model.py:353) only re-normalises while training: the forward pass
    convolves with the stored kernel `w`.
Layout: audio [B, T] (the reference's [B, T, 1]); every feature map is returned channels-last like the reference:
DiscriminatorP [B, H, period, C], DiscriminatorS [B, L, C]; the score is the flattened last map.
"""
from __future__ import annotations

import numpy as np

LRELU_SLOPE = 0.1                       # HiFiGAN_TF/model.py:14
MPD_PERIODS = (2, 3, 5, 7, 11)          # model.py:250-256

# (c_in, c_out, kernel, stride, groups, padding)
P_CONVS = [(1, 32, 5, 3, 1, "same"), (32, 128, 5, 3, 1, "same"), (128, 512, 5, 3, 1, "same"),
           (512, 1024, 5, 3, 1, "same"), (1024, 1024, 5, 1, 1, "same")]                        # model.py:186-212
P_POST = (1024, 1, 3, 1, 1, "same")                                                             # model.py:213-216
S_CONVS = [(1, 128, 15, 1, 1, "valid"), (128, 128, 41, 2, 4, "same"), (128, 256, 41, 2, 16, "same"),
           (256, 512, 41, 4, 16, "same"), (512, 1024, 41, 4, 16, "same"), (1024, 1024, 41, 1, 16, "same"),
           (1024, 1024, 5, 1, 1, "same")]                                                       # model.py:284-327
S_POST = (1024, 1, 3, 1, 1, "same")                                                             # model.py:328-332


def layer_specs(kind):
    """-> list of (name, spec) in forward order for every sub-discriminator of `kind` ("mpd" / "msd")."""
    n = len(MPD_PERIODS) if kind == "mpd" else 3
    convs, post = (P_CONVS, P_POST) if kind == "mpd" else (S_CONVS, S_POST)
    out = []
    for i in range(n):
        for l, spec in enumerate(convs):
            out.append((f"discriminators.{i}.convs.{l}", spec))
        out.append((f"discriminators.{i}.conv_post", post))
    return out


def kernel_shape(spec):
    """Keras kernel with the unit width axis of Conv2D((k, 1)) squeezed: [k, c_in / groups, c_out]."""
    cin, cout, k, _, groups, _ = spec
    return (k, cin // groups, cout)


def fold(weights, name, dtype=np.float64):
    """-> (w [k, cin_g, cout], b [cout]).  `name.w` is used as stored (SpectralNormalization at inference, or an already
    folded kernel); otherwise tfa WeightNormalization: w = v * g / ||v|| with the norm over every axis but the last."""
    b = np.asarray(weights[name + ".b"], dtype)
    if name + ".w" in weights:
        return np.asarray(weights[name + ".w"], dtype), b
    v = np.asarray(weights[name + ".v"], dtype)
    g = np.asarray(weights[name + ".g"], dtype).reshape(-1)
    nrm = np.sqrt(np.sum(np.square(v), axis=tuple(range(v.ndim - 1))))
    return v * (g / nrm), b


def same_pads(n, k, s, compat, padding):
    """-> (left, right, out_len) of one strided conv / pool over n samples."""
    if padding == "valid" and compat == "reference":
        return 0, 0, (n - k) // s + 1
    if compat == "upstream":
        p = (k - 1) // 2
        return p, p, (n + 2 * p - k) // s + 1
    out = -(-n // s)
    total = max((out - 1) * s + k - n, 0)
    return total // 2, total - total // 2, out


def _conv(x, w, b, spec, compat):
    """x: torch [N, C, L]; w numpy [k, cin_g, cout]."""
    import torch
    import torch.nn.functional as F
    cin, cout, k, s, groups, padding = spec
    left, right, out_len = same_pads(x.shape[-1], k, s, compat, padding)
    xp = F.pad(x, (left, right))
    wt = torch.from_numpy(np.ascontiguousarray(np.transpose(w, (2, 1, 0)))).to(x.dtype)      # [cout, cin_g, k]
    y = F.conv1d(xp, wt, torch.from_numpy(np.asarray(b)).to(x.dtype), stride=s, groups=groups)
    assert y.shape[-1] == out_len, (y.shape, out_len)
    return y


def _lrelu(x):
    import torch
    return torch.where(x >= 0, x, x * LRELU_SLOPE)


def disc_p_forward(weights, idx, period, y, compat="reference", dtype="float32"):
    """One DiscriminatorP (model.py:222-243): y numpy [B, T] -> (score [B, H*period], fmaps [[B, H, period, C], ...])."""
    import torch
    td = getattr(torch, dtype)
    x = torch.from_numpy(np.asarray(y, np.float64)).to(td)
    B, T = x.shape
    if T % period:                                              # model.py:228-233: reflect-pad the END
        n_pad = period - T % period
        x = torch.nn.functional.pad(x[:, None, :], (0, n_pad), mode="reflect")[:, 0]
        T += n_pad
    H = T // period
    # [B, H, period] -> period independent sequences per utterance: [B * period, 1, H]
    x = x.reshape(B, H, period).permute(0, 2, 1).reshape(B * period, 1, H)
    fmaps = []
    npd = np.float64 if dtype == "float64" else np.float32
    for l, spec in enumerate(P_CONVS):
        w, b = fold(weights, f"discriminators.{idx}.convs.{l}", npd)
        x = _lrelu(_conv(x, w, b, spec, compat))
        fmaps.append(x)
    w, b = fold(weights, f"discriminators.{idx}.conv_post", npd)
    x = _conv(x, w, b, P_POST, compat)
    fmaps.append(x)
    outs = [f.reshape(B, period, f.shape[1], f.shape[2]).permute(0, 3, 1, 2).contiguous().numpy() for f in fmaps]   # [B, H, period, C]
    return outs[-1].reshape(B, -1), outs


def disc_s_forward(weights, idx, y, compat="reference", dtype="float32"):
    """One DiscriminatorS (model.py:334-346): y numpy [B, T] -> (score [B, L], fmaps [[B, L, C], ...])."""
    import torch
    td = getattr(torch, dtype)
    x = torch.from_numpy(np.asarray(y, np.float64)).to(td)[:, None, :]
    fmaps = []
    npd = np.float64 if dtype == "float64" else np.float32
    for l, spec in enumerate(S_CONVS):
        w, b = fold(weights, f"discriminators.{idx}.convs.{l}", npd)
        x = _lrelu(_conv(x, w, b, spec, compat))
        fmaps.append(x)
    w, b = fold(weights, f"discriminators.{idx}.conv_post", npd)
    x = _conv(x, w, b, S_POST, compat)
    fmaps.append(x)
    outs = [f.permute(0, 2, 1).contiguous().numpy() for f in fmaps]
    return outs[-1].reshape(outs[-1].shape[0], -1), outs


def mean_pool(y, compat="reference"):
    """AveragePooling1D(4, 2, "same") (model.py:357-364) over numpy [B, T], in float64."""
    y = np.asarray(y, np.float64)
    n = y.shape[1]
    if compat == "upstream":                                    # AvgPool1d(4, 2, padding=1): zeros count
        left, out = 1, (n + 2 - 4) // 2 + 1
    else:
        left, _, out = same_pads(n, 4, 2, "reference", "same")
    res = np.zeros((y.shape[0], out))
    for o in range(out):
        lo, hi = o * 2 - left, o * 2 - left + 4
        a, b = max(lo, 0), min(hi, n)
        res[:, o] = y[:, a:b].sum(axis=1) / (4 if compat == "upstream" else (b - a))
    return res


def mpd_forward(weights, y, compat="reference", dtype="float32"):
    """MultiPeriodDiscriminator on ONE input (model.py:259-274 runs it on y and on y_hat): -> (scores, fmaps) lists."""
    scores, fmaps = [], []
    for i, p in enumerate(MPD_PERIODS):
        s, f = disc_p_forward(weights, i, p, y, compat, dtype)
        scores.append(s); fmaps.append(f)
    return scores, fmaps


def msd_forward(weights, y, compat="reference", dtype="float32"):
    """MultiScaleDiscriminator on ONE input (model.py:367-389): scale i sees the input pooled i times."""
    scores, fmaps = [], []
    y = np.asarray(y, np.float64)
    for i in range(3):
        if i:
            y = mean_pool(y, compat)
            if dtype == "float32":
                y = y.astype(np.float32).astype(np.float64)     # the reference pools in fp32 storage
        s, f = disc_s_forward(weights, i, y, compat, dtype)
        scores.append(s); fmaps.append(f)
    return scores, fmaps


def synthetic_weights(kind, variant="A", seed=4321):
    """Seeded Glorot-uniform weights (Keras' default initialiser) for every layer of `kind`; variant "B" perturbs g and
    uses non-zero biases; the first DiscriminatorS (spectral norm) gets a plain kernel `w`."""
    out = {}
    for n, (name, spec) in enumerate(layer_specs(kind)):
        cin, cout, k, _, groups, _ = spec
        rng = np.random.default_rng(seed + n)
        shape = kernel_shape(spec)
        lim = np.sqrt(6.0 / (k * (cin // groups) + k * cout))
        v = rng.uniform(-lim, lim, size=shape).astype(np.float32)
        b = np.zeros(cout, np.float32)
        g = np.sqrt(np.sum(np.square(v.astype(np.float64)), axis=(0, 1)))
        if variant == "B":
            g = g * rng.uniform(0.5, 1.5, size=g.shape)
            b = rng.uniform(-0.05, 0.05, size=cout).astype(np.float32)
        if kind == "msd" and name.startswith("discriminators.0."):
            out[name + ".w"] = (v * (g / np.sqrt(np.sum(np.square(v.astype(np.float64)), axis=(0, 1))))).astype(np.float32)
        else:
            out[name + ".v"] = v
            out[name + ".g"] = g.astype(np.float32)
        out[name + ".b"] = b
    return out


# ---------------------------------------------------------------- loss terms (reference HiFiGAN_TF/losses.py), float64
def feature_loss(fmap_r, fmap_g):
    """losses.py:8-17: 2 * sum over sub-discriminators and their maps of mean |real - generated|."""
    loss = 0.0
    for dr, dg in zip(fmap_r, fmap_g):
        for rl, gl in zip(dr, dg):
            loss += float(np.mean(np.abs(np.asarray(rl, np.float64) - np.asarray(gl, np.float64))))
    return loss * 2


def discriminator_loss(disc_real_outputs, disc_generated_outputs):
    """losses.py:20-33 -> (loss, r_losses, g_losses)."""
    loss, r_losses, g_losses = 0.0, [], []
    for dr, dg in zip(disc_real_outputs, disc_generated_outputs):
        r = float(np.mean((1 - np.asarray(dr, np.float64)) ** 2))
        g = float(np.mean(np.asarray(dg, np.float64) ** 2))
        loss += r + g
        r_losses.append(r); g_losses.append(g)
    return loss, r_losses, g_losses


def generator_loss(disc_outputs):
    """losses.py:36-46 -> (loss, gen_losses)."""
    gen_losses = [float(np.mean((1 - np.asarray(dg, np.float64)) ** 2)) for dg in disc_outputs]
    return float(sum(gen_losses)), gen_losses
