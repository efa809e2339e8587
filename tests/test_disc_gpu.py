"""Discriminator forward passes on the GPU (hfg_disc_*, csrc/disc.cu) against the CPU oracle (oracle/disc_oracle.py) and
the committed fixtures of the vendored torch discriminators."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 2e-4            # max |cuda - oracle| per feature map, relative to that map's max |value|.  The FFMA kernels differ from
                      # torch-CPU fp32 by summation order only (measured <= 4e-6, TOL_FFMA); the default build runs every conv
                      # but the first and conv_post on the tensor cores as 3xTF32, whose fp32 accumulator truncates instead of
                      # rounding: measured <= 6.2e-5 over up to six such layers (plain TF32, TensorFlow's GPU default, is ~1e-3)
TOL_FFMA = 2e-5


def _waveform(b, t, seed=99):
    rng = np.random.default_rng(seed)
    n = np.arange(t)[None, :]
    f = rng.uniform(0.01, 0.2, size=(b, 1))
    return (0.6 * np.sin(2 * np.pi * f * n) + 0.2 * rng.standard_normal((b, t))).astype(np.float32)


def _positions(n):
    return (np.arange(97) * max(n // 97, 1)) % n


def _model(kind, compat, variant="B"):
    from neuraltexttospeech_b200.discriminators import MultiPeriodDiscriminator, MultiScaleDiscriminator
    from oracle import disc_oracle as D
    w = D.synthetic_weights(kind, variant)
    cls = MultiPeriodDiscriminator if kind == "mpd" else MultiScaleDiscriminator
    return cls(compat=compat).load_weights(w), w


def _compare(scores, fmaps, o_scores, o_fmaps, tag, tol=TOL):
    worst = 0.0
    for i, (s, fm, os_, ofm) in enumerate(zip(scores, fmaps, o_scores, o_fmaps)):
        assert len(fm) == len(ofm)
        for l, (f, of) in enumerate(zip(fm, ofm)):
            f = np.asarray(f)
            assert tuple(f.shape) == tuple(of.shape), (tag, i, l, f.shape, of.shape)
            d = float(np.abs(f - of).max())
            scale = float(np.abs(of).max())                   # random-init maps shrink with depth: judge each map on its own scale
            assert scale > 1e-7, (tag, i, l)                  # maps are not degenerate
            worst = max(worst, d / scale)
            assert d <= tol * scale, (tag, i, l, d, scale)
        assert tuple(np.asarray(s).shape) == tuple(os_.shape)
        assert np.array_equal(np.asarray(s), np.asarray(fm[-1]).reshape(s.shape[0], -1))   # the score is the flattened last map
    print(tag, "worst max-abs difference / map max: %.2e" % worst)
    return worst


@pytest.mark.parametrize("kind", ["mpd", "msd"])
@pytest.mark.parametrize("compat", ["reference", "upstream"])
@pytest.mark.parametrize("B,T", [(2, 1031), (1, 2048), (3, 515)])
def test_forward_matches_oracle(kind, compat, B, T):
    from oracle import disc_oracle as D
    model, w = _model(kind, compat)
    y = _waveform(B, T, seed=7 + T)
    scores, fmaps = model.forward(y)                          # numpy in -> host path (hfg_disc_forward_host)
    o_scores, o_fmaps = (D.mpd_forward if kind == "mpd" else D.msd_forward)(w, y, compat=compat)
    _compare(scores, fmaps, o_scores, o_fmaps, (kind, compat, B, T))


@pytest.mark.parametrize("kind", ["mpd", "msd"])
def test_upstream_mode_matches_the_vendored_torch_fixture(kind):
    fx = np.load(os.path.join(GOLDEN, f"disc_upstream_{kind}.npz"))
    model, _ = _model(kind, "upstream")
    y = _waveform(int(fx["B"]), int(fx["T"]))
    scores, fmaps = model.forward(y)
    for i, (s, fm) in enumerate(zip(scores, fmaps)):
        np.testing.assert_allclose(s, fx[f"score_{i}"], rtol=0, atol=TOL * np.abs(fx[f"score_{i}"]).max())
        for l, f in enumerate(fm):
            assert tuple(f.shape) == tuple(fx[f"shape_{i}_{l}"])
            flat = np.ascontiguousarray(f).astype(np.float64).reshape(-1)
            samples = fx[f"samples_{i}_{l}"]
            assert np.abs(flat[_positions(flat.size)] - samples).max() <= TOL * np.abs(samples).max()


@pytest.mark.parametrize("kind", ["mpd", "msd"])
def test_reference_call_convention_and_device_path(kind):
    """mpd((y, y_hat)) -> y_d_rs, y_d_gs, fmap_rs, fmap_gs (model.py:259-274 / 367-389) on torch CUDA tensors [B, T, 1];
    the device path must give the bits of the host path, and the pair call the bits of two single calls."""
    import torch
    model, _ = _model(kind, "reference", "A")
    y, y_hat = _waveform(2, 777, seed=1), _waveform(2, 777, seed=2)
    yt, yht = torch.from_numpy(y).cuda()[:, :, None], torch.from_numpy(y_hat).cuda()[:, :, None]
    y_d_rs, y_d_gs, fmap_rs, fmap_gs = model((yt, yht))
    s_r, f_r = model.forward(y)
    s_g, f_g = model.forward(y_hat)
    assert len(y_d_rs) == len(y_d_gs) == len(fmap_rs) == len(fmap_gs) == model.N_SUB
    for i in range(model.N_SUB):
        assert np.array_equal(y_d_rs[i].cpu().numpy(), s_r[i]) and np.array_equal(y_d_gs[i].cpu().numpy(), s_g[i])
        assert len(fmap_rs[i]) == model.N_LAYERS
        for l in range(model.N_LAYERS):
            assert np.array_equal(fmap_rs[i][l].cpu().numpy(), f_r[i][l])
            assert np.array_equal(fmap_gs[i][l].cpu().numpy(), f_g[i][l])
    if kind == "mpd":
        assert tuple(fmap_rs[2][0].shape)[2] == 5            # [B, len, period, C] like the reference's reshape
        assert fmap_rs[2][0].shape[3] == 32


def test_training_segment_shape_matches_oracle():
    """segment_size 8192 (config_v1.json) -- the shape gan.py feeds the discriminators; B = 2 keeps the oracle quick."""
    from oracle import disc_oracle as D
    y = _waveform(2, 8192, seed=5)
    for kind in ("mpd", "msd"):
        model, w = _model(kind, "reference", "A")
        scores, fmaps = model.forward(y)
        o_scores, o_fmaps = (D.mpd_forward if kind == "mpd" else D.msd_forward)(w, y, compat="reference")
        _compare(scores, fmaps, o_scores, o_fmaps, (kind, 8192))


def test_errors():
    from neuraltexttospeech_b200.discriminators import MultiScaleDiscriminator
    from oracle import disc_oracle as D
    m = MultiScaleDiscriminator()
    with pytest.raises(RuntimeError):                         # weights not finalised
        m.forward(_waveform(1, 512))
    w = D.synthetic_weights("msd", "A")
    bad = dict(w); bad["discriminators.1.convs.2.v"] = np.zeros((41, 8, 255), np.float32)
    with pytest.raises(ValueError):
        m.load_weights(bad)
    m.load_weights(w)
    with pytest.raises(ValueError):                           # shorter than the "valid" k = 15 first conv
        m.forward(_waveform(1, 10))


@pytest.mark.parametrize("kind", ["mpd", "msd"])
def test_tensor_core_layers_match_the_ffma_path(kind, monkeypatch):
    """Strided, grouped and dense convs run as 3xTF32 on the tensor cores by default; HFG_DISC_TC=0 keeps them on the FFMA
    kernel.  Both must sit inside their oracle tolerance, and within TOL of each other on every map."""
    from oracle import disc_oracle as D
    y = _waveform(2, 4099, seed=11)
    model_tc, w = _model(kind, "reference", "B")
    monkeypatch.setenv("HFG_DISC_TC", "0")
    model_ff, _ = _model(kind, "reference", "B")
    monkeypatch.delenv("HFG_DISC_TC")
    s_tc, f_tc = model_tc.forward(y)
    s_ff, f_ff = model_ff.forward(y)
    o_scores, o_fmaps = (D.mpd_forward if kind == "mpd" else D.msd_forward)(w, y, compat="reference")
    _compare(s_tc, f_tc, o_scores, o_fmaps, (kind, "3xTF32"))
    _compare(s_ff, f_ff, o_scores, o_fmaps, (kind, "ffma"), tol=TOL_FFMA)
    differs = False
    for fa, fb in zip(f_tc, f_ff):
        for l, (a, b) in enumerate(zip(fa, fb)):
            if l == 0:
                assert np.array_equal(a, b)                   # c_in = 1: the same FFMA kernel in both
            else:
                differs = differs or not np.array_equal(a, b)
                assert np.abs(a - b).max() <= TOL * np.abs(b).max()
    assert differs                                            # i.e. the knob really switches kernels


@pytest.mark.parametrize("kind", ["mpd", "msd"])
def test_pair_losses_match_the_reference_formulas(kind):
    """hfg_disc_pair_losses (losses.py:8-43 reduced on the GPU) against the oracle's formulas applied (a) to the maps the
    CUDA forward itself returned -- only the reduction differs -- and (b) to the oracle's own maps."""
    from oracle import disc_oracle as D
    model, w = _model(kind, "reference", "B")
    y, y_hat = _waveform(3, 2051, seed=21), _waveform(3, 2051, seed=22)
    got = model.pair_losses((y, y_hat))
    again = model.pair_losses((y[:, :, None], y_hat[:, :, None]))
    assert got == again                                       # deterministic, and [B, T, 1] is accepted
    y_d_rs, y_d_gs, fmap_rs, fmap_gs = model((y, y_hat))
    fwd = D.mpd_forward if kind == "mpd" else D.msd_forward
    (o_rs, o_fr), (o_gs, o_fg) = fwd(w, y), fwd(w, y_hat)
    for tag, (rs, gs, fr, fg), tol in (("own maps", (y_d_rs, y_d_gs, fmap_rs, fmap_gs), 2e-6), ("oracle maps", (o_rs, o_gs, o_fr, o_fg), 2e-4)):
        fm = D.feature_loss(fr, fg)
        dl, r_l, g_l = D.discriminator_loss(rs, gs)
        gl, gen_l = D.generator_loss(gs)
        assert abs(got["feature_loss"] - fm) <= tol * fm, (tag, got["feature_loss"], fm)
        assert abs(got["discriminator_loss"] - dl) <= tol * dl, tag
        assert abs(got["generator_loss"] - gl) <= tol * gl, tag
        np.testing.assert_allclose(got["r_losses"], r_l, rtol=tol)
        np.testing.assert_allclose(got["g_losses"], g_l, rtol=max(tol, 1e-5), atol=1e-12)
        np.testing.assert_allclose(got["gen_losses"], gen_l, rtol=tol)
    assert got["feature_loss"] > 0 and len(got["r_losses"]) == model.N_SUB
