"""The discriminator oracle (oracle/disc_oracle.py) against the fixtures generated from the torch discriminators the
reference vendors (tests/golden/make_golden_disc.py), plus known-answer checks of the Keras-only semantics."""
import os

import numpy as np
import pytest

from oracle import disc_oracle as D

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _waveform(b, t, seed=99):                # same generator as make_golden_disc.waveform
    rng = np.random.default_rng(seed)
    n = np.arange(t)[None, :]
    f = rng.uniform(0.01, 0.2, size=(b, 1))
    return (0.6 * np.sin(2 * np.pi * f * n) + 0.2 * rng.standard_normal((b, t))).astype(np.float32)


def _positions(n):
    return (np.arange(97) * max(n // 97, 1)) % n


@pytest.mark.parametrize("kind", ["mpd", "msd"])
def test_oracle_upstream_mode_reproduces_the_vendored_torch_discriminators(kind):
    fx = np.load(os.path.join(GOLDEN, f"disc_upstream_{kind}.npz"))
    y = _waveform(int(fx["B"]), int(fx["T"]))
    weights = D.synthetic_weights(kind, "B")
    fwd = D.mpd_forward if kind == "mpd" else D.msd_forward
    for dtype, tol in (("float64", 5e-8), ("float32", 2e-5)):   # the fixture ran on fp32-rounded folded kernels
        scores, fmaps = fwd(weights, y, compat="upstream", dtype=dtype)
        for i, (s, fm) in enumerate(zip(scores, fmaps)):
            want = fx[f"score_{i}"]
            assert s.shape == want.shape
            scale = max(np.abs(want).max(), 1e-6)
            assert np.abs(s - want).max() <= tol * max(scale, 1.0), (kind, i, dtype)
            for l, f in enumerate(fm):
                assert tuple(f.shape) == tuple(fx[f"shape_{i}_{l}"]), (kind, i, l)
                flat = f.astype(np.float64).reshape(-1)
                samples = fx[f"samples_{i}_{l}"]
                assert np.abs(flat[_positions(flat.size)] - samples).max() <= tol * max(np.abs(samples).max(), 1.0)
                sums = fx[f"sums_{i}_{l}"]
                assert abs(np.abs(flat).sum() - sums[1]) <= tol * 50 * max(sums[1], 1.0)


def test_tensorflow_same_padding_rule():
    # kernel 5, stride 3 (DiscriminatorP): the left pad depends on the length modulo the stride
    assert D.same_pads(9, 5, 3, "reference", "same") == (1, 1, 3)
    assert D.same_pads(10, 5, 3, "reference", "same") == (2, 2, 4)
    assert D.same_pads(11, 5, 3, "reference", "same") == (1, 2, 4)
    # kernel 41, stride 2 / 4 (DiscriminatorS)
    assert D.same_pads(100, 41, 2, "reference", "same") == (19, 20, 50)
    assert D.same_pads(101, 41, 2, "reference", "same") == (20, 20, 51)
    assert D.same_pads(100, 41, 4, "reference", "same") == (18, 19, 25)
    # the first DiscriminatorS conv is "valid" in the reference, padded by 7 upstream
    assert D.same_pads(100, 15, 1, "reference", "valid") == (0, 0, 86)
    assert D.same_pads(100, 15, 1, "upstream", "valid") == (7, 7, 100)
    # torch: (k - 1) / 2 on both sides
    assert D.same_pads(11, 5, 3, "upstream", "same") == (2, 2, 4)


def test_mean_pool_ignores_padding_in_reference_mode():
    y = np.arange(1, 8, dtype=np.float64)[None, :]             # 7 samples: SAME pads 1 left, 2 right, 4 outputs
    ref = D.mean_pool(y, "reference")
    assert ref.shape == (1, 4)
    np.testing.assert_allclose(ref[0], [(1 + 2 + 3) / 3, (2 + 3 + 4 + 5) / 4, (4 + 5 + 6 + 7) / 4, (6 + 7) / 2])
    up = D.mean_pool(np.arange(1, 9, dtype=np.float64)[None, :], "upstream")   # AvgPool1d(4, 2, padding=1) counts the zeros
    np.testing.assert_allclose(up[0], [(1 + 2 + 3) / 4, (2 + 3 + 4 + 5) / 4, (4 + 5 + 6 + 7) / 4, (6 + 7 + 8) / 4])


def test_reference_mode_shapes_follow_keras():
    """DiscriminatorS loses 14 samples in its "valid" first conv (model.py:286-289); DiscriminatorP pads the end by
    reflection up to a multiple of the period (model.py:228-233); every strided layer yields ceil(n / s)."""
    y = _waveform(1, 301)
    w = D.synthetic_weights("msd", "A")
    _, fm = D.disc_s_forward(w, 1, y, "reference")
    lens = [f.shape[1] for f in fm]
    n = 301 - 14
    want = [n]
    for s in (2, 2, 4, 4, 1, 1, 1):
        n = -(-n // s); want.append(n)
    assert lens == want
    wp = D.synthetic_weights("mpd", "A")
    score, fm = D.disc_p_forward(wp, 2, 5, y, "reference")
    H = -(-301 // 5)
    hs = [H]
    for s in (3, 3, 3, 3, 1, 1):
        hs.append(-(-hs[-1] // s))
    assert [f.shape[1] for f in fm] == hs[1:] and all(f.shape[2] == 5 for f in fm)
    assert score.shape == (1, hs[-1] * 5)
    # reflect padding: the padded tail mirrors the samples before the last one
    yp = np.concatenate([y, y[:, -2:-6:-1]], axis=1)           # 301 -> 305 by hand
    score2, _ = D.disc_p_forward(wp, 2, 5, yp, "reference")
    np.testing.assert_allclose(score, score2, rtol=0, atol=1e-7)


def test_package_and_oracle_weight_generators_agree():
    """bench.py / tools may not import oracle/: the package carries its own copy of the seeded generator."""
    from neuraltexttospeech_b200.synthetic import synthetic_disc_weights
    for kind in ("mpd", "msd"):
        for variant in "AB":
            a, b = synthetic_disc_weights(kind, variant), D.synthetic_weights(kind, variant)
            assert a.keys() == b.keys()
            assert all(np.array_equal(a[k], b[k]) for k in a)
