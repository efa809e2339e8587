"""Operator-level parity on the GPU: each tapped-conv kernel (direct FFMA and tcgen05) against a
numpy restatement of the same op, over the channel shapes the three stock configs use."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def ref_tapped_conv(x, w, bias, tap_off, out_stride=1, out_off=0, in_slope=1.0, res=None):
    """numpy fp64 restatement of kernels.cuh ConvArgs semantics."""
    B, rows, cin = x.shape
    xa = np.where(x >= 0, x, x * in_slope).astype(np.float64)
    out = np.zeros((B, rows * out_stride, w.shape[2]))
    val = np.zeros((B, rows, w.shape[2]))
    for j, off in enumerate(tap_off):
        lo, hi = max(0, -off), min(rows, rows - off)
        if hi > lo:
            val[:, lo:hi] += xa[:, lo + off:hi + off] @ w[j].astype(np.float64)
    val += bias
    out[:, out_off::out_stride] = val
    if res is not None:
        out[:, out_off::out_stride] += res[:, out_off::out_stride]
    return out


SHAPES = [  # (c_in, c_out, taps, B, rows)
    (32, 32, [-2, -1, 0], 2, 700),                       # causal k=3 (reference quirk Q1)
    (32, 32, [5 * j - 25 for j in range(11)], 1, 1300),  # k=11, d=5
    (64, 64, [3 * j - 9 for j in range(7)], 2, 520),
    (128, 128, [j - 5 for j in range(11)], 1, 300),
    (256, 256, [5 * j - 25 for j in range(11)], 1, 270),
    (512, 256, [-1, 0], 1, 140),                         # one phase of ups[0]
    (256, 128, [0, 1], 1, 200),
    (128, 64, [-1, 0], 1, 600),
    (64, 32, [0, 1], 2, 515),
    (128, 256, [j - 3 for j in range(7)], 2, 130),       # conv_pre (mel padded 80 -> 128), half of C_out=512
    (128, 512, [j - 3 for j in range(7)], 1, 130),       # conv_pre, N split over grid.y
    (128, 128, [12 * j - 36 for j in range(7)], 1, 400), # V3 k=7 d=12
]


def lrelu(v, s):
    return np.where(v >= 0, v, v * np.float32(s)).astype(np.float32)


@pytest.mark.parametrize("cin,cout,taps,B,rows", SHAPES)
@pytest.mark.parametrize("engine,dtype", [("ffma", "float32"), ("tc", "float16"), ("tc", "bfloat16"), ("ffma", "float16")])
def test_tapped_conv(cin, cout, taps, B, rows, engine, dtype):
    import torch
    from neuraltexttospeech_b200.ops import tapped_conv
    rng = np.random.default_rng(cin * 1000 + cout + len(taps))
    tdt = getattr(torch, dtype)
    x = rng.standard_normal((B, rows, cin)).astype(np.float32)
    w = (rng.standard_normal((len(taps), cin, cout)) / np.sqrt(cin * len(taps))).astype(np.float32)
    bias = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
    res = rng.standard_normal((B, rows, cout)).astype(np.float32)
    tol = {"float32": 2e-5, "float16": 4e-3, "bfloat16": 3e-2}[dtype]   # output rounding of O(1) values
    if engine == "ffma":
        xt = torch.from_numpy(x).cuda().to(tdt)
        rt = torch.from_numpy(res).cuda().to(tdt)
        xq, rq = xt.float().cpu().numpy(), rt.float().cpu().numpy()
        ref = ref_tapped_conv(xq, w, bias, taps, in_slope=0.1, res=rq)       # the direct kernel keeps fp32 weights
        out = tapped_conv(xt, w, bias, taps, engine="ffma", in_slope=0.1, res=rt).float().cpu().numpy()
        assert np.abs(out - ref).max() <= tol
        return
    # tcgen05 kernel: x and res are stored already activated (slope 0.1); output stored activated with 0.3
    xt = torch.from_numpy(lrelu(x, 0.1)).cuda().to(tdt)
    rt = torch.from_numpy(lrelu(res, 0.1)).cuda().to(tdt)
    xq, rq = xt.float().cpu().numpy(), rt.float().cpu().numpy()
    wq = torch.from_numpy(w).to(tdt).float().numpy()
    res_raw = np.where(rq >= 0, rq, rq * np.float32(10.0))
    ref = lrelu(ref_tapped_conv(xq, wq, bias, taps, in_slope=1.0, res=res_raw), 0.3)
    out = tapped_conv(xt, w, bias, taps, engine="tc", res=rt, res_inv=10.0, out_slope=0.3).float().cpu().numpy()
    assert np.abs(out - ref).max() <= tol


def test_tc_mrf_sum_and_scale():
    """res2 (running MRF sum) and out_scale = 1/num_kernels of the tcgen05 epilogue (reference model.py:161-167)."""
    import torch
    from neuraltexttospeech_b200.ops import tapped_conv
    rng = np.random.default_rng(9)
    B, rows, c = 2, 333, 64
    x = torch.from_numpy(rng.standard_normal((B, rows, c)).astype(np.float32)).cuda().half()
    res = torch.from_numpy(rng.standard_normal((B, rows, c)).astype(np.float32)).cuda().half()
    s2 = torch.from_numpy(rng.standard_normal((B, rows, c)).astype(np.float32)).cuda().half()
    w = (rng.standard_normal((3, c, c)) / 14).astype(np.float32)
    bias = rng.uniform(-0.1, 0.1, c).astype(np.float32)
    taps = [-2, -1, 0]
    wq = torch.from_numpy(w).half().float().numpy()
    ref = (ref_tapped_conv(x.float().cpu().numpy(), wq, bias, taps, res=res.float().cpu().numpy()) + s2.float().cpu().numpy()) / 3
    out = tapped_conv(x, w, bias, taps, engine="tc", res=res, res_inv=1.0, res2=s2, out_scale=1.0 / 3).float().cpu().numpy()
    assert np.abs(out - ref).max() <= 4e-3
    # in place on the sum buffer (how the engine accumulates xs)
    s2b = s2.clone()
    tapped_conv(x, w, bias, taps, engine="tc", res=res, res_inv=1.0, res2=s2b, out=s2b)
    assert np.abs(s2b.float().cpu().numpy() - ref * 3).max() <= 8e-3


@pytest.mark.parametrize("engine,dtype", [("ffma", "float32"), ("tc", "float16")])
def test_transposed_phases(engine, dtype):
    import torch
    from neuraltexttospeech_b200.ops import tapped_conv
    rng = np.random.default_rng(5)
    tdt = getattr(torch, dtype)
    B, rows, cin, cout, u = 2, 300, 64, 32, 2
    x = torch.from_numpy(rng.standard_normal((B, rows, cin)).astype(np.float32)).cuda().to(tdt)
    bias = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
    out = torch.zeros((B, rows * u, cout), dtype=tdt, device="cuda")
    ref = np.zeros((B, rows * u, cout))
    xq = x.float().cpu().numpy()
    for p in range(u):
        w = (rng.standard_normal((2, cin, cout)) / 12).astype(np.float32)
        taps = [-1, 0] if p == 0 else [0, 1]
        tapped_conv(x, w, bias, taps, engine=engine, out_stride=u, out_off=p, out=out)
        wq = w if engine == "ffma" else torch.from_numpy(w).to(tdt).float().numpy()
        ref += ref_tapped_conv(xq, wq, bias, taps, out_stride=u, out_off=p)      # other phases' rows stay 0
    tol = 2e-5 if dtype == "float32" else 4e-3
    assert np.abs(out.float().cpu().numpy() - ref).max() <= tol


def test_ffma_mrf_modes():
    # MRF: acc = v0; acc += v1; out = (acc + v2) / 3   (reference model.py:161-167)
    import torch
    from neuraltexttospeech_b200.ops import tapped_conv
    rng = np.random.default_rng(6)
    B, rows, cin, cout = 2, 300, 64, 32
    x = torch.from_numpy(rng.standard_normal((B, rows, cin)).astype(np.float32)).cuda()
    bias = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
    acc = torch.zeros((B, rows, cout), dtype=torch.float32, device="cuda")
    ws = [(rng.standard_normal((3, cin, cout)) / 14).astype(np.float32) for _ in range(3)]
    taps = [-1, 0, 1]
    tapped_conv(x, ws[0], bias, taps, engine="ffma", acc=acc, mrf_mode=1)
    tapped_conv(x, ws[1], bias, taps, engine="ffma", acc=acc, mrf_mode=2)
    fin = tapped_conv(x, ws[2], bias, taps, engine="ffma", acc=acc, mrf_mode=3, mrf_div=3.0, out_slope=0.3)
    xq = x.cpu().numpy()
    tot = sum(ref_tapped_conv(xq, w, bias, taps) for w in ws) / 3
    assert np.abs(fin.cpu().numpy() - lrelu(tot, 0.3)).max() <= 2e-5


# ---------------------------------------------------------------- fused ResBlock1 iteration (resblock_tc.cu)
def ref_resblock_fused(xq, w1, b1, off1, w2, b2, off2, mid_slope, res_inv, res2, out_scale, out_slope, mid_dtype):
    """numpy fp64 restatement of kernels.cuh RbArgs semantics on the SAME 16-bit inputs; the intermediate is rounded to
    the storage type like the kernel's shared-memory tile."""
    import torch
    xt = ref_tapped_conv(xq, w1, b1, off1)                                        # conv1 on the stored (activated) x
    xt = np.where(xt >= 0, xt, xt * mid_slope)
    xt = torch.from_numpy(xt.astype(np.float32)).to(mid_dtype).float().numpy()    # 16-bit xt tile
    v = ref_tapped_conv(xt, w2, b2, off2)
    v += np.where(xq >= 0, xq, xq * np.float32(res_inv))                          # skip connection on the un-activated x
    if res2 is not None:
        v += res2
    v *= out_scale
    return np.where(v >= 0, v, v * out_slope)


def folded_taps(k, d, pad_left, f):
    return sorted({(s + j * d - pad_left) // f for s in range(f) for j in range(k)})


RB_CASES = [  # (k, dilation, fold, B, rows): the ResBlock1 iterations of V1 / V2 stages with C*fold = 64
    (3, 1, 1, 2, 700),      # causal conv1 AND causal conv2 (reference quirk Q1): taps [-2,-1,0]
    (3, 3, 1, 1, 1500),
    (7, 5, 1, 2, 900),
    (11, 5, 1, 1, 2100),    # k = 11: resident weights leave room for one xt tile only / one block per tile
    (11, 1, 2, 2, 640),     # C = 32 folded by 2
    (11, 5, 2, 1, 1111),    # 17 + 7 kernel-view taps
    (7, 3, 8, 3, 130),      # C = 8 folded by 8 (V2 stage 3); utterance shorter than a tile
    (3, 5, 4, 2, 257),
]


@pytest.mark.parametrize("k,d,f,B,rows", RB_CASES)
@pytest.mark.parametrize("dtype", ["float16", "bfloat16"])
def test_resblock_fused_matches_two_convs(k, d, f, B, rows, dtype):
    import torch
    from neuraltexttospeech_b200.ops import resblock_fused
    from neuraltexttospeech_b200.utils import get_padding
    rng = np.random.default_rng(k * 100 + d * 10 + f)
    tdt = getattr(torch, dtype)

    def pad_left(kk, dd):                       # reference HiFiGAN_TF/utils.py:29-37 -> keras padding mode
        return dd * (kk - 1) if get_padding(kk, dd) == "causal" else dd * (kk - 1) // 2

    off1, off2 = folded_taps(k, d, pad_left(k, d), f), folded_taps(k, 1, pad_left(k, 1), f)
    x = rng.standard_normal((B, rows, 64)).astype(np.float32)
    xt = torch.from_numpy(lrelu(x, 0.1)).cuda().to(tdt)
    xq = xt.float().cpu().numpy()
    w1 = (rng.standard_normal((len(off1), 64, 64)) / np.sqrt(64 * len(off1))).astype(np.float32)
    w2 = (rng.standard_normal((len(off2), 64, 64)) / np.sqrt(64 * len(off2))).astype(np.float32)
    b1 = rng.uniform(-0.1, 0.1, 64).astype(np.float32); b2 = rng.uniform(-0.1, 0.1, 64).astype(np.float32)
    w1q = torch.from_numpy(w1).to(tdt).float().numpy(); w2q = torch.from_numpy(w2).to(tdt).float().numpy()
    tol = {"float16": 6e-3, "bfloat16": 5e-2}[dtype]
    # plain iteration: output stored activated with 0.1
    ref = ref_resblock_fused(xq, w1q, b1, off1, w2q, b2, off2, 0.1, 10.0, None, 1.0, 0.1, tdt)
    out = resblock_fused(xt, w1, b1, off1, w2, b2, off2).float().cpu().numpy()
    err = np.abs(out - ref)
    assert err.max() <= tol, (err.max(), np.unravel_index(err.argmax(), err.shape))
    # last iteration of the last resblock: + running MRF sum, / num_kernels, stored for conv_post (slope 0.3)
    s2 = torch.from_numpy(rng.standard_normal((B, rows, 64)).astype(np.float32)).cuda().to(tdt)
    ref = ref_resblock_fused(xq, w1q, b1, off1, w2q, b2, off2, 0.1, 10.0, s2.float().cpu().numpy(), 1.0 / 3, 0.3, tdt)
    out = resblock_fused(xt, w1, b1, off1, w2, b2, off2, res2=s2, out_scale=1.0 / 3, out_slope=0.3).float().cpu().numpy()
    assert np.abs(out - ref).max() <= tol
    # in place on the sum buffer (xs += rb_j(x)), raw output
    s2b = s2.clone()
    resblock_fused(xt, w1, b1, off1, w2, b2, off2, res2=s2b, out_slope=1.0, out=s2b)
    ref = ref_resblock_fused(xq, w1q, b1, off1, w2q, b2, off2, 0.1, 10.0, s2.float().cpu().numpy(), 1.0, 1.0, tdt)
    assert np.abs(s2b.float().cpu().numpy() - ref).max() <= 2 * tol


RB2_CASES = [  # (k, (d_a, d_b), fold, B, rows): the ResBlock2s of V3 stages 1-2 (C*fold = 64)
    (3, (1, 2), 1, 2, 700),      # causal c_a (k=3, d=1), "same" c_b
    (5, (2, 6), 1, 1, 1500),
    (7, (3, 12), 1, 2, 900),     # span of c_b = 72 rows: 184 valid rows per 256-row tile
    (7, (3, 12), 2, 1, 1111),    # C = 32 folded by 2
    (3, (1, 2), 2, 3, 130),
]


@pytest.mark.parametrize("k,ds,f,B,rows", RB2_CASES)
def test_resblock2_fused_matches_two_convs_with_skips(k, ds, f, B, rows):
    """mode 2 of the fused kernel = a whole ResBlock2 (reference model.py:106-112)."""
    import torch
    from neuraltexttospeech_b200.ops import resblock_fused
    from neuraltexttospeech_b200.utils import get_padding
    rng = np.random.default_rng(k * 100 + ds[1] * 10 + f)
    tdt = torch.float16

    def pad_left(kk, dd):
        return dd * (kk - 1) if get_padding(kk, dd) == "causal" else dd * (kk - 1) // 2

    off1, off2 = folded_taps(k, ds[0], pad_left(k, ds[0]), f), folded_taps(k, ds[1], pad_left(k, ds[1]), f)
    x = rng.standard_normal((B, rows, 64)).astype(np.float32)
    xt = torch.from_numpy(lrelu(x, 0.1)).cuda().to(tdt)
    xq = xt.float().cpu().numpy()
    w1 = (rng.standard_normal((len(off1), 64, 64)) / np.sqrt(64 * len(off1))).astype(np.float32)
    w2 = (rng.standard_normal((len(off2), 64, 64)) / np.sqrt(64 * len(off2))).astype(np.float32)
    b1 = rng.uniform(-0.1, 0.1, 64).astype(np.float32); b2 = rng.uniform(-0.1, 0.1, 64).astype(np.float32)
    w1q = torch.from_numpy(w1).to(tdt).float().numpy(); w2q = torch.from_numpy(w2).to(tdt).float().numpy()
    s2 = torch.from_numpy(rng.standard_normal((B, rows, 64)).astype(np.float32)).cuda().to(tdt)

    def ref(res2, out_scale, out_slope):
        x1 = ref_tapped_conv(xq, w1q, b1, off1) + np.where(xq >= 0, xq, xq * np.float32(10.0))
        a1 = torch.from_numpy(np.where(x1 >= 0, x1, x1 * 0.1).astype(np.float32)).to(tdt).float().numpy()   # 16-bit xt tile
        v = ref_tapped_conv(a1, w2q, b2, off2) + np.where(a1 >= 0, a1, a1 * np.float32(10.0))              # x1 as stored in the tile
        if res2 is not None:
            v += res2
        v *= out_scale
        return np.where(v >= 0, v, v * out_slope)

    out = resblock_fused(xt, w1, b1, off1, w2, b2, off2, mode=2).float().cpu().numpy()
    err = np.abs(out - ref(None, 1.0, 0.1))
    assert err.max() <= 8e-3, (err.max(), np.unravel_index(err.argmax(), err.shape))
    out = resblock_fused(xt, w1, b1, off1, w2, b2, off2, mode=2, res2=s2, out_scale=1.0 / 3, out_slope=0.3).float().cpu().numpy()
    assert np.abs(out - ref(s2.float().cpu().numpy(), 1.0 / 3, 0.3)).max() <= 8e-3


# ---------------------------------------------------------------- TF32 tensor-core kernel (conv_tf32.cu)
def tf32_round(a):
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    u = (u + np.uint32(0x0FFF) + ((u >> np.uint32(13)) & np.uint32(1))) & np.uint32(0xFFFFE000)
    return u.view(np.float32)


TF32_SHAPES = [  # (c_in, c_out, taps, B, rows)
    (32, 32, [-2, -1, 0], 2, 700),
    (64, 64, [3 * j - 9 for j in range(7)], 2, 520),
    (128, 128, [5 * j - 25 for j in range(11)], 1, 300),
    (512, 256, [-1, 0], 1, 140),                         # one phase of ups[0], two N tiles
    (96, 512, [j - 3 for j in range(7)], 1, 130),        # conv_pre: 80 mel channels padded to 96
    (16, 16, [-1, 0, 1], 3, 260),                        # 64-byte swizzle rows
    (8, 16, [0, 1], 2, 300),                             # 32-byte swizzle rows, K = 8 (one MMA per tap)
    (128, 64, [-1, 0], 1, 600),
]


@pytest.mark.parametrize("cin,cout,taps,B,rows", TF32_SHAPES)
def test_tapped_conv_tf32(cin, cout, taps, B, rows):
    import torch
    from neuraltexttospeech_b200.ops import tapped_conv
    rng = np.random.default_rng(cin * 1000 + cout + len(taps))
    x = tf32_round(lrelu(rng.standard_normal((B, rows, cin)).astype(np.float32), 0.1))     # stored activated, TF32-rounded
    w = (rng.standard_normal((len(taps), cin, cout)) / np.sqrt(cin * len(taps))).astype(np.float32)
    bias = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
    res = lrelu(rng.standard_normal((B, rows, cout)).astype(np.float32), 0.1)
    res_raw = np.where(res >= 0, res, res * np.float32(10.0))
    ref = lrelu(ref_tapped_conv(x, tf32_round(w), bias, taps, in_slope=1.0, res=res_raw), 0.3)
    out = tapped_conv(torch.from_numpy(x).cuda(), w, bias, taps, engine="tf32", res=torch.from_numpy(res).cuda(), res_inv=10.0,
                      out_slope=0.3).cpu().numpy()
    assert np.abs(out - ref).max() <= 3e-3            # fp32 accumulate of exact TF32 products; the stored output is TF32-rounded
    assert np.array_equal(out, tf32_round(out))       # ... to 10 mantissa bits
