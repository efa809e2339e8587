"""hfg_op_tapped_conv_noalloc (include/hifigan_b200.h): the direct tapped convolution whose weight / bias scratch is a
static device array instead of a per-call cudaMalloc.  Checked through the C ABI against the existing entry point
(hfg_op_tapped_conv, engine 0) on calls in a row that share the scratch.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

SCRATCH_BYTES = 4 << 20          # HFG_NOALLOC_SCRATCH_BYTES
TOL = {"float32": 2e-5, "float16": 4e-3}   # the tolerances of tests/test_ops_gpu.py for the same kernel


def _call_noalloc(lib, x, w, bias, taps, out, *, out_stride=1, out_off=0, in_slope=1.0, res=None, out_slope=1.0):
    import torch
    act = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[x.dtype]
    B, rows, c_in = (int(v) for v in x.shape)
    tap_arr = (C.c_int * len(taps))(*[int(t) for t in taps])
    err = C.create_string_buffer(512)
    rc = lib.hfg_op_tapped_conv_noalloc(
        act, C.c_void_p(x.data_ptr()), B, rows, c_in, w.ctypes.data_as(C.c_void_p), bias.ctypes.data_as(C.c_void_p),
        int(w.shape[2]), len(taps), tap_arr, out_stride, out_off, float(in_slope),
        C.c_void_p(res.data_ptr()) if res is not None else None, float(out_slope), C.c_void_p(out.data_ptr()),
        C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream), err, len(err))
    return rc, err.value.decode()


CASES = [  # (dtype, c_in, c_out, taps, B, rows, out_stride, out_off, in_slope, out_slope)
    ("float32", 64, 64, [3 * j - 9 for j in range(7)], 2, 520, 1, 0, 0.1, 1.0),     # k=7 d=3 resblock conv, 115 KB of scratch
    ("float32", 32, 32, [-2, -1, 0], 2, 700, 1, 0, 0.1, 0.1),                      # causal k=3: smaller, sits inside call 1's bytes
    ("float16", 128, 64, [-1, 0], 1, 600, 8, 3, 0.1, 1.0),                         # one phase of a transposed conv, 16-bit storage
    ("float32", 256, 256, [5 * j - 25 for j in range(11)], 1, 270, 1, 0, 1.0, 1.0),  # 2.9 MB: most of the scratch
    ("float32", 64, 64, [3 * j - 9 for j in range(7)], 2, 520, 1, 0, 0.1, 1.0),     # the first layer again, after a larger one
]


@pytest.mark.gpu
def test_noalloc_matches_the_allocating_entry_point_on_calls_in_a_row():
    import torch
    from neuraltexttospeech_b200 import _lib
    from neuraltexttospeech_b200.ops import tapped_conv
    lib = _lib.load()
    worst = 0.0
    for n, (dtype, cin, cout, taps, B, rows, stride, off, in_slope, out_slope) in enumerate(CASES):
        rng = np.random.default_rng(1234 + n)
        tdt = getattr(torch, dtype)
        x = torch.from_numpy(rng.standard_normal((B, rows, cin)).astype(np.float32)).cuda().to(tdt)
        res = torch.from_numpy(rng.standard_normal((B, rows * stride, cout)).astype(np.float32)).cuda().to(tdt)
        w = (rng.standard_normal((len(taps), cin, cout)) / np.sqrt(cin * len(taps))).astype(np.float32)
        bias = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
        want = tapped_conv(x, w, bias, taps, engine="ffma", out_stride=stride, out_off=off, in_slope=in_slope, res=res,
                           out_slope=out_slope)
        got = torch.zeros_like(want)                         # allocated by the caller, before the call
        rc, msg = _call_noalloc(lib, x, w, bias, taps, got, out_stride=stride, out_off=off, in_slope=in_slope, res=res,
                                out_slope=out_slope)
        assert rc == 0, msg
        torch.cuda.synchronize()
        a, b = got.float().cpu().numpy(), want.float().cpu().numpy()
        assert np.isfinite(a).all()
        assert np.abs(b).max() > 0.1                         # the comparison is not between two empty buffers
        d = float(np.abs(a - b).max())
        worst = max(worst, d)
        assert d <= TOL[dtype], f"case {n}: max |noalloc - allocating| = {d}"
    print(f"noalloc vs allocating entry point: worst max-abs difference {worst:g} over {len(CASES)} calls in a row")


@pytest.mark.gpu
def test_noalloc_back_to_back_without_a_synchronise_between_calls():
    """Two calls enqueued on one stream with no host synchronise in between: stream order is what protects the scratch."""
    import torch
    from neuraltexttospeech_b200 import _lib
    from neuraltexttospeech_b200.ops import tapped_conv
    lib = _lib.load()
    rng = np.random.default_rng(77)
    layers = []
    for cin, cout, taps in [(64, 32, [0, 1]), (32, 32, [-2, -1, 0])]:
        w = (rng.standard_normal((len(taps), cin, cout)) / np.sqrt(cin * len(taps))).astype(np.float32)
        layers.append((w, rng.uniform(-0.1, 0.1, cout).astype(np.float32), taps))
    x = torch.from_numpy(rng.standard_normal((2, 515, 64)).astype(np.float32)).cuda()
    mid = torch.zeros((2, 515, 32), dtype=torch.float32, device="cuda")
    out = torch.zeros_like(mid)
    torch.cuda.synchronize()
    rc1, m1 = _call_noalloc(lib, x, *layers[0], mid, in_slope=0.1)
    rc2, m2 = _call_noalloc(lib, mid, *layers[1], out, in_slope=0.1)
    assert rc1 == 0 and rc2 == 0, (m1, m2)
    torch.cuda.synchronize()
    want_mid = tapped_conv(x, layers[0][0], layers[0][1], layers[0][2], engine="ffma", in_slope=0.1)
    want = tapped_conv(want_mid, layers[1][0], layers[1][1], layers[1][2], engine="ffma", in_slope=0.1)
    assert float((out - want).abs().max()) <= TOL["float32"]
    assert float(want.abs().max()) > 0.1


def test_noalloc_rejects_layers_beyond_the_static_scratch():
    """The size check comes before any CUDA call, so it also runs on a box without a GPU."""
    from neuraltexttospeech_b200 import _lib
    lib = _lib.load()
    cin = cout = 512
    taps = (C.c_int * 7)(*range(-3, 4))
    assert cout * 4 + 7 * cin * cout * 4 > SCRATCH_BYTES
    w = np.zeros(1, np.float32)
    err = C.create_string_buffer(512)
    dummy = C.c_void_p(256)                                  # never dereferenced: the call fails on the size
    rc = lib.hfg_op_tapped_conv_noalloc(0, dummy, 1, 64, cin, w.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p), cout, 7,
                                        taps, 1, 0, 1.0, None, 1.0, dummy, None, err, len(err))
    assert rc == -4                                          # HFG_ERR_NOMEM
    msg = err.value.decode()
    assert "HFG_NOALLOC_SCRATCH_BYTES" in msg and str(SCRATCH_BYTES) in msg
    # unsupported channel counts and bad arguments say so too
    rc = lib.hfg_op_tapped_conv_noalloc(0, dummy, 1, 64, 30, w.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p), 32, 7,
                                        taps, 1, 0, 1.0, None, 1.0, dummy, None, err, len(err))
    assert rc == -5 and "c_in" in err.value.decode()
    rc = lib.hfg_op_tapped_conv_noalloc(0, None, 1, 64, 32, w.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p), 32, 7,
                                        taps, 1, 0, 1.0, None, 1.0, dummy, None, None, 0)
    assert rc == -1
