"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol the
header declares, and the host logic (config marshalling, padding rule, topology counts) is right.
No compute calls are made here (no GPU in this container)."""
import os
import re

import numpy as np
import pytest

import __graft_entry__ as entry
from neuraltexttospeech_b200 import HParams, _lib
from neuraltexttospeech_b200.generator import _config_from_hparams
from neuraltexttospeech_b200.topology import (boundary_bytes_per_frame, count_macs_per_frame, count_params,
                                              enumerate_layers)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    entry.build()
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "hifigan_b200.h")).read()
    declared = set(re.findall(r"\b(hfg_[a-z0-9_]+)\s*\(", header))
    declared -= {"hfg_ctx"}
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    assert lib.hfg_version() >= 1000


def test_sass_contains_blackwell_tensor_core_and_bulk_copy_ops():
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", os.path.join(ROOT, "build", "conv_tc.o")],
                          capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass          # tcgen05.mma
    assert "LDTM" in sass             # tcgen05.ld
    assert "UBLKCP" in sass           # cp.async.bulk (weight stages)
    assert "UTMALDG" in sass          # cp.async.bulk.tensor loads (TMA halo slabs)
    assert "UTMASTG" in sass          # TMA tensor stores (epilogue)
    assert "USETMAXREG.DEALLOC" in sass and "USETMAXREG.TRY_ALLOC" in sass   # setmaxnreg: producer/MMA/store warps hand registers to the epilogue groups
    assert "UTCHMMA.2CTA" in sass     # cta_group::2 pair kernels of the wide stages
    assert "sm_100a" in sass or "SM100" in sass.upper()


def test_create_fails_loudly_without_gpu_or_with_bad_config(lib):
    import ctypes as C
    import torch
    cfg = _config_from_hparams(HParams.builtin("1"), "reference")
    h = C.c_void_p()
    if not torch.cuda.is_available():
        rc = lib.hfg_create(C.byref(cfg), 0, C.byref(h))
        assert rc == _lib.HFG_ERR_CUDA and not h
        assert "no CPU fallback" in _lib.last_error(None)
    bad = _config_from_hparams(HParams.builtin("1"), "reference")
    bad.resblock = 7
    rc = lib.hfg_create(C.byref(bad), 0, C.byref(h))
    assert rc == _lib.HFG_ERR_INVALID and "resblock" in _lib.last_error(None)
    bad = _config_from_hparams(HParams.builtin("1"), "reference")
    bad.resblock_kernel_sizes[0] = 2          # k=2,d=1 -> 'valid' conv: the reference's residual add breaks
    rc = lib.hfg_create(C.byref(bad), 0, C.byref(h))
    assert rc == _lib.HFG_ERR_INVALID


def test_config_marshalling():
    cfg = _config_from_hparams(HParams.builtin("3"), "reference")
    assert cfg.resblock == 2 and cfg.num_upsamples == 3 and list(cfg.upsample_rates)[:3] == [8, 8, 4]
    assert cfg.num_kernels == 3 and list(cfg.num_dilations)[:3] == [2, 2, 2]
    assert [list(r)[:2] for r in cfg.resblock_dilation_sizes][:3] == [[1, 2], [2, 6], [3, 12]]
    cfg = _config_from_hparams(HParams.builtin("1"), "upstream")
    assert cfg.resblock == 1 and cfg.compat == 1 and cfg.upsample_initial_channel == 512


def test_hparams_construction_idiom_and_dict():
    import json
    from neuraltexttospeech_b200.hparams import coerce_hparams
    path = os.path.join(ROOT, "neuraltexttospeech_b200", "configs", "config_v2.json")
    h = HParams(**json.load(open(path)))            # reference HiFiGAN_TF/train.py:44-47
    assert h.upsample_initial_channel == 128
    d = coerce_hparams({k: getattr(h, k) for k in ("resblock", "upsample_rates", "upsample_kernel_sizes",
                                                     "upsample_initial_channel", "resblock_kernel_sizes",
                                                     "resblock_dilation_sizes")})
    assert d.num_mels == 80
    with pytest.raises(ValueError):
        coerce_hparams({"resblock": "1"})
    with pytest.raises(TypeError):
        HParams(resblock="1")                        # all 25 fields are required, like the reference


def test_topology_invariants_match_survey():
    # SURVEY.md section 8a invariants
    expect = {"1": (307052544, 13926017), "2": (19255296, 925985), "3": (22482944, 1462273)}
    for v, (macs, params) in expect.items():
        h = HParams.builtin(v)
        assert count_macs_per_frame(h) == macs
        assert count_params(h) == params
    h = HParams.builtin("1")
    names = [L.name for L in enumerate_layers(h)]
    assert names[0] == "conv_pre" and names[1] == "ups.0" and names[2] == "resblocks.0.convs1.0" and names[-1] == "conv_post"
    causal = [L.name for L in enumerate_layers(h) if L.padding == "causal"]
    # reference quirk Q1: k=3,d=1 convs are causal: convs1[0] and all convs2 of the k=3 resblock of every stage
    assert len(causal) == 4 * 4 and all(".convs2." in n or n.endswith("convs1.0") for n in causal)
    assert boundary_bytes_per_frame(h, 2) * 862 * 64 / 1e9 == pytest.approx(138.2, rel=0.01)


def test_generator_requires_cuda_and_never_computes_on_cpu():
    import torch
    from neuraltexttospeech_b200 import Generator
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Generator(HParams.builtin("1"))
