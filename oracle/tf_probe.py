"""Run-time probe for the REAL reference (TensorFlow 2 + tensorflow-addons) -- TEST INFRASTRUCTURE ONLY.

SURVEY.md section 8c: the reference's arithmetic lives in tensorflow / keras / tensorflow-addons, none of which is
installable in the build container or on the GPU boxes, which is why the oracle's parity is "unpinned" at the TF
boundary.  The day a box does carry them (a driver-provided install under baseline/_ref, or HFG_TF_REFERENCE_DIR
pointing at a checkout of the reference's HiFiGAN_TF directory), `tf_generator_forward` builds the reference's own
`Generator` (HiFiGAN_TF/model.py:115-173), loads the named weights into its WeightNormalization wrappers by attribute
(`layer.v`, `layer.g`, `layer.layer.bias`; never by get_weights() order) and runs it on the CPU, so that
`__graft_entry__.smoke()` and `bench.py --impl reference` pin the oracle against it by themselves and record
kind "tf".  Without TensorFlow every entry point returns None and the callers say "tf unavailable".
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_dir():
    cands = [os.environ.get("HFG_TF_REFERENCE_DIR"), os.path.join(ROOT, "baseline", "_ref", "HiFiGAN_TF"),
             os.path.join(ROOT, "baseline", "_ref")]
    for c in cands:
        if c and os.path.isfile(os.path.join(c, "model.py")) and os.path.isfile(os.path.join(c, "hparams.py")):
            return c
    return None


def available():
    """(ok, why): ok only when tensorflow, tensorflow_addons AND the reference's model.py are all reachable."""
    try:
        import tensorflow  # noqa: F401
        import tensorflow_addons  # noqa: F401
    except Exception as e:                       # ImportError, or a broken install
        return False, f"tensorflow / tensorflow_addons not importable ({type(e).__name__})"
    d = _reference_dir()
    if d is None:
        return False, "tensorflow is importable but no HiFiGAN_TF/model.py under baseline/_ref or HFG_TF_REFERENCE_DIR"
    return True, d


def _build(hparams, weights, num_mels):
    import tensorflow as tf
    d = _reference_dir()
    if d not in sys.path:
        sys.path.insert(0, d)
    import importlib
    model = importlib.import_module("model")                  # the reference's HiFiGAN_TF/model.py, unmodified
    gen = model.Generator(hparams)
    gen(tf.zeros((1, 4, num_mels), tf.float32))               # lazy Keras build; sets tfa's `initialized`

    def put(wrapper, name):                                   # tfa.layers.WeightNormalization: .v, .g, .layer.bias
        wrapper.v.assign(np.asarray(weights[f"{name}.v"], np.float32))
        wrapper.g.assign(np.asarray(weights[f"{name}.g"], np.float32).reshape(wrapper.g.shape))
        wrapper.layer.bias.assign(np.asarray(weights[f"{name}.b"], np.float32))

    put(gen.conv_pre, "conv_pre")
    put(gen.conv_post, "conv_post")
    for i, up in enumerate(gen.ups):
        put(up, f"ups.{i}")
    for n, rb in enumerate(gen.resblocks):
        if hasattr(rb, "convs1"):
            for m in range(len(rb.convs1)):
                put(rb.convs1[m], f"resblocks.{n}.convs1.{m}")
                put(rb.convs2[m], f"resblocks.{n}.convs2.{m}")
        else:
            for m in range(len(rb.convs)):
                put(rb.convs[m], f"resblocks.{n}.convs.{m}")
    return gen


def tf_generator_forward(hparams, weights, mel):
    """The reference's TF2 generator on the CPU: mel [B,T,num_mels] -> audio [B, T*prod(rates), 1], or None when
    TensorFlow / the reference checkout is not reachable."""
    ok, _ = available()
    if not ok:
        return None
    import tensorflow as tf
    with tf.device("/CPU:0"):
        gen = _build(hparams, weights, int(np.asarray(mel).shape[-1]))
        return np.asarray(gen(tf.constant(np.asarray(mel, np.float32))))


def pin_oracle(hparams, weights, mel, oracle_out):
    """-> dict for logs / bench lines: {"kind": "tf", "oracle_vs_tf_max_abs": ...} or {"kind": "port", "tf": why}."""
    ok, why = available()
    if not ok:
        return {"kind": "port", "tf": "unavailable: " + why}
    y = tf_generator_forward(hparams, weights, mel)
    return {"kind": "tf", "oracle_vs_tf_max_abs": float(np.abs(np.asarray(oracle_out) - y).max())}
