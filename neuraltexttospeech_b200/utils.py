"""Small helpers whose *semantics* the generator must reproduce.

``get_padding`` restates reference HiFiGAN_TF/utils.py:29-37: the integer pad
``(k*d - d) // 2`` is mapped onto a Keras padding-mode **string**; pad 1 (only
k=3, d=1 in the stock configs) therefore becomes a *causal* convolution, which is
a deliberate quirk of the reference that this drop-in keeps.
"""

LRELU_SLOPE = 0.1          # reference HiFiGAN_TF/model.py:14
FINAL_LRELU_SLOPE = 0.3    # keras LeakyReLU() default, reference HiFiGAN_TF/model.py:152


def get_padding(kernel_size: int, dilation: int = 1) -> str:
    pad = int((kernel_size * dilation - dilation) / 2)
    return {0: "valid", 1: "causal"}.get(pad, "same")


def conv_pads(kernel_size: int, dilation: int, mode: str):
    """(left, right) zero padding in samples for a stride-1 Conv1D in `mode`."""
    span = dilation * (kernel_size - 1)
    if mode == "valid":
        return 0, 0
    if mode == "causal":
        return span, 0
    if mode == "same":
        # TF SAME for stride 1: total = span, left = span // 2
        return span // 2, span - span // 2
    raise ValueError(mode)
