"""GPU parity tests of the steps either side of the generator, through the C ABI:
16-bit PCM output fused into the conv_post epilogue, and the mel front-end kernel."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
from oracle import adjacent_oracle as A
from oracle import hifigan_oracle as O


@pytest.mark.parametrize("v", "123")
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_pcm16_is_the_quantised_fp32_output_bit_exact(v, precision):
    h = HParams.builtin(v)
    g = Generator(h, precision=precision).load_weights(synthetic_weights(h, "B"))
    mel = synthetic_mel(3, 19, h.num_mels)
    audio = g(mel)
    pcm = g.synthesize_pcm16(mel)
    assert pcm.dtype == np.int16 and pcm.shape == audio.shape[:2]
    assert (pcm == A.pcm16(audio[..., 0])).all()               # same kernel, same tanh: integer work is bit-exact
    # and within the path's waveform tolerance of the quantised oracle
    ref = A.pcm16(O.generator_forward(h, synthetic_weights(h, "B"), mel, "torch32")[..., 0])
    tol = 1 + int(32768 * (1e-5 if precision == "fp32" else 2e-3))
    assert np.abs(pcm.astype(np.int32) - ref.astype(np.int32)).max() <= tol
    g.close()


def test_pcm16_saturates_and_device_path_matches_host_path():
    import torch
    h = HParams.builtin("2")
    w = synthetic_weights(h, "B")
    w["conv_post.b"] = np.array([30.0], np.float32)            # tanh -> exactly 1.0f everywhere: must give 32767, not wrap
    g = Generator(h, precision="fp16").load_weights(w)
    mel = synthetic_mel(2, 8, h.num_mels)
    pcm = g.synthesize_pcm16(mel)
    assert (pcm == 32767).all()
    w["conv_post.b"] = np.array([-30.0], np.float32)
    g.load_weights(w)
    assert (g.synthesize_pcm16(mel) == -32768).all()
    w["conv_post.b"] = np.array([0.01], np.float32)
    g.load_weights(w)
    host = g.synthesize_pcm16(mel)
    dev = g.synthesize_pcm16(torch.from_numpy(mel).cuda())
    assert dev.is_cuda and dev.dtype == torch.int16
    assert (dev.cpu().numpy() == host).all()
    assert g.synthesize_pcm16(np.zeros((0, 8, 80), np.float32)).shape == (0, 8 * 256)
    g.close()
