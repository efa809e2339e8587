"""B200-native HiFi-GAN generator forward pass (drop-in for the reference's
HiFiGAN_TF/model.py `Generator`).  See DESIGN.md."""
from .hparams import HParams, coerce_hparams          # noqa: F401
from .utils import get_padding, LRELU_SLOPE           # noqa: F401
from .generator import Generator                      # noqa: F401
from .audio_processing import STFT                    # noqa: F401
from .pcm import write_wav, write_wavs                # noqa: F401
from .discriminators import MultiPeriodDiscriminator, MultiScaleDiscriminator   # noqa: F401

__all__ = ["Generator", "HParams", "STFT", "MultiPeriodDiscriminator", "MultiScaleDiscriminator", "get_padding", "LRELU_SLOPE", "write_wav", "write_wavs"]
