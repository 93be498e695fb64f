from .mel import log_mel_spectrogram  # noqa: F401
from .vocos import ISTFTHead  # noqa: F401
