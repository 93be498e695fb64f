from .istftnet import Decoder, build_decoder  # noqa: F401
from .kokoro import Model, ModelConfig, SyntheticFrontend  # noqa: F401
from .pipeline import KokoroPipeline  # noqa: F401
