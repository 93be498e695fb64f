from .istftnet import Decoder, build_decoder  # noqa: F401
