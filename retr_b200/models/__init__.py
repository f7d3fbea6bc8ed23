from . import backbone, caption, position_encoding, transformer, utils  # noqa: F401
