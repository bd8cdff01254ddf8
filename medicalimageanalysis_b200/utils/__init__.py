from . import rigid
from .creation import CreateImageFromMask

__all__ = ["rigid", "CreateImageFromMask"]
