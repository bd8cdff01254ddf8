from . import rigid
from . import deformable
from .creation import CreateImageFromMask

__all__ = ["rigid", "deformable", "CreateImageFromMask"]
