from .image import Image
from .rigid import Rigid, Display

__all__ = ["Image", "Rigid", "Display"]
