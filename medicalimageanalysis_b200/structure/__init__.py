from .image import Image
from .dose import Dose
from .display import VolumeDisplay
from .rigid import Rigid, Display

__all__ = ["Image", "Dose", "VolumeDisplay", "Rigid", "Display"]
