"""medicalimageanalysis_b200 -- B200-native (sm_100a) rigid resample + SSD / joint-histogram-MI hot path behind
the MedicalImageAnalysis (``mia``) API.  See DESIGN.md for scope and INTEGRATION.md for the drop-in boundary.

Mirrors the reference's top-level names that sit on this path (``medicalimageanalysis/__init__.py:2-10``):
``Data``, ``Rigid``, ``utils``; everything else of the reference is out of scope.
"""
from .data import Data
from .structure.image import Image
from .structure.rigid import Rigid
from . import utils
from . import read
from . import geometry
from .registration import RegistrationConfig, RegistrationResult

__version__ = "0.1.0"
__all__ = ["Data", "Image", "Rigid", "utils", "read", "geometry", "RegistrationConfig", "RegistrationResult"]
