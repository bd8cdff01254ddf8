from .mhd import MhdReader, read_mhd_header, read_mhd, write_mhd
from .npy import load_npy_to_device, save_npy_from_device

__all__ = ["MhdReader", "read_mhd_header", "read_mhd", "write_mhd", "load_npy_to_device", "save_npy_from_device"]
