"""Timings of the SURVEY 8(f) rows N2 / N3 / N4 at clinical sizes (not the bench contract):
N2 off-axis view reslice of a 512x512x300 int16 CT; N3 dose (2.5 mm grid) -> CT lattice under an ROI mask + DVH
statistics; N4 MetaImage / npy payloads between a tmpfs file and HBM."""
import json
import os
import statistics
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medicalimageanalysis_b200 import Data, geometry as G, kernels      # noqa: E402
from medicalimageanalysis_b200.read import read_mhd, write_mhd            # noqa: E402
from medicalimageanalysis_b200.structure import Dose                      # noqa: E402
from medicalimageanalysis_b200.utils import CreateImageFromMask           # noqa: E402

dev = torch.device("cuda")
PEAK = 6547.5


def timed(fn, n=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e3


class _Src(object):
    pass


res = {}
shape, sp = (300, 512, 512), (0.9765625, 0.9765625, 2.5)
z, y, x = (torch.arange(n, device=dev, dtype=torch.float32) for n in shape)
ct = (600 * torch.sin(x[None, None, :] / 17) * torch.cos(y[None, :, None] / 23) * torch.sin(z[:, None, None] / 19)).to(torch.int16)
img = CreateImageFromMask(ct, (-250.0, -250.0, -375.0), sp, "CT 01").add_image()

# N2: interactive view rotation (whole-volume off-axis reslice)
us = timed(lambda: img.update_rotation(3, -2, 4))
nvox = img.display.secondary_array.numel()
res["n2_offaxis_update_rotation"] = {"us": round(us, 1), "out_shape": list(img.display.secondary_array.shape),
                                     "gvoxel_s": round(nvox / us / 1e3, 1)}

for ang in (30, 60, 90):                                   # beyond the standard box: the tall-box family of the store kernel
    us = timed(lambda: img.update_rotation(0, 0, ang))
    res["n2_offaxis_update_rotation_z%d" % ang] = {"us": round(us, 1), "gvoxel_s": round(img.display.secondary_array.numel() / us / 1e3, 1)}
img.reset_array()

# N3: dose on the CT lattice under a mask
dshape, dsp = (150, 200, 200), (2.5, 2.5, 5.0)
dz, dy, dx = (torch.linspace(-1, 1, n, device=dev) for n in dshape)
dose = (70 * torch.exp(-2.5 * (dx[None, None, :] ** 2 + 1.4 * dy[None, :, None] ** 2 + 0.8 * dz[:, None, None] ** 2))).float()
src = _Src()
src.array, src.dose_name, src.spacing, src.origin = dose, "Dose 01", dsp, (-250.0, -250.0, -375.0)
src.image_matrix, src.dimensions = np.identity(3, dtype=np.float32), dshape
d = Dose(src)
zz, yy, xx = torch.meshgrid(z, y, x, indexing="ij")
mask = ((((xx - 256) / 60) ** 2 + ((yy - 256) / 45) ** 2 + ((zz - 150) / 40) ** 2) < 1).to(torch.uint8)
del zz, yy, xx
A = G.compose_index_matrix(img.grid, d.grid, np.identity(4), convention="itk")
n_roi = int(mask.sum())
us_sample = timed(lambda: kernels.resample_masked(d.array, A, mask))
sample = kernels.resample_masked(d.array, A, mask)
us_stats = timed(lambda: kernels.sample_statistics(sample, percentiles=[100 - q for q in range(1, 100, 4)], levels=list(range(0, 155, 5))))
us_all = timed(lambda: d.compute_roi_dose_statistics(img, mask))
us_full = timed(lambda: kernels.resample(d.array, A, shape, background=0.0, mode="itk")[mask > 0])
res["n3_dose_dvh"] = {"roi_voxels": n_roi, "roi_fraction": round(n_roi / mask.numel(), 4),
                      "masked_sample_us": round(us_sample, 1), "statistics_us": round(us_stats, 1),
                      "compute_roi_dose_statistics_us": round(us_all, 1),
                      "full_resample_then_mask_us": round(us_full, 1),
                      "mask_bytes_gbs": round(2 * mask.numel() / us_sample / 1e3, 1)}

# N4: file <-> device
tmp = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
p = os.path.join(tmp, "mia_perf_vol.mhd")
t0 = time.perf_counter(); write_mhd(p, ct, spacing=sp); t_w = time.perf_counter() - t0
ts = []
for _ in range(5):
    t0 = time.perf_counter(); back, _h = read_mhd(p, device=dev); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
assert torch.equal(back, ct)
t0 = time.perf_counter(); ref = torch.from_numpy(np.fromfile(p[:-4] + ".raw", dtype=np.int16).reshape(shape)).to(dev); torch.cuda.synchronize()
t_np = time.perf_counter() - t0
nbytes = ct.numel() * 2
res["n4_mhd_io"] = {"bytes": nbytes, "read_to_device_gbs": round(nbytes / min(ts) / 1e9, 2), "write_from_device_gbs": round(nbytes / t_w / 1e9, 2),
                    "numpy_fromfile_then_to_device_gbs": round(nbytes / t_np / 1e9, 2), "where": tmp}
for f in (p, p[:-4] + ".raw"):
    os.remove(f)
print(json.dumps(res))
