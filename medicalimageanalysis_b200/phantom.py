"""Deterministic synthetic CT / MR phantoms.  There is no network and no DICOM data in the
image, so every test and benchmark volume comes from here: air outside a body ellipse, soft tissue with
a smooth low-frequency field and noise, bone and lung ellipsoids, a couch line.  ``world_matrix`` lets
the same anatomy be generated under a ground-truth rigid offset (the "moving" image of a pair)."""
from __future__ import annotations

import numpy as np

__all__ = ["make_phantom"]


def _structures(rng):
    bones = [(rng.uniform(-0.25, 0.25, 3), rng.uniform(0.03, 0.09, 3), rng.uniform(700, 1400)) for _ in range(6)]
    lungs = [(rng.uniform(-0.22, 0.22, 3), rng.uniform(0.06, 0.14, 3)) for _ in range(4)]
    waves = [(rng.uniform(0.5, 2.5, 3), rng.uniform(0, 2 * np.pi)) for _ in range(3)]
    return bones, lungs, waves


def make_phantom(shape, modality="CT", seed=0, spacing=(1.0, 1.0, 1.0), world_matrix=None,
                 dtype=None, noise=True):
    """Return a ``[z, y, x]`` volume.  CT: int16 in [-1024, 3071]; MR: float32 in [0, 4095].

    ``world_matrix`` (4x4, mm) maps this image's physical coordinates (relative to the volume centre)
    to anatomy coordinates; identity = the canonical pose."""
    nz, ny, nx = (int(s) for s in shape)
    rng = np.random.default_rng(seed)
    bones, lungs, waves = _structures(rng)
    noise_rng = np.random.default_rng(seed + 7919)
    sp = np.asarray(spacing, dtype=np.float64)
    ext = np.array([nx, ny, nz]) * sp                      # physical size, mm
    W = np.identity(4) if world_matrix is None else np.asarray(world_matrix, dtype=np.float64)
    if dtype is None:
        dtype = np.int16 if modality == "CT" else np.float32
    out = np.empty((nz, ny, nx), dtype=dtype)
    x = ((np.arange(nx) + 0.5) * sp[0] - ext[0] / 2)[None, None, :]
    y = ((np.arange(ny) + 0.5) * sp[1] - ext[1] / 2)[None, :, None]
    zc = 16
    for z0 in range(0, nz, zc):
        z1 = min(nz, z0 + zc)
        z = ((np.arange(z0, z1) + 0.5) * sp[2] - ext[2] / 2)[:, None, None]
        # anatomy coordinates, normalised by the physical extent so every shape looks alike
        ax = (W[0, 0] * x + W[0, 1] * y + W[0, 2] * z + W[0, 3]) / ext[0]
        ay = (W[1, 0] * x + W[1, 1] * y + W[1, 2] * z + W[1, 3]) / ext[1]
        az = (W[2, 0] * x + W[2, 1] * y + W[2, 2] * z + W[2, 3]) / ext[2]
        ax, ay, az = (np.broadcast_to(a, (z1 - z0, ny, nx)).astype(np.float32) for a in (ax, ay, az))
        body = (ax / 0.42) ** 2 + (ay / 0.32) ** 2 <= 1.0
        field = np.zeros_like(ax)
        for k, ph in waves:
            field += np.cos(2 * np.pi * (k[0] * ax + k[1] * ay + k[2] * az) + ph).astype(np.float32)
        field *= 10.0
        if modality == "CT":
            v = np.where(body, 40.0 + field, -1000.0).astype(np.float32)
            for c, r, val in bones:
                e = ((ax - c[0]) / r[0]) ** 2 + ((ay - c[1]) / r[1]) ** 2 + ((az - c[2]) / (r[2] * 3)) ** 2 <= 1.0
                v = np.where(e & body, np.float32(val), v)
            for c, r in lungs:
                e = ((ax - c[0]) / r[0]) ** 2 + ((ay - c[1]) / r[1]) ** 2 + ((az - c[2]) / (r[2] * 3)) ** 2 <= 1.0
                v = np.where(e & body, np.float32(-800.0), v)
            couch = (np.abs(ay - 0.40) < 0.008) & (np.abs(ax) < 0.45)
            v = np.where(couch, np.float32(200.0), v)
            if noise:
                v = v + noise_rng.normal(0.0, 15.0, v.shape).astype(np.float32)
            v = np.clip(np.rint(v), -1024, 3071)
        else:  # MR-like: non-monotone remap of the same anatomy
            v = np.where(body, 450.0 + 4.0 * field, 0.0).astype(np.float32)
            fat = body & ((ax / 0.42) ** 2 + (ay / 0.32) ** 2 > 0.85)
            v = np.where(fat, np.float32(900.0), v)
            for c, r, val in bones:
                e = ((ax - c[0]) / r[0]) ** 2 + ((ay - c[1]) / r[1]) ** 2 + ((az - c[2]) / (r[2] * 3)) ** 2 <= 1.0
                v = np.where(e & body, np.float32(80.0), v)
            for c, r in lungs:
                e = ((ax - c[0]) / r[0]) ** 2 + ((ay - c[1]) / r[1]) ** 2 + ((az - c[2]) / (r[2] * 3)) ** 2 <= 1.0
                v = np.where(e & body, np.float32(30.0), v)
            if noise:
                n1 = noise_rng.normal(0.0, 12.0, v.shape).astype(np.float32)
                n2 = noise_rng.normal(0.0, 12.0, v.shape).astype(np.float32)
                v = np.sqrt((v + n1) ** 2 + n2 ** 2)        # Rician-ish
            v = np.clip(v, 0, 4095)
        out[z0:z1] = v.astype(dtype)
    return out


def make_phantom_device(shape, modality="CT", seed=0, spacing=(1.0, 1.0, 1.0), world_matrix=None, device="cuda",
                        z_range=None, noise=True):
    """``make_phantom`` on the device, slab by slab (config 5 is 2.5 GB -- building it with NumPy on the host
    takes minutes).  Same anatomy for the same ``seed`` (the structure list comes from the same NumPy generator); the
    noise comes from a device generator seeded per 16-plane chunk, so a z-slab (``z_range=(z0, z1)``, multiples of 16)
    generated on its own equals the same planes of the whole volume.  Returns a ``[z, y, x]`` torch tensor, int16 (CT)
    or float32 (MR); values differ from the host phantom's only through the noise stream."""
    import torch
    nz, ny, nx = (int(s) for s in shape)
    z_lo, z_hi = (0, nz) if z_range is None else (int(z_range[0]), int(z_range[1]))
    if z_lo % 16 != 0:
        raise ValueError("z_range must start on a multiple of 16 (the noise is seeded per 16-plane chunk)")
    rng = np.random.default_rng(seed)
    bones, lungs, waves = _structures(rng)
    sp = np.asarray(spacing, dtype=np.float64)
    ext = np.array([nx, ny, nz]) * sp
    W = np.identity(4) if world_matrix is None else np.asarray(world_matrix, dtype=np.float64)
    dev = torch.device(device)
    dtype = torch.int16 if modality == "CT" else torch.float32
    out = torch.empty((z_hi - z_lo, ny, nx), dtype=dtype, device=dev)
    f32 = dict(dtype=torch.float32, device=dev)
    x = (((torch.arange(nx, **f32) + 0.5) * float(sp[0])) - float(ext[0]) / 2)[None, None, :]
    y = (((torch.arange(ny, **f32) + 0.5) * float(sp[1])) - float(ext[1]) / 2)[None, :, None]
    gen = torch.Generator(device=dev)

    def ellipsoid(ax, ay, az, c, r):
        return ((ax - float(c[0])) / float(r[0])) ** 2 + ((ay - float(c[1])) / float(r[1])) ** 2 + \
               ((az - float(c[2])) / (float(r[2]) * 3)) ** 2 <= 1.0

    for z0 in range(z_lo, z_hi, 16):
        z1 = min(z_hi, z0 + 16)
        z = (((torch.arange(z0, z1, **f32) + 0.5) * float(sp[2])) - float(ext[2]) / 2)[:, None, None]
        ax = (float(W[0, 0]) * x + float(W[0, 1]) * y + float(W[0, 2]) * z + float(W[0, 3])) / float(ext[0])
        ay = (float(W[1, 0]) * x + float(W[1, 1]) * y + float(W[1, 2]) * z + float(W[1, 3])) / float(ext[1])
        az = (float(W[2, 0]) * x + float(W[2, 1]) * y + float(W[2, 2]) * z + float(W[2, 3])) / float(ext[2])
        body = (ax / 0.42) ** 2 + (ay / 0.32) ** 2 <= 1.0
        field = torch.zeros((z1 - z0, ny, nx), **f32)
        for k, ph in waves:
            field += torch.cos(2 * np.pi * (float(k[0]) * ax + float(k[1]) * ay + float(k[2]) * az) + float(ph))
        field *= 10.0
        gen.manual_seed(int(seed) * 1000003 + 7919 + z0)
        if modality == "CT":
            v = torch.where(body, 40.0 + field, torch.full_like(field, -1000.0))
            for c, r, val in bones:
                v = torch.where(ellipsoid(ax, ay, az, c, r) & body, torch.full_like(v, float(val)), v)
            for c, r in lungs:
                v = torch.where(ellipsoid(ax, ay, az, c, r) & body, torch.full_like(v, -800.0), v)
            couch = ((ay - 0.40).abs() < 0.008) & (ax.abs() < 0.45)
            v = torch.where(couch, torch.full_like(v, 200.0), v)
            if noise:
                v = v + 15.0 * torch.randn(v.shape, generator=gen, **f32)
            out[z0 - z_lo:z1 - z_lo] = torch.clamp(torch.round(v), -1024, 3071).to(torch.int16)
        else:
            v = torch.where(body, 450.0 + 4.0 * field, torch.zeros_like(field))
            fat = body & ((ax / 0.42) ** 2 + (ay / 0.32) ** 2 > 0.85)
            v = torch.where(fat, torch.full_like(v, 900.0), v)
            for c, r, val in bones:
                v = torch.where(ellipsoid(ax, ay, az, c, r) & body, torch.full_like(v, 80.0), v)
            for c, r in lungs:
                v = torch.where(ellipsoid(ax, ay, az, c, r) & body, torch.full_like(v, 30.0), v)
            if noise:
                n1 = 12.0 * torch.randn(v.shape, generator=gen, **f32)
                n2 = 12.0 * torch.randn(v.shape, generator=gen, **f32)
                v = torch.sqrt((v + n1) ** 2 + n2 ** 2)
            out[z0 - z_lo:z1 - z_lo] = torch.clamp(v, 0, 4095)
    return out
