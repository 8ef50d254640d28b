"""Shared helpers of the test-suite: small seeded configs and input batches."""
import ctypes as C

import numpy as np

from pyvspg import _abi as A
from pyvspg import configs


def small(name, **kw):
    """The BASELINE configs at sizes the CPU oracle finishes in seconds."""
    if name == "C1":
        return configs.config1(res=kw.get("res", 48), spp=kw.get("spp", 4))
    if name == "C1G":   # C1 under pbrt's default pixel filter: gaussian, radius 1.5, sigma 0.5 (filters.cpp GaussianFilter::Create)
        cfg = configs.config1(res=kw.get("res", 48), spp=kw.get("spp", 4))
        for cam in (cfg.scene.desc.camera, cfg.scene.camera):
            cam.filter_type, cam.filter_sigma = 1, 0.5
            cam.filter_radius[0] = cam.filter_radius[1] = 1.5
        cfg.name = "C1-gaussian-filter"
        return cfg
    if name == "C1S":
        return configs.config1_spheres(res=kw.get("res", 48), spp=kw.get("spp", 4))
    if name == "C1E":
        return configs.config1_emitters(res=kw.get("res", 48), spp=kw.get("spp", 4))
    if name == "C1T":
        return configs.config1_triangles(res=kw.get("res", 48), spp=kw.get("spp", 4))
    common = dict(grid_n=kw.get("grid_n", 32), vol_leaves=kw.get("vol_leaves", 256), surf_leaves=kw.get("surf_leaves", 128))
    if name == "C2":
        return configs.config2(res=kw.get("res", 48), spp=kw.get("spp", 4), **common)
    if name == "C2K":   # C2 with reflectance-0 patches among the lit surfaces: vertices that take no NEE share warps with ones that do
        return configs.config2(res=kw.get("res", 48), spp=kw.get("spp", 4), black_patches=True, **common)
    if name == "C3":
        return configs.config3(res=kw.get("res", 48), spp=kw.get("spp", 4), **common)
    if name == "C4":
        return configs.config4(width=kw.get("res", 64), height=kw.get("res", 64) * 9 // 16, spp=kw.get("spp", 2), **common)
    if name == "C5":
        common["grid_n"] = kw.get("grid_n", 24)
        return configs.config5(width=kw.get("res", 64), height=kw.get("res", 64) * 9 // 16, spp=kw.get("spp", 2), **common)
    raise KeyError(name)


def random_rays(rng, n, eye, lo=(-0.3, -0.3, -0.3), hi=(1.3, 1.3, 1.3), tmax_range=(0.2, 4.0)):
    """n x 7 (o, d normalised, tMax) in render space aimed at the unit medium box (world [0,1]^3)."""
    o = rng.uniform(lo, hi, size=(n, 3)) - eye
    target = rng.uniform(0.0, 1.0, size=(n, 3)) - eye
    d = target - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    # a few axis-aligned / zero-component directions: the DDA's inf / -0 handling (media.h:158-160, 186-190)
    for i in range(0, min(n, 24)):
        k = i % 3
        d[i] = 0
        d[i, k] = 1.0 if (i // 3) % 2 == 0 else -1.0
        if i >= 12:
            d[i, (k + 1) % 3] = -0.0
    d32 = d.astype(np.float32)
    d32 /= np.sqrt((d32.astype(np.float64) ** 2).sum(1, keepdims=True)).astype(np.float32)
    t = rng.uniform(*tmax_range, size=(n, 1))
    return np.concatenate([o, d32, t], axis=1).astype(np.float32)


def pcg_state(rng):
    return int(rng.integers(0, 2 ** 63)), int(rng.integers(0, 2 ** 62)) * 2 + 1


def struct_array_to_np(arr, field, n=None):
    out = [getattr(x, field) for x in arr]
    if hasattr(out[0], "__len__"):
        return np.array([list(v) for v in out], dtype=np.float64)
    return np.array(out)


def rel_err(a, b, floor=1e-6):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)
