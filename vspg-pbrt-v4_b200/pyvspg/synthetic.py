"""Seeded synthetic inputs: fBm density grids, a synthetic OpenPGL-0.7.0-shaped guiding cache (KD tree
of vMF mixtures + 8-NN neighbour tables), image-space VSP buffer and transmittance buffer.  Host-side data generation
only (numpy); none of this is on the measured path.
"""
import ctypes as C

import numpy as np

from . import _abi as A


# --------------------------------------------------------------------------- procedural density
def _hash3(ix, iy, iz, seed):
    h = (ix.astype(np.uint64) * np.uint64(0x9E3779B185EBCA87)) ^ (iy.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)) ^ \
        (iz.astype(np.uint64) * np.uint64(0x165667B19E3779F9)) ^ np.uint64(seed)
    h ^= h >> np.uint64(33)
    h *= np.uint64(0xFF51AFD7ED558CCD)
    h ^= h >> np.uint64(33)
    return ((h >> np.uint64(40)).astype(np.float32)) * np.float32(1.0 / (1 << 24))


def _value_noise(x, y, z, seed):
    ix, iy, iz = np.floor(x).astype(np.int64), np.floor(y).astype(np.int64), np.floor(z).astype(np.int64)
    fx, fy, fz = (x - ix).astype(np.float32), (y - iy).astype(np.float32), (z - iz).astype(np.float32)
    sx, sy, sz = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy), fz * fz * (3 - 2 * fz)
    out = np.zeros_like(fx)
    for dz in (0, 1):
        for dy in (0, 1):
            for dx in (0, 1):
                w = (sx if dx else 1 - sx) * (sy if dy else 1 - sy) * (sz if dz else 1 - sz)
                out += w * _hash3(ix + dx, iy + dy, iz + dz, seed)
    return out


def _fbm_torch(n, seed, octaves, base_freq, gain, lacunarity, scale_, offset, slab=32):
    """The same arithmetic as the numpy path below, op for op, on torch tensors (CUDA when present): the 512^3 grids
    of C4 take minutes in single-threaded numpy and about a second here."""
    import torch
    dev = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
    f32, i64 = torch.float32, torch.int64

    def u64(v):   # two's-complement int64 with the same bits as the uint64 constant
        return v - (1 << 64) if v >= (1 << 63) else v

    def lsr(h, k):   # logical shift right on int64
        return (h >> k) & ((1 << (64 - k)) - 1)

    def hash3(ix, iy, iz, sd):
        h = (ix * u64(0x9E3779B185EBCA87)) ^ (iy * u64(0xC2B2AE3D27D4EB4F)) ^ (iz * u64(0x165667B19E3779F9)) ^ sd
        h = h ^ lsr(h, 33)
        h = h * u64(0xFF51AFD7ED558CCD)
        h = h ^ lsr(h, 33)
        return lsr(h, 40).to(f32) * (1.0 / (1 << 24))

    def value_noise(x, y, z, sd):
        fl = [torch.floor(v) for v in (x, y, z)]
        ix, iy, iz = (v.to(i64) for v in fl)
        fx, fy, fz = x - fl[0], y - fl[1], z - fl[2]
        sx, sy, sz = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy), fz * fz * (3 - 2 * fz)
        out = torch.zeros_like(fx)
        for dz in (0, 1):
            for dy in (0, 1):
                for dx in (0, 1):
                    w = (sx if dx else 1 - sx) * (sy if dy else 1 - sy) * (sz if dz else 1 - sz)
                    out += w * hash3(ix + dx, iy + dy, iz + dz, sd)
        return out

    out = np.empty((n, n, n), dtype=np.float32)
    xs = (torch.arange(n, dtype=f32, device=dev) + 0.5) / n
    norm = sum(gain ** o for o in range(octaves))
    for z0 in range(0, n, slab):
        zs = (torch.arange(z0, min(n, z0 + slab), dtype=f32, device=dev) + 0.5) / n
        Z, Y, X = torch.meshgrid(zs, xs, xs, indexing="ij")
        acc = torch.zeros_like(X)
        amp, freq = 1.0, base_freq
        for o in range(octaves):
            acc += float(np.float32(amp)) * value_noise(X * freq, Y * freq, Z * freq, seed + 101 * o)
            amp *= gain
            freq *= lacunarity
        d = torch.clamp(acc / float(np.float32(norm)) * float(np.float32(scale_)) + float(np.float32(offset)), 0, 1)
        out[z0:z0 + len(zs)] = d.cpu().numpy()
    return out


def fbm_grid(n, seed=1337, octaves=5, base_freq=4.0, gain=0.5, lacunarity=2.0, scale_=1.6, offset=-0.3, slab=16, backend=None):
    """density[nz][ny][nx] = clamp(fBm * scale + offset, 0, 1); generated slab by slab to bound memory.
    backend: "numpy" | "torch" | None (= torch for n >= 192, where numpy takes minutes)."""
    if backend == "torch" or (backend is None and n >= 192):
        return _fbm_torch(n, seed, octaves, base_freq, gain, lacunarity, scale_, offset)
    nx = ny = nz = n
    out = np.empty((nz, ny, nx), dtype=np.float32)
    xs = (np.arange(nx, dtype=np.float32) + 0.5) / nx
    ys = (np.arange(ny, dtype=np.float32) + 0.5) / ny
    norm = sum(gain ** o for o in range(octaves))
    for z0 in range(0, nz, slab):
        zs = (np.arange(z0, min(nz, z0 + slab), dtype=np.float32) + 0.5) / nz
        Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
        acc = np.zeros_like(X)
        amp, freq = 1.0, base_freq
        for o in range(octaves):
            acc += np.float32(amp) * _value_noise(X * freq, Y * freq, Z * freq, seed + 101 * o)
            amp *= gain
            freq *= lacunarity
        out[z0:z0 + len(zs)] = np.clip(acc / np.float32(norm) * np.float32(scale_) + np.float32(offset), 0, 1)
    return out


def cloud_grid(n, seed=4242):
    """Thick cloud-like medium for C4: fBm shaped by a height profile and a smoothstep threshold."""
    d = fbm_grid(n, seed=seed, octaves=5, base_freq=3.0, scale_=1.0, offset=0.0)
    y = (np.arange(n, dtype=np.float32) + 0.5) / n
    prof = np.clip(1.0 - np.abs(y - 0.45) / 0.45, 0, 1)[None, :, None]
    t = np.clip((d * prof - 0.28) / 0.25, 0, 1)
    return (t * t * (3 - 2 * t)).astype(np.float32)


def rgb_grids(n, seed=777):
    """C5 chromatic medium: sigma_a = (0.02,0.05,0.2)*d, sigma_s = (4,3,2)*d  -> two [n][n][n][3] grids."""
    d = fbm_grid(n, seed=seed, octaves=4, base_freq=3.0, scale_=1.8, offset=-0.4)
    sa = d[..., None] * np.array([0.02, 0.05, 0.2], dtype=np.float32)
    ss = d[..., None] * np.array([4.0, 3.0, 2.0], dtype=np.float32)
    return np.ascontiguousarray(sa, dtype=np.float32), np.ascontiguousarray(ss, dtype=np.float32)


# --------------------------------------------------------------------------- guiding inputs
class Guiding:
    """Owns the arrays behind a vspg_guiding_desc."""

    def __init__(self):
        self.desc = A.GuidingDesc()
        self._keep = []
        self.fields = {}
        self.points = {}     # per field: the regions' sample means (the KNN points), for pgl_writer

    def set_field(self, which, field_struct, keep):
        self._keep.append((field_struct, keep))
        self.fields[which] = field_struct
        if isinstance(keep, tuple) and len(keep) == 4:
            self.points[which] = keep[3]
        setattr(self.desc, which, C.pointer(field_struct))

    def set_isg(self, vsp):
        vsp = np.ascontiguousarray(vsp, dtype=np.float32)
        self._keep.append(vsp)
        self.isg = vsp
        self.desc.isg_vsp = vsp.ctypes.data_as(A.PF)
        self.desc.isg_height, self.desc.isg_width = vsp.shape

    def set_tr(self, tr_rgb):
        tr = np.ascontiguousarray(tr_rgb, dtype=np.float32)
        self._keep.append(tr)
        self.tr = tr
        self.desc.tr_rgb = tr.ctypes.data_as(A.PF)
        self.desc.tr_height, self.desc.tr_width = tr.shape[:2]


def _mean_cos_from_kappa(k):
    k = np.asarray(k, dtype=np.float64)
    return np.where(k > 1e-3, 1.0 / np.tanh(np.maximum(k, 1e-3)) - 1.0 / np.maximum(k, 1e-3), 0.0)


def synthetic_field(bmin, bmax, n_leaves, rng, lobes=8, kappa_range=(1.0, 200.0), dist_range=(0.2, 2.0), max_lobes_jitter=True):
    """A valid vspg_field: balanced KD tree over [bmin,bmax] to n_leaves (power of two) leaves; one region per leaf with
    a `lobes`-component parallax-aware vMF mixture (weights Dirichlet(1), mu uniform on S^2, kappa log-uniform,
    N = kappa/(2 pi (1-e^{-2kappa})), distances/vsp weights uniform); pivot = sample mean = leaf centre; 8-NN neighbour
    tables over the region centres (self included), as Field::deserialize rebuilds them (field/Field.h:345-348)."""
    bmin, bmax = np.asarray(bmin, dtype=np.float32), np.asarray(bmax, dtype=np.float32)
    depth = int(round(np.log2(n_leaves)))
    assert 1 << depth == n_leaves
    nodes = []          # (split, dim, idx)
    centres = np.zeros((n_leaves, 3), dtype=np.float32)
    leaf_counter = [0]

    def build(lo, hi, d, me):
        if d == depth:
            li = leaf_counter[0]
            leaf_counter[0] += 1
            centres[li] = (lo + hi) * np.float32(0.5)
            nodes[me] = (0.0, 3, li)
            return
        ax = int(np.argmax(hi - lo))
        t = np.float32(rng.uniform(0.4, 0.6))
        split = np.float32(lo[ax] + (hi[ax] - lo[ax]) * t)
        left = len(nodes)
        nodes.append(None)
        nodes.append(None)
        nodes[me] = (float(split), ax, left)
        hi_l, lo_r = hi.copy(), lo.copy()
        hi_l[ax] = split
        lo_r[ax] = split
        build(lo, hi_l, d + 1, left)
        build(lo_r, hi, d + 1, left + 1)

    nodes.append(None)
    build(bmin.copy(), bmax.copy(), 0, 0)
    nd = (A.KdNode * len(nodes))()
    for i, (s, dim, idx) in enumerate(nodes):
        nd[i].split = s
        nd[i].dim_idx = (dim << 30) | idx
    regs = (A.Mixture * n_leaves)()
    ra = np.frombuffer(regs, dtype=np.float32).reshape(n_leaves, C.sizeof(A.Mixture) // 4)
    ri = np.frombuffer(regs, dtype=np.int32).reshape(n_leaves, C.sizeof(A.Mixture) // 4)
    L = A.MAX_LOBES
    for r in range(n_leaves):
        n = lobes if not max_lobes_jitter else int(rng.integers(max(1, lobes - 3), min(L, lobes + 3) + 1))
        w = rng.dirichlet(np.ones(n)).astype(np.float32)
        mu = rng.normal(size=(n, 3))
        mu = (mu / np.linalg.norm(mu, axis=1, keepdims=True)).astype(np.float32)
        kappa = np.exp(rng.uniform(np.log(kappa_range[0]), np.log(kappa_range[1]), size=n)).astype(np.float32)
        e2k = np.exp(-2.0 * kappa.astype(np.float64)).astype(np.float32)
        norm = (kappa.astype(np.float64) / (2 * np.pi * (1.0 - e2k.astype(np.float64)))).astype(np.float32)
        mc = _mean_cos_from_kappa(kappa).astype(np.float32)
        dist = rng.uniform(dist_range[0], dist_range[1], size=n).astype(np.float32)
        vspw = rng.uniform(0.05, 0.95, size=n).astype(np.float32)
        row = ra[r]
        # padding lanes: weight 0, kappa 0, mu = (0,0,1), N = 1/4pi, e^-2k = 1 (the reference's cleared lanes)
        row[0 * L:1 * L] = 0
        row[1 * L:2 * L] = 0
        row[2 * L:3 * L] = 0
        row[3 * L:4 * L] = 0
        row[4 * L:5 * L] = 1
        row[5 * L:6 * L] = np.float32(1 / (4 * np.pi))
        row[6 * L:7 * L] = 1
        row[7 * L:8 * L] = 0
        row[8 * L:9 * L] = 0
        row[9 * L:10 * L] = 0
        for j, arr in enumerate((w, kappa, mu[:, 0], mu[:, 1], mu[:, 2], norm, e2k, mc, dist, vspw)):
            row[j * L:j * L + n] = arr
        row[10 * L:10 * L + 3] = centres[r]
        ri[r, 10 * L + 3] = n
        ri[r, 10 * L + 4] = 1
    neigh = build_neighbours(centres)
    f = A.Field()
    f.usable = 1
    f.bounds_min[:] = [float(x) for x in bmin]
    f.bounds_max[:] = [float(x) for x in bmax]
    f.n_nodes, f.nodes = len(nodes), nd
    f.n_regions, f.regions = n_leaves, regs
    f.neighbours = neigh
    f.stochastic_nn, f.is_nn = 1, 0
    return f, (nd, regs, neigh, centres)


def build_neighbours(points):
    """spatial/KNN.h:343-393 buildRegionNeighbours: for each region the 8 nearest region points (self included),
    written in ascending-distance order; unused slots id = ~0."""
    from scipy.spatial import cKDTree
    pts = np.asarray(points, dtype=np.float32)
    n = len(pts)
    k = min(8, n)
    _, idx = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=k)
    idx = idx.reshape(n, k)
    out = (A.Neighbours * n)()
    for i in range(n):
        nb = out[i]
        for s in range(8):
            if s < k:
                j = int(idx[i, s])
                nb.ids[s] = j
                nb.px[s], nb.py[s], nb.pz[s] = float(pts[j, 0]), float(pts[j, 1]), float(pts[j, 2])
            else:
                nb.ids[s] = 0xFFFFFFFF
                nb.px[s] = nb.py[s] = nb.pz[s] = 0.0
        nb.size = k
    return out


def synthetic_isg(width, height, rng, invalid_fraction=0.02):
    """Image-space VSP buffer (layer VSP.Y): smooth field in [0.05, 0.95], ~2 % pixels = -1 (invalid)."""
    y, x = np.mgrid[0:height, 0:width].astype(np.float32)
    v = 0.5 + 0.45 * np.sin(x / width * 5.1 + 0.3) * np.cos(y / height * 3.7 - 0.2)
    v = np.clip(v, 0.05, 0.95).astype(np.float32)
    mask = rng.random((height, width)) < invalid_fraction
    v[mask] = -1.0
    return v


def synthetic_tr(width, height):
    """Transmittance buffer Transmittance.{R,G,B}: smooth per-pixel RGB in (0,1)."""
    y, x = np.mgrid[0:height, 0:width].astype(np.float32)
    base = 0.5 + 0.4 * np.cos(x / width * 4.0) * np.sin(y / height * 4.5 + 0.5)
    tr = np.stack([base, np.clip(base * 0.9 + 0.05, 0, 1), np.clip(base * 0.8 + 0.1, 0, 1)], axis=-1)
    return np.clip(tr, 0.02, 0.98).astype(np.float32)
