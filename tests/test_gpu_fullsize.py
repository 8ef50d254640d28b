"""BASELINE.json's full sizes (C1: 256^2 x 16 spp; C2 / C3: 512^2 x 64 spp = 16.8 M paths, fBm 128^3; C4: 1080p, 512^3;
C5: 1080p, 256^3 rgbgrid) on the GPU, checked through
size-independent properties -- sample accounting, sharding linearity, capacity independence -- and against the CPU
oracle on a random sample of whole pixels (the oracle renders single-pixel windows at full spp: the per-path streams
do not depend on what else is rendered, so a pixel of the full-size film must equal the oracle's pixel)."""
import numpy as np
import pytest

import helpers as H
from pyvspg import configs

pytestmark = pytest.mark.gpu


def _setup(ctx, cfg):
    ctx.set_scene(cfg.scene)
    ctx.set_guiding(cfg.guiding)
    ctx.set_integrator(cfg.ip)
    ctx.set_capacity(0)


@pytest.fixture(scope="module")
def c2_full():
    return configs.config2()       # 512 x 512, 64 spp, 128^3 grid, guiding cache + ISG buffer


def test_c2_full_size_sample_accounting_and_sharding(ctx, c2_full):
    cfg = c2_full
    assert (cfg.width, cfg.height, cfg.spp) == (512, 512, 64)
    _setup(ctx, cfg)
    ctx.film_clear(); ctx.counters_reset()
    ctx.render(0, cfg.spp, seed=1)
    whole, cnt = ctx.film_download(), ctx.counters()
    n = cfg.width * cfg.height * cfg.spp
    assert cnt["paths"] == n == 16777216
    assert np.array_equal(whole[..., 3], np.full((cfg.height, cfg.width), float(cfg.spp)))     # box filter: weight 1 per sample
    assert np.isfinite(whole).all() and (whole[..., :3] >= 0).all() and whole[..., :3].mean() > 0
    # the medium is traversed: several density queries per path on average, bounded by the 10000-collision cap
    assert 2 * n < cnt["density_queries"] < 2000 * n
    # interleaved sample passes (the multi-GPU partition) and row bands sum to the same film up to fp64 atomic ordering
    ctx.film_clear()
    for r in range(4):
        ctx.render(r, cfg.spp, spp_stride=4, seed=1)
    assert np.allclose(ctx.film_download(), whole, rtol=1e-11, atol=0)
    ctx.film_clear()
    for y0 in range(0, cfg.height, 128):
        ctx.render(0, cfg.spp, window=(0, y0, cfg.width, y0 + 128), seed=1)
    assert np.allclose(ctx.film_download(), whole, rtol=1e-11, atol=0)
    # wavefront capacity does not change a single path
    ctx.set_capacity(300000)
    ctx.film_clear()
    ctx.render(0, cfg.spp, seed=1)
    assert np.allclose(ctx.film_download(), whole, rtol=1e-11, atol=0)
    ctx.set_capacity(0)


def _full(name, c2_full):
    """BASELINE.json's full film and grid sizes; C4 / C5 at 4 spp (their 256 / 1024 spp only repeat the same per-sample work)."""
    if name == "C1":
        return configs.config1()
    if name == "C2":
        return c2_full
    if name == "C3":
        return configs.config3()              # 512 x 512, 64 spp, NDS+ with the loaded Tr buffer
    if name == "C4":
        return configs.config4(spp=4)         # 1920 x 1080, 512^3 cloud (4.3 GB of cell records), distant + infinite light
    return configs.config5(spp=4)             # 1920 x 1080, 256^3 rgbgrid, emissive quads + boxes


@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4", "C5"])
def test_full_size_pixels_equal_the_oracle(ctx, orc, name, c2_full):
    """Whole pixels of the full-size film, NEE on, against single-pixel oracle renders at the same spp.  Paths are
    bit-identical (tests/test_ref_pbrt.py), so a pixel's fp64 sums differ by the order of its atomic adds only."""
    cfg = _full(name, c2_full)
    _setup(ctx, cfg)
    ctx.film_clear()
    ctx.render(0, cfg.spp, seed=5)
    film = ctx.film_download()
    rng = np.random.default_rng(8)
    n_px = 48 if name in ("C1", "C2", "C3") else 24
    xs, ys = rng.integers(0, cfg.width, n_px), rng.integers(0, cfg.height, n_px)
    scratch = np.zeros((cfg.height, cfg.width, 4), dtype=np.float64)
    bad = []
    for x, y in zip(xs, ys):
        scratch[y, x] = 0
        f, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, window=(int(x), int(y), int(x) + 1, int(y) + 1), seed=5, nthreads=1, film=scratch)
        a, b = film[y, x], f[y, x]
        assert a[3] == b[3] == cfg.spp
        if not np.allclose(a[:3], b[:3], rtol=1e-11, atol=0):
            bad.append((int(x), int(y), a[:3].tolist(), b[:3].tolist()))
    assert not bad, f"{len(bad)} of {n_px} pixels differ from the oracle: {bad[:3]}"


def test_c4_cell_records_and_majorants_at_512_cubed(ctx, orc):
    """The 512^3 grid goes through k_build_cells into 4.3 GB of 32-byte cell records: density fetches and majorant
    segments on it are bit-exact against the oracle reading the plain float grid."""
    cfg = configs.config4(spp=1)
    _setup(ctx, cfg)
    med = cfg.scene.desc.media[0]
    assert (med.nx, med.ny, med.nz) == (512, 512, 512)
    rng = np.random.default_rng(5)
    pts = (rng.uniform(-0.05, 1.05, size=(3000, 3)) - cfg.scene.eye).astype(np.float32)
    got = ctx.debug_density(0, pts)
    want = np.array([orc.density_lookup(med, p) for p in pts], dtype=np.float32)
    assert np.array_equal((got[:, 0] + got[:, 3]).view(np.uint32), want.view(np.uint32))
    assert (want > 0).sum() > 300
    rays = H.random_rays(rng, 2048, cfg.scene.eye)
    seg_g, cnt_g = ctx.debug_majorant_segments(0, rays, 64)
    seg_o, cnt_o = orc.majorant_segments(med, rays, 64)
    assert np.array_equal(cnt_g, cnt_o)
    for i in range(len(cnt_o)):
        c = min(int(cnt_o[i]), 64)
        assert np.array_equal(seg_g[i, :c].view(np.uint32), seg_o[i, :c].view(np.uint32)), f"ray {i}"
