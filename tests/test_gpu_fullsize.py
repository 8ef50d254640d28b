"""BASELINE.json's full sizes (C1: 256^2 x 16 spp, C2: 512^2 x 64 spp = 16.8 M paths, fBm 128^3) on the GPU, checked through
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


@pytest.mark.parametrize("name", ["C1", "C2"])
def test_full_size_pixels_equal_the_oracle(ctx, orc, name, c2_full):
    cfg = configs.config1() if name == "C1" else c2_full
    old = cfg.ip.use_nee
    cfg.ip.use_nee = 0      # without NEE every decision comes from the path's own stream: pixels agree path by path
    try:
        _setup(ctx, cfg)
        ctx.film_clear()
        ctx.render(0, cfg.spp, seed=5)
        film = ctx.film_download()
        rng = np.random.default_rng(8)
        n_px = 48
        xs, ys = rng.integers(0, cfg.width, n_px), rng.integers(0, cfg.height, n_px)
        close = 0
        for x, y in zip(xs, ys):
            f, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, window=(int(x), int(y), int(x) + 1, int(y) + 1), seed=5, nthreads=1)
            a, b = film[y, x], f[y, x]
            assert a[3] == b[3] == cfg.spp
            err = np.abs(a[:3] - b[:3]) / np.maximum(np.abs(b[:3]), 1e-3 * cfg.spp)
            close += err.max() < 1e-3
        assert close >= n_px - 2, f"{close} of {n_px} pixels agree"      # a last-ulp libm difference may flip a rare comparison
    finally:
        cfg.ip.use_nee = old
