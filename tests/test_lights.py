"""The reference's default light sampler ("bvh", lightsamplers.h:258-404): oracle properties on CPU, and the CUDA
restatement (host-built tree, decoded nodes) against the oracle (quantised nodes decoded per call) on the GPU."""
import numpy as np
import pytest

import helpers as H
from pyvspg import _abi as A


def _many_lights_cfg():
    """C5 has 3 emissive quads; add a point light so the tree has 4 leaves, plus the distant/infinite-style lights of
    the other configs are covered by C2/C4 (one area light + unbounded lights)."""
    cfg = H.small("C5")
    return cfg


def _contexts(cfg, rng, n):
    eye = cfg.scene.eye.astype(np.float32)
    p = rng.uniform(-0.8, 1.8, size=(n, 3)).astype(np.float32) - eye
    ns = rng.normal(size=(n, 3)).astype(np.float32)
    ns /= np.linalg.norm(ns, axis=1, keepdims=True)
    ns[::3] = 0                                    # medium vertices: LightSampleContext::ns = 0
    u = rng.random(n).astype(np.float32)
    return p, ns, u


@pytest.mark.parametrize("name", ["C5", "C2", "C1"])
def test_bvh_sampler_pmf_is_a_distribution_and_matches_sample(orc, name):
    cfg = H.small(name)
    cfg.ip.light_sampler = A.LIGHTSAMPLER_BVH
    nl = cfg.scene.desc.n_lights
    rng = np.random.default_rng(3)
    n = 400
    p, ns, u = _contexts(cfg, rng, n)
    total = np.zeros(n, np.float64)
    for li in range(nl):
        light, pmf, pmf_of = orc.light_sample(cfg.scene, cfg.ip, p, ns, u, np.full(n, li, np.int32))
        total += pmf_of
        hit = light == li
        # Sample() and PMF() walk the same tree with the same importances
        assert np.allclose(pmf[hit], pmf_of[hit], rtol=2e-6)
    ok = light >= 0                                # contexts where every subtree has zero importance return no light
    assert np.allclose(total[ok], 1.0, atol=1e-5), (total[ok].min(), total[ok].max())
    assert ok.mean() > 0.5


def test_bvh_sampler_frequencies_follow_the_pmf(orc):
    cfg = H.small("C5")
    cfg.ip.light_sampler = A.LIGHTSAMPLER_BVH
    nl = cfg.scene.desc.n_lights
    eye = cfg.scene.eye.astype(np.float32)
    p0 = (np.float32([0.5, 0.2, 0.3]) - eye)[None].repeat(20000, 0)
    ns = np.float32([0, 1, 0])[None].repeat(20000, 0)
    u = np.random.default_rng(9).random(20000).astype(np.float32)
    light, pmf, _ = orc.light_sample(cfg.scene, cfg.ip, p0, ns, u, np.zeros(20000, np.int32))
    for li in range(nl):
        sel = light == li
        if sel.any():
            assert abs(sel.mean() - pmf[sel][0]) < 0.015, (li, sel.mean(), pmf[sel][0])


def test_light_samplers_agree_in_expectation(orc):
    """uniform / power / bvh only change the variance: the mean image must agree (SURVEY 4 unbiasedness)."""
    means = {}
    for tag, ls in (("power", A.LIGHTSAMPLER_POWER), ("bvh", A.LIGHTSAMPLER_BVH)):
        cfg = H.small("C5", res=32, spp=48)
        cfg.ip.light_sampler = ls
        film, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, seed=5, nthreads=8)
        px = film[..., :3].mean(axis=2) / cfg.spp
        means[tag] = (px.mean(), px.std(ddof=1) / np.sqrt(px.size) * 3)
    (a, sa), (b, sb) = means["power"], means["bvh"]
    assert abs(a - b) <= 4 * np.hypot(sa, sb) + 0.02 * a, means


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["C5", "C2", "C4", "C1"])
def test_light_sampling_bit_exact_against_the_oracle(ctx, orc, name):
    cfg = H.small(name)
    nl = cfg.scene.desc.n_lights
    rng = np.random.default_rng(17)
    n = 4000
    p, ns, u = _contexts(cfg, rng, n)
    light_of = rng.integers(0, nl, n).astype(np.int32)
    for ls in (A.LIGHTSAMPLER_BVH, A.LIGHTSAMPLER_POWER, A.LIGHTSAMPLER_UNIFORM):
        cfg.ip.light_sampler = ls
        ctx.set_scene(cfg.scene)
        ctx.set_guiding(cfg.guiding)
        ctx.set_integrator(cfg.ip)
        got = ctx.debug_light_sample(p, ns, u, light_of)
        want = orc.light_sample(cfg.scene, cfg.ip, p, ns, u, light_of)
        assert np.array_equal(got[0], want[0]), f"sampler {ls}: {np.sum(got[0] != want[0])} of {n} lights differ"
        assert np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32)), f"sampler {ls}: pmf bits differ"
        assert np.array_equal(got[2].view(np.uint32), want[2].view(np.uint32)), f"sampler {ls}: PMF(ctx, light) bits differ"
