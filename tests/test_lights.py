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


@pytest.mark.parametrize("name", ["C5", "C2", "C1", "C1E"])
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
        # a light in a subtree whose two children both have zero importance gets 0/0 from BVHLightSampler::PMF
        # (lightsamplers.h:330-340; the reference only DCHECKs ci[child] > 0): such a light is never sampled
        assert not np.isnan(pmf_of[light == li]).any()
        total += np.nan_to_num(pmf_of, nan=0.0)
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
@pytest.mark.parametrize("name", ["C5", "C2", "C4", "C1", "C1E"])
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
        nan = np.isnan(want[2])                       # 0/0 for a light under a zero-importance subtree: NaN payload bits differ by platform
        assert np.array_equal(np.isnan(got[2]), nan)
        assert np.array_equal(got[2][~nan].view(np.uint32), want[2][~nan].view(np.uint32)), f"sampler {ls}: PMF(ctx, light) bits differ"


# ------------------------------------------------------------------ Light::SampleLi / PDF_Li (area lights)
def _area_light_scene():
    """Cornell-style room with a rectangular light (spherical-rectangle branch, shapes.cpp:1280-1327), a small far
    rectangle (falls back to area sampling below MinSphericalSampleArea) and a non-rectangular emissive patch
    (bilinear area sampling)."""
    from pyvspg import api
    from pyvspg.scenes import Scene, perspective_camera
    cam, eye = perspective_camera([0.5, 0.5, -1.4], [0.5, 0.5, 0.5], [0, 1, 0], 40.0, 16, 16)
    sc = Scene(cam, eye)
    w = sc.add_material([.7, .7, .7])
    sc.add_quad([0.3, 0.999, 0.3], [0.4, 0, 0], [0, 0, 0.4], material=w, emission=[17, 12, 4])                 # big ceiling light, faces down
    sc.add_quad([30.0, 40.0, 30.0], [0.2, 0, 0], [0, 0, 0.2], material=w, emission=[5, 5, 5], two_sided=True)  # tiny solid angle
    sc.add_quad(None, None, None, material=w, emission=[3, 2, 1], two_sided=True,
                corners=[[1.5, -0.3, 0.2], [1.9, -0.3, 0.2], [1.5, -0.3, 0.7], [2.1, -0.3, 0.9]])                 # planar, not a rectangle
    sc.add_triangles([[-0.4, 0.1, 0.2], [-0.4, 0.9, 0.3], [-0.45, 0.4, 0.9], [60.0, 10.0, 5.0], [60.0, 10.2, 5.0], [60.0, 10.0, 5.2]],
                     [(0, 1, 2), (3, 4, 5)], material=w, emission=[6, 7, 8], two_sided=True)               # lights 3 (near) and 4 (tiny solid angle)
    sc.add_sphere([0.2, 0.35, 1.4], 0.17, material=w, emission=[9, 3, 3])                                     # light 5: seen from outside (cone sampling)
    sc.add_sphere([0.5, 0.4, 0.5], 80.0, material=w, emission=[0.2, 0.2, 0.3], two_sided=True)                # light 6: reference points inside
    return sc.finalize(), api.integrator_defaults()


N_TEST_LIGHTS = 7


def _li_inputs(scene, rng, n):
    eye = scene.eye.astype(np.float32)
    p = rng.uniform(-0.2, 1.2, size=(n, 3)).astype(np.float32)
    p[:, 1] = rng.uniform(-0.2, 0.95, size=n)          # below the ceiling light
    p -= eye
    nn = rng.normal(size=(n, 3)).astype(np.float32)
    nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    nn[::2] = 0                                         # medium vertices
    ctx9 = np.concatenate([p, nn, nn], axis=1)
    return ctx9, rng.integers(0, N_TEST_LIGHTS, n).astype(np.int32), rng.random((n, 2)).astype(np.float32)


def test_area_light_sample_pdf_matches_pdf_li(orc):
    """shapes_test.cpp's Sample-vs-PDF consistency: the pdf SampleLi returns equals PDF_Li along the sampled direction
    (InvertSphericalRectangleSample must undo SampleSphericalRectangle through the cos-theta warp)."""
    scene, ip = _area_light_scene()
    rng = np.random.default_rng(3)
    ctx9, light, u2 = _li_inputs(scene, rng, 14000)
    out = orc.sample_li(scene, ip, ctx9, light, u2)
    ok = (out[:, 0] == 1) & (out[:, 11] >= 0)
    for l in range(N_TEST_LIGHTS):
        assert (ok & (light == l)).sum() > 300, l
    rel = np.abs(out[ok, 11] - out[ok, 1]) / out[ok, 1]
    assert np.quantile(rel, 0.99) < 2e-3 and np.median(rel) < 1e-4, (np.quantile(rel, 0.99), np.median(rel))
    # sampled points lie on the light and the direction points at them
    d = out[ok, 5:8] - ctx9[ok, :3]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    assert np.abs((d * out[ok, 2:5]).sum(1) - 1).max() < 1e-5


def test_spherical_rectangle_sampling_is_uniform_in_solid_angle(orc):
    """From a medium vertex (no warp) the pdf is 1 / solid angle for every sample and the samples cover the
    rectangle: E[1/pdf] = solid angle and the mean sampled direction matches a brute-force solid-angle average."""
    scene, ip = _area_light_scene()
    rng = np.random.default_rng(4)
    n = 20000
    p = np.float32([0.45, 0.3, 0.55]) - scene.eye.astype(np.float32)
    ctx9 = np.zeros((n, 9), np.float32); ctx9[:, :3] = p
    out = orc.sample_li(scene, ip, ctx9, np.zeros(n, np.int32), rng.random((n, 2)).astype(np.float32))
    assert (out[:, 0] == 1).all()
    pdf = out[:, 1]
    assert np.ptp(pdf) <= 1e-6 * pdf[0]
    # brute force: area-uniform points on the light weighted by cos / r^2
    q = scene.desc.quads[0]
    p00, p10, p01 = (np.array(list(getattr(q, k)), np.float64) for k in ("p00", "p10", "p01"))
    uv = rng.random((400000, 2))
    pts = p00 + uv[:, :1] * (p10 - p00) + uv[:, 1:] * (p01 - p00)
    nrm = np.cross(p10 - p00, p01 - p00); area = np.linalg.norm(nrm); nrm /= area
    w = pts - p.astype(np.float64)
    r2 = (w * w).sum(1); w /= np.sqrt(r2)[:, None]
    g = np.abs(w @ nrm) / r2
    omega = area * g.mean()
    assert abs(1 / pdf[0] - omega) < 0.01 * omega
    mean_dir = (w * g[:, None]).sum(0) / g.sum()
    got_dir = out[:, 2:5].astype(np.float64).mean(0)
    assert np.abs(mean_dir - got_dir).max() < 0.01


@pytest.mark.gpu
def test_area_light_sampling_matches_the_oracle(ctx, orc):
    scene, ip = _area_light_scene()
    rng = np.random.default_rng(5)
    ctx9, light, u2 = _li_inputs(scene, rng, 20000)
    ctx.set_scene(scene)
    ctx.set_guiding(None)
    ctx.set_integrator(ip)
    got = ctx.debug_sample_li(ctx9, light, u2)
    want = orc.sample_li(scene, ip, ctx9, light, u2)
    assert np.array_equal(got[:, 0], want[:, 0])
    ok = want[:, 0] == 1
    assert ok.mean() > 0.5
    # sinf / cosf / asinf / atan2f differ from glibc by an ulp or two: 1e-5 relative on pdf, 1e-5 absolute on wi / pLight
    assert (np.abs(got[ok, 1] - want[ok, 1]) <= 1e-4 * np.abs(want[ok, 1])).mean() > 0.999
    assert np.median(np.abs(got[ok, 1] - want[ok, 1]) / np.abs(want[ok, 1])) < 1e-6
    dw = np.abs(got[ok, 2:5] - want[ok, 2:5]).max(axis=1)     # ill-conditioned spherical-triangle inversions amplify the last-ulp differences
    assert np.quantile(dw, 0.999) < 1e-4 and dw.max() < 5e-3, (np.quantile(dw, 0.999), dw.max())
    assert np.array_equal(got[ok, 8:11], want[ok, 8:11])
    both = ok & (want[:, 11] >= 0) & (got[:, 11] >= 0)
    assert both.sum() > 0.4 * ok.sum()
    rel = np.abs(got[both, 11] - want[both, 11]) / want[both, 11]
    assert np.quantile(rel, 0.99) < 1e-3 and np.median(rel) < 1e-5
