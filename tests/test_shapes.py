"""Shape intersection: bilinear patches, watertight triangles and interval-arithmetic spheres.

CPU: the oracle's Sphere::BasicIntersect restatement against the analytic double-precision solution and the
properties the reference's shapes_test.cpp checks (hit point on the surface, re-intersection from the spawned
origin does not self-hit, rays from inside hit the far root, tMax cuts).  GPU: the CUDA BVH walk + shape tests +
InteractionFromIntersection against the oracle's brute-force closest hit, bit for bit, through the C ABI.
"""
import numpy as np
import pytest

import helpers as H
from pyvspg import _abi as A, api
from pyvspg.scenes import Scene, perspective_camera


def _sphere_scene(center, radius, extra=()):
    cam, eye = perspective_camera([0.5, 0.5, -1.4], [0.5, 0.5, 0.5], [0, 1, 0], 40.0, 16, 16)
    sc = Scene(cam, eye)
    m = sc.add_material([.5, .5, .5])
    sc.add_sphere(center, radius, material=m)
    for c, r in extra:
        sc.add_sphere(c, r, material=m)
    return sc.finalize(), eye


def _rays_at_sphere(rng, n, c, r, spread=1.6):
    o = c + rng.normal(size=(n, 3)) * r * 3
    target = c + rng.uniform(-spread, spread, size=(n, 3)) * r
    d = target - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d, np.full((n, 1), np.inf)], axis=1).astype(np.float32)


def _analytic(rays, c, r):
    o, d = rays[:, :3].astype(np.float64) - c, rays[:, 3:6].astype(np.float64)
    a = (d * d).sum(1); b = 2 * (o * d).sum(1); cc = (o * o).sum(1) - r * r
    disc = b * b - 4 * a * cc
    ok = disc > 0
    sq = np.sqrt(np.where(ok, disc, 0))
    t0, t1 = (-b - sq) / (2 * a), (-b + sq) / (2 * a)
    t = np.where(t0 > 0, t0, t1)
    return ok & (t > 0), t, disc / (4 * a * r * r)


def test_oracle_sphere_matches_the_analytic_solution(orc):
    rng = np.random.default_rng(11)
    c_world, r = np.array([0.55, 0.5, 0.5]), 0.28
    sc, eye = _sphere_scene(c_world, r)
    c = np.float32(c_world - eye).astype(np.float64)   # render-space centre as the desc carries it
    rays = _rays_at_sphere(rng, 4000, c, r)
    prim, out = orc.intersect(sc, api.integrator_defaults(), rays)
    hit, t, margin = _analytic(rays, c, np.float64(np.float32(r)))
    clear = np.abs(margin) > 1e-4            # away from grazing incidence the two must agree on hit / miss
    assert np.array_equal((prim >= 0)[clear], hit[clear])
    assert (prim >= 0).sum() > 1000 and (prim < 0).sum() > 500
    k = (prim >= 0) & clear
    assert np.abs(out[k, 0] - t[k]).max() <= 2e-5 * np.abs(t[k]).max()
    p = out[k, 1:4].astype(np.float64)
    # refined hit point lies on the sphere to float precision, the error bound covers the residual, n is radial
    resid = np.abs(np.linalg.norm(p - c, axis=1) - r)
    assert resid.max() < 1e-6
    assert (out[k, 4:7] > 0).all() and (np.linalg.norm(out[k, 4:7], axis=1) >= resid * 0.5).all()
    radial = (p - c) / np.linalg.norm(p - c, axis=1, keepdims=True)
    assert np.abs((radial * out[k, 7:10]).sum(1) - 1).max() < 1e-5


def test_oracle_sphere_inside_rays_tmax_and_misses(orc):
    rng = np.random.default_rng(12)
    c_world, r = np.array([0.3, 0.4, 0.9]), 0.5
    sc, eye = _sphere_scene(c_world, r)
    c = np.float32(c_world - eye).astype(np.float64)
    n = 500
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    inside = np.concatenate([c + rng.uniform(-0.25, 0.25, size=(n, 3)), d, np.full((n, 1), np.inf)], axis=1).astype(np.float32)
    prim, out = orc.intersect(sc, api.integrator_defaults(), inside)
    assert (prim == 0).all()                                     # every ray from inside leaves through the far root
    hit, t, _ = _analytic(inside, c, np.float64(np.float32(r)))
    assert np.abs(out[:, 0] - t).max() < 1e-5
    assert ((out[:, 7:10] * inside[:, 3:6]).sum(1) > 0).all()    # outward normal faces along the exiting ray
    cut = inside.copy(); cut[:, 6] = out[:, 0] * 0.999
    assert (orc.intersect(sc, api.integrator_defaults(), cut)[0] == -1).all()
    keep = inside.copy(); keep[:, 6] = out[:, 0] * 1.001
    assert (orc.intersect(sc, api.integrator_defaults(), keep)[0] == 0).all()
    away = inside.copy(); away[:, :3] = np.float32(c + d * 2 * r); away[:, 6] = np.inf   # outside, pointing away
    assert (orc.intersect(sc, api.integrator_defaults(), away)[0] == -1).all()


def test_oracle_sphere_spawned_rays_do_not_self_intersect(orc):
    """shapes_test.cpp 'Reintersect' idea: a ray spawned from a hit (OffsetRayOrigin with pi.Error(), ray.h:75-92)
    along the outward normal must not hit the same sphere again."""
    rng = np.random.default_rng(13)
    c_world, r = np.array([0.55, 0.5, 0.5]), 0.28
    sc, eye = _sphere_scene(c_world, r)
    c = np.float32(c_world - eye).astype(np.float64)
    rays = _rays_at_sphere(rng, 3000, c, r, spread=0.9)
    prim, out = orc.intersect(sc, api.integrator_defaults(), rays)
    k = prim >= 0
    p, e, n = out[k, 1:4], out[k, 4:7], out[k, 7:10]
    w = n + rng.normal(size=n.shape).astype(np.float32) * 0.7
    w[(w * n).sum(1) <= 0.05] = n[(w * n).sum(1) <= 0.05]
    w /= np.linalg.norm(w, axis=1, keepdims=True)
    dd = (np.abs(n) * e).sum(1, keepdims=True)
    off = dd * n
    o = p + off
    o = np.where(off > 0, np.nextafter(o, np.float32(np.inf)), np.where(off < 0, np.nextafter(o, np.float32(-np.inf)), o)).astype(np.float32)
    again = np.concatenate([o, w, np.full((o.shape[0], 1), np.inf)], axis=1).astype(np.float32)
    assert (orc.intersect(sc, api.integrator_defaults(), again)[0] == -1).all()


def _scene_rays(rng, cfg, n):
    eye = cfg.scene.eye
    o = rng.uniform(-0.1, 1.1, size=(n, 3)) - eye
    t = rng.uniform(-0.2, 1.2, size=(n, 3)) - eye
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    tmax = np.where(rng.random(n) < 0.3, rng.uniform(0.05, 1.5, size=n), np.inf)
    return np.concatenate([o, d, tmax[:, None]], axis=1).astype(np.float32)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["C1", "C1T", "C1S"])
def test_gpu_intersection_is_bit_identical_to_the_oracle(name):
    import orc as O
    cfg = H.small(name)
    rng = np.random.default_rng(21)
    rays = _scene_rays(rng, cfg, 20000)
    ctx = api.Context()
    try:
        ctx.set_scene(cfg.scene)
        prim_g, out_g = ctx.debug_intersect(rays)
    finally:
        ctx.close()
    prim_o, out_o = O.intersect(cfg.scene, cfg.ip, rays)
    assert (prim_o >= 0).mean() > 0.5
    kinds = np.unique(prim_o[prim_o >= 0])
    assert len(kinds) >= min(6, cfg.scene.desc.n_quads + cfg.scene.desc.n_triangles + cfg.scene.desc.n_spheres)
    assert np.array_equal(prim_g, prim_o)
    assert np.array_equal(out_g.view(np.uint32), out_o.view(np.uint32))


@pytest.mark.gpu
def test_gpu_sphere_intersection_bits_on_dense_rays():
    """Many spheres of very different sizes and offsets (large |centre| stresses the Point3fi error terms)."""
    import orc as O
    rng = np.random.default_rng(22)
    extra = [(rng.uniform(-3, 3, size=3), float(r)) for r in (1e-3, 0.02, 0.4, 2.5, 40.0)]
    sc, eye = _sphere_scene(np.array([0.5, 0.5, 0.5]), 0.3, extra=extra)
    spheres = [(np.array([0.5, 0.5, 0.5]), 0.3)] + extra
    rays = np.concatenate([_rays_at_sphere(rng, 6000, c - eye, r) for c, r in spheres])
    ctx = api.Context()
    try:
        ctx.set_scene(sc)
        prim_g, out_g = ctx.debug_intersect(rays)
    finally:
        ctx.close()
    prim_o, out_o = O.intersect(sc, api.integrator_defaults(), rays)
    assert len(np.unique(prim_o[prim_o >= 0])) == len(spheres)
    assert np.array_equal(prim_g, prim_o)
    assert np.array_equal(out_g.view(np.uint32), out_o.view(np.uint32))
