"""The pbrt half of the path, pinned to the reference renderer itself.

tests/golden/pbrt/*.npz hold what the REFERENCE computes -- its own code, compiled from /root/reference by
oracle/build_ref_pbrt.sh and driven by oracle/ref_harness/pbrt_harness.cpp, running with the path's portable libm
(include/vspg_libm.h via oracle/libportm.so) -- on the scene files next to them: camera matrices and rays, closest hits,
SampleT_maj callbacks, light picks, Light::SampleLi / PDF_Li, SampleLd, and Li for 1500 (pixel, sample) pairs per scene.

CPU tests (here): the C++ front end reproduces the reference's camera bit for bit, and the oracle restatement reproduces
every one of those vectors BIT FOR BIT -- including the whole-path radiance.  GPU tests (-m gpu): the CUDA path does the
same through the C ABI.  The scenes cover every BASELINE config family (tests/ref_scenes.py).
"""
import ctypes as C
import os

import numpy as np
import pytest

from pyvspg import _abi as A
from pyvspg import api
from ref_scenes import SCENES

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pbrt")
PF, PI = C.POINTER(C.c_float), C.POINTER(C.c_int)


def fp(a):
    return a.ctypes.data_as(PF)


def bits(a):
    """Bit patterns, with every NaN mapped to one pattern (the reference's light-BVH PMF is 0/0 = NaN for a light whose
    importance is zero at every node on its path; payload bits of a NaN carry no meaning)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    return np.where(np.isnan(a), np.uint32(0x7fc00000), a.view(np.uint32))


def load(name):
    hs = api.HostScene(path=os.path.join(GOLD, name + ".pbrt"))
    return hs, np.load(os.path.join(GOLD, name + ".npz"))


def guiding_ptr(hs):
    return C.byref(hs.guiding.desc) if hs.guiding is not None else None


@pytest.fixture(scope="module")
def orc():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
    import orc as o   # the checker
    L = o.lib()
    L.orc_sample_tmaj.argtypes = [C.c_void_p, C.c_int, PF, C.c_float, C.c_float, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, PF, PF]
    L.orc_sample_ld.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, PF, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, PF, C.c_int, PF]
    return o


# ------------------------------------------------------------------------------------------------ front end
@pytest.mark.parametrize("name", SCENES)
def test_front_end_camera_matrices_equal_the_references(name):
    """cameraFromRaster / renderFromCamera as ProjectiveCamera / CameraTransform build them (cameras.h:259-278,
    cameras.cpp:26-57) -- composed by csrc/host_xform.h with the reference's compensated float arithmetic."""
    hs, g = load(name)
    c = hs.desc.camera
    assert np.array_equal(bits(np.array(c.camera_from_raster.m, dtype=np.float32).reshape(4, 4)), bits(g["cam_camera_from_raster"]))
    assert np.array_equal(bits(np.array(c.render_from_camera.m, dtype=np.float32).reshape(4, 4)), bits(g["cam_render_from_camera"]))
    hs.close()


@pytest.mark.parametrize("name", SCENES)
def test_oracle_camera_rays_are_bit_equal(name, orc):
    hs, g = load(name)
    smp, rays = g["samples"], g["rays"]
    for (px, py, s), r in zip(smp[:400], rays[:400]):
        o, d, ch = orc.camera_ray(hs, int(px), int(py), int(s), hs.seed)
        assert np.array_equal(bits(o), bits(r[:3])) and np.array_equal(bits(d), bits(r[3:6])) and ch == int(r[6]), (px, py, s)
    hs.close()


# ------------------------------------------------------------------------------------------------ oracle, stage by stage
@pytest.mark.parametrize("name", SCENES)
def test_oracle_closest_hits_are_bit_equal(name, orc):
    """Integrator::Intersect -> tHit, p() (the midpoint of the outward-rounded Point3fi), n."""
    hs, g = load(name)
    prim, out = orc.intersect(hs, hs.ip, g["hit_rays"])
    ref = g["hit_out"]
    assert np.array_equal(prim >= 0, g["hit_valid"] == 1)
    m = prim >= 0
    assert np.array_equal(bits(out[m, 0:4]), bits(ref[m, 0:4]))     # tHit, p
    assert np.array_equal(bits(out[m, 7:10]), bits(ref[m, 4:7]))    # n
    hs.close()


@pytest.mark.parametrize("name", [s for s in SCENES])
def test_oracle_sample_t_maj_is_bit_equal(name, orc):
    """SampleT_maj (media_sampleTMaj.h:39-117) over homogeneous / uniformgrid / rgbgrid media: every tentative collision
    (p, sigma_a, sigma_s, sigma_maj, T_maj), the returned T_maj and the tracker's next random number."""
    hs, g = load(name)
    for mi in range(hs.desc.n_media):
        rays, tmax, u, ch, n_ref, rec_ref, fin_ref = (g[f"tm{mi}_{k}"] for k in ("ray", "tmax", "u", "ch", "n", "rec", "final"))
        cap = rec_ref.shape[1]
        for i in range(len(rays)):
            rec, fin = np.zeros((48, 15), np.float32), np.zeros(4, np.float32)
            n = orc.lib().orc_sample_tmaj(C.byref(hs.desc), mi, fp(np.ascontiguousarray(rays[i])), float(tmax[i]), float(u[i]), 1000 + i, 17, int(ch[i]),
                                          -1, 48, fp(rec), fp(fin))
            assert n == n_ref[i], (mi, i, n, n_ref[i])
            k = min(n, cap)
            assert np.array_equal(bits(rec[:k]), bits(rec_ref[i, :k])), (mi, i)
            assert np.array_equal(bits(fin), bits(fin_ref[i])), (mi, i)
    hs.close()


@pytest.mark.parametrize("name", SCENES)
def test_oracle_light_sampler_and_sample_li_are_bit_equal(name, orc):
    """LightSampler::Sample / PMF (uniform, power = alias table over Phi / lambda.PDF(), BVH) and Light::SampleLi
    (point, distant, spherical rectangles / triangles, cone-sampled spheres) with Light::PDF_Li at the sampled direction."""
    hs, g = load(name)
    if "li_ctx" not in g:
        pytest.skip("scene has no lights")
    ctx, light = g["li_ctx"], g["li_light"]
    pick, pmf, pmf_of = orc.light_sample(hs, hs.ip, ctx[:, :3], ctx[:, 6:9], g["ls_u"], light)
    assert np.array_equal(pick, g["ls_pick"])
    assert np.array_equal(bits(pmf), bits(g["ls_pmf"])) and np.array_equal(bits(pmf_of), bits(g["ls_pmf_of"]))
    li, ref = orc.sample_li(hs, hs.ip, ctx, light, g["li_u2"]), g["li_out"]
    assert np.array_equal(bits(li[:, :11]), bits(ref[:, :11]))
    cmpb = li[:, 11] != -1   # the oracle evaluates PDF_Li only where the sampled ray's first scene hit is the light's own shape
    assert np.array_equal(bits(li[cmpb, 11]), bits(ref[cmpb, 11]))
    hs.close()


@pytest.mark.parametrize("name", SCENES)
def test_oracle_sample_ld_is_bit_equal(name, orc):
    """GuidedVolPathVSPGIntegrator::SampleLd (vspg.cpp:1136-1252) at seeded surface and volume vertices: light pick,
    SampleLi, guided scatter pdf, the shadow-ray chain through interfaces with ratio tracking + RR, MIS weights."""
    hs, g = load(name)
    rp = np.ones(3, np.float32)
    for kind in ("vol", "surf"):
        if f"ld_{kind}_in" not in g:
            continue
        rays, rc_ref, out_ref, ch = g[f"ld_{kind}_in"], g[f"ld_{kind}_rc"], g[f"ld_{kind}_out"], g[f"ld_{kind}_ch"]
        for i in range(len(rays)):
            out = np.zeros(3, np.float32)
            rc = orc.lib().orc_sample_ld(C.byref(hs.desc), guiding_ptr(hs), C.byref(hs.ip), 1 if kind == "surf" else 0, fp(np.ascontiguousarray(rays[i])),
                                         -1 if kind == "surf" else 0, 3, 4, i, 9, hs.seed, fp(rp), int(ch[i]), fp(out))
            assert rc == rc_ref[i], (kind, i, rc, rc_ref[i])
            if rc == 0:
                assert np.array_equal(bits(out), bits(out_ref[i])), (kind, i, out, out_ref[i])
    hs.close()


# ------------------------------------------------------------------------------------------------ oracle, whole path
@pytest.mark.parametrize("name", SCENES)
def test_oracle_li_is_bit_equal_to_the_references_li(name, orc):
    """RayIntegrator::EvaluatePixelSample -> GuidedVolPathVSPGIntegrator::Li for 1500 (pixel, sample) pairs per scene."""
    hs, g = load(name)
    smp, Lref = g["samples"], g["L"]
    out = np.zeros(3, np.float32)
    bad = []
    for i, (px, py, s) in enumerate(smp):
        orc.lib().orc_render_pixel_sample(C.byref(hs.desc), guiding_ptr(hs), C.byref(hs.ip), int(px), int(py), int(s), hs.seed, fp(out))
        if not np.array_equal(bits(out), bits(Lref[i])):
            bad.append((int(px), int(py), int(s), out.copy(), Lref[i]))
    assert not bad, f"{len(bad)} of {len(smp)} samples differ from the reference, first: {bad[0]}"
    assert (Lref != 0).any(axis=1).sum() > 400
    hs.close()


# ------------------------------------------------------------------------------------------------ CUDA path (through the C ABI)
@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
def test_gpu_li_is_bit_equal_to_the_references_li(name, ctx):
    """vspg_render, one sample index at a time: the film then holds w * L of exactly one path per pixel, which must be
    the reference's Li for that (pixel, sample) -- bit for bit, NEE shadow chains included."""
    hs, g = load(name)
    ctx.set_scene(hs)
    if hs.guiding is not None:
        ctx.set_guiding(hs.guiding)
    ctx.set_integrator(hs.ip)
    smp, Lref, w = g["samples"], g["L"], g["rays"][:, 7]
    films = {}
    for s in range(hs.spp):
        ctx.film_clear()
        ctx.render(s, s + 1, seed=hs.seed)
        films[s] = ctx.film_download().copy()
    got = np.stack([films[int(s)][int(py), int(px), :3] for px, py, s in smp]).astype(np.float32)
    wsum = np.array([films[int(s)][int(py), int(px), 3] for px, py, s in smp], dtype=np.float32)
    want = (w[:, None] * Lref).astype(np.float32)   # RGBFilm::AddSample: rgbSum += weight * rgb (film.h:251-267)
    assert np.array_equal(bits(wsum), bits(w))
    same = (bits(got) == bits(want)).all(axis=1)
    assert same.all(), f"{(~same).sum()} of {len(smp)} samples differ from the reference, first: {smp[~same][0]}, {got[~same][0]} vs {want[~same][0]}"
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
def test_gpu_hits_light_picks_and_sample_li_are_bit_equal(name, ctx):
    hs, g = load(name)
    ctx.set_scene(hs)
    ctx.set_integrator(hs.ip)
    prim, out = ctx.debug_intersect(g["hit_rays"])
    ref = g["hit_out"]
    assert np.array_equal(prim >= 0, g["hit_valid"] == 1)
    m = prim >= 0
    assert np.array_equal(bits(out[m, 0:4]), bits(ref[m, 0:4])) and np.array_equal(bits(out[m, 7:10]), bits(ref[m, 4:7]))
    if "li_ctx" in g:
        c9, light = g["li_ctx"], g["li_light"]
        pick, pmf, pmf_of = ctx.debug_light_sample(c9[:, :3], c9[:, 6:9], g["ls_u"], light)
        assert np.array_equal(pick, g["ls_pick"])
        assert np.array_equal(bits(pmf), bits(g["ls_pmf"])) and np.array_equal(bits(pmf_of), bits(g["ls_pmf_of"]))
        li, lref = ctx.debug_sample_li(c9, light, g["li_u2"]), g["li_out"]
        assert np.array_equal(bits(li[:, :11]), bits(lref[:, :11]))
        cmpb = li[:, 11] != -1
        assert np.array_equal(bits(li[cmpb, 11]), bits(lref[cmpb, 11]))
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
def test_gpu_sample_ld_is_bit_equal_to_the_references_sample_ld(name, ctx):
    """vspg_debug_sample_ld (nee_presample + scatter pdf + shadow_trace: the code of k_lookup / k_shade / k_volscatter /
    k_shadow) against GuidedVolPathVSPGIntegrator::SampleLd run by the reference itself at the same seeded vertices:
    decisions identical, Ld bit for bit (gray and RGB trackers, interface chains, delta and area lights)."""
    from pyvspg import _abi as A
    hs, g = load(name)
    ctx.set_scene(hs)
    if hs.guiding is not None:
        ctx.set_guiding(hs.guiding)
    ctx.set_integrator(hs.ip)
    checked = 0
    for kind in ("vol", "surf"):
        if f"ld_{kind}_in" not in g:
            continue
        rays, rc_ref, out_ref, ch = g[f"ld_{kind}_in"], g[f"ld_{kind}_rc"], g[f"ld_{kind}_out"], g[f"ld_{kind}_ch"]
        inp = (A.LdInput * len(rays))()
        for i, r in enumerate(rays):
            I = inp[i]
            I.is_surface = 1 if kind == "surf" else 0
            I.ray_o[:] = [float(x) for x in r[:3]]
            I.ray_d[:] = [float(x) for x in r[3:6]]
            I.medium = -1 if kind == "surf" else 0
            I.px, I.py, I.sample_index, I.dim, I.seed = 3, 4, i, 9, hs.seed
            I.r_p[:] = [1.0, 1.0, 1.0]
            I.channel = int(ch[i])
        status, Ld = ctx.debug_sample_ld(inp)
        assert np.array_equal(status, rc_ref.astype(np.int32)), (kind, status, rc_ref)
        ok = rc_ref == 0
        same = (bits(Ld[ok]) == bits(out_ref[ok])).all(axis=1)
        assert same.all(), f"{kind}: {(~same).sum()} of {ok.sum()} differ, first {Ld[ok][~same][0]} vs {out_ref[ok][~same][0]}"
        checked += int(ok.sum())
    assert checked > 0 or name.startswith("c1")
    hs.close()
