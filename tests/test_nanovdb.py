"""NanoVDB media (reference: media.h:657-753 NanoVDBMedium, media.cpp:549-734).

tests/golden/nvdb_small.nvdb was written by the reference's own vendored NanoVDB (oracle/ref_harness/nvdb_writer.cpp,
driven by tests/golden/make_golden_pbrt.py write_nvdb) from a seeded array that is regenerated here.  The whole-path
pinning of the medium (Li, SampleT_maj, SampleLd of scenes "cn" / "cnu" against the reference renderer) lives in
tests/test_ref_pbrt.py; this file covers the reader, the front end and the per-kernel hooks."""
import os

import numpy as np
import pytest

import helpers as H
from pyvspg import _abi as A
from pyvspg import api
from pyvspg.scenes import Scene, perspective_camera

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NVDB = os.path.join(GOLD, "nvdb_small.nvdb")
I0, VOXEL, ORIGIN = (-5, 3, 100), 0.05, (0.22, -0.15, -4.95)     # the writer's arguments (make_golden_pbrt.py write_nvdb)


def source_array():
    rng = np.random.default_rng(77)
    nx, ny, nz = 20, 18, 16
    g = rng.random((nz, ny, nx), dtype=np.float32)
    z, y, x = np.mgrid[0:nz, 0:ny, 0:nx]
    mask = ((x - 10) ** 2 / 90 + (y - 9) ** 2 / 70 + (z - 8) ** 2 / 50) < 1.0
    return np.where(mask, g * 2.0 + 0.05, 0).astype(np.float32)


def nvdb_scene(density_offset=0.02, majorant_scale=1.25, filename=NVDB, gridname="density"):
    cam, eye = perspective_camera([0.5, 0.6, -2.2], [0.5, 0.45, 0.5], [0, 1, 0], 40.0, 40, 24)
    sc = Scene(cam, eye)
    white = sc.add_material([.6, .6, .6])
    sc.add_quad([-2, -0.05, -2], [5, 0, 0], [0, 0, 5], material=white)
    med = sc.add_nanovdb_medium(filename, [0.1, 0.1, 0.1], [3.0, 3.0, 3.0], 4.0, 0.5, gridname=gridname, density_offset=density_offset,
                                majorant_scale=majorant_scale)
    sc.add_box([0, 0, 0], [1, 1, 1], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    sc.add_point_light([0.2, 1.8, -0.4], [9.0, 8.0, 7.0])
    return sc.finalize()


def test_reader_flattens_the_reference_written_tree_exactly():
    sc = nvdb_scene()
    m = sc.desc.media[0]
    assert m.type == A.MEDIUM_NANOVDB
    src = source_array()
    zs, ys, xs = np.nonzero(src)
    lo = np.array([xs.min() + I0[0], ys.min() + I0[1], zs.min() + I0[2]])
    hi = np.array([xs.max() + I0[0], ys.max() + I0[1], zs.max() + I0[2]])
    assert list(m.index_min) == lo.tolist() and [m.nx, m.ny, m.nz] == (hi - lo + 1).tolist()    # Tree::bbox() = active voxels
    dense = np.ctypeslib.as_array(m.density, (m.nz, m.ny, m.nx))
    want = src[lo[2] - I0[2]:hi[2] - I0[2] + 1, lo[1] - I0[1]:hi[1] - I0[1] + 1, lo[0] - I0[0]:hi[0] - I0[0] + 1]
    assert np.array_equal(dense.view(np.uint32), want.view(np.uint32))
    assert m.background == 0.0
    # index -> world: p = voxel * ijk + origin; the file's world bounding box covers voxels [lo, hi + 1)
    assert np.allclose(list(m.index_from_medium), np.diag([1 / VOXEL] * 3).reshape(-1), rtol=1e-6)
    assert np.allclose(list(m.index_origin), ORIGIN, rtol=1e-6)
    assert np.allclose(list(m.bounds_min), lo * VOXEL + np.array(ORIGIN), atol=1e-5)
    assert np.allclose(list(m.bounds_max), (hi + 1) * VOXEL + np.array(ORIGIN), atol=1e-5)
    assert list(m.maj_res) == [64, 64, 64]


def test_front_end_majorant_grid_equals_the_oracles_restatement(orc):
    """media.cpp:605-670 twice: csrc/host_scene.cpp (product front end) and oracle/orc_media.h; (max + offset) * scale."""
    sc = nvdb_scene()
    m = sc.desc.media[0]
    got = np.ctypeslib.as_array(m.majorant, (64 * 64 * 64,))
    want = orc.build_majorant_grid(m)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert got.min() == np.float32((0 + np.float32(0.02)) * np.float32(1.25)) and got.max() > 2.0
    # conservative: no density sample exceeds the majorant of its cell
    rng = np.random.default_rng(2)
    bmin, bmax = np.array(list(m.bounds_min)), np.array(list(m.bounds_max))
    pts = rng.uniform(bmin, bmax, size=(4000, 3))
    cell = np.minimum(((pts - bmin) / (bmax - bmin) * 64).astype(int), 63)
    sig = np.float32(m.sigma_a[0] + m.sigma_s[0])
    for p, c in zip(pts, cell):
        d = orc.density_lookup(m, (p - sc.eye).astype(np.float32))
        assert d <= got[c[0] + 64 * (c[1] + 64 * c[2])] * sig * (1 + 1e-6)


def test_oracle_sampler_reproduces_trilinear_interpolation_of_the_source(orc):
    sc = nvdb_scene(density_offset=0.0)
    m = sc.desc.media[0]
    src = source_array()
    rng = np.random.default_rng(3)
    sig = float(np.float32(m.sigma_a[0]) + np.float32(m.sigma_s[0]))
    for _ in range(500):
        ijk = rng.uniform([I0[0] + 1, I0[1] + 1, I0[2] + 1], [I0[0] + 18, I0[1] + 16, I0[2] + 14])
        p = ijk * VOXEL + np.array(ORIGIN)
        f = ijk - np.floor(ijk)
        b = (np.floor(ijk).astype(int) - np.array(I0))
        c = src[b[2]:b[2] + 2, b[1]:b[1] + 2, b[0]:b[0] + 2].astype(np.float64)
        want = (((c[0, 0, 0] * (1 - f[0]) + c[0, 0, 1] * f[0]) * (1 - f[1]) + (c[0, 1, 0] * (1 - f[0]) + c[0, 1, 1] * f[0]) * f[1]) * (1 - f[2])
                + ((c[1, 0, 0] * (1 - f[0]) + c[1, 0, 1] * f[0]) * (1 - f[1]) + (c[1, 1, 0] * (1 - f[0]) + c[1, 1, 1] * f[0]) * f[1]) * f[2])
        got = orc.density_lookup(m, (p - sc.eye).astype(np.float32)) / sig
        assert abs(got - want) < 2e-3 * max(1.0, want)    # float index coordinates: ~1e-5 of a voxel


def test_front_end_errors():
    with pytest.raises(api.VspgError) as e:
        nvdb_scene(filename=os.path.join(GOLD, "does_not_exist.nvdb")).desc
    assert "cannot open" in str(e.value)
    with pytest.raises(api.VspgError) as e:
        nvdb_scene(gridname="temperature").desc
    assert "no grid named" in str(e.value)
    with pytest.raises(api.VspgError) as e:
        nvdb_scene(filename=os.path.join(GOLD, "isg_min.exr")).desc
    assert "not a NanoVDB file" in str(e.value)


# ------------------------------------------------------------------------------------------------ GPU
def _setup(ctx, sc):
    ctx.set_scene(sc)
    ctx.set_guiding(None)
    ctx.set_integrator(api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0))


@pytest.mark.gpu
def test_gpu_density_fetch_on_the_nanovdb_grid_is_bit_exact(ctx, orc):
    sc = nvdb_scene()
    _setup(ctx, sc)
    m = sc.desc.media[0]
    rng = np.random.default_rng(5)
    pts = (rng.uniform(-0.1, 1.1, size=(4000, 3)) - sc.eye).astype(np.float32)
    got = ctx.debug_density(0, pts)
    want = np.array([orc.density_lookup(m, p) for p in pts], dtype=np.float32)
    assert np.array_equal((got[:, 0] + got[:, 3]).view(np.uint32), want.view(np.uint32))
    assert (want > np.float32(0.02) * np.float32(12.4)).sum() > 500      # inside the blob, not only offset-valued background


@pytest.mark.gpu
def test_gpu_majorant_segments_on_the_nanovdb_grid_are_bit_exact(ctx, orc):
    sc = nvdb_scene()
    _setup(ctx, sc)
    m = sc.desc.media[0]
    rays = H.random_rays(np.random.default_rng(11), 4096, sc.eye)
    seg_g, cnt_g = ctx.debug_majorant_segments(0, rays, 200)       # 64^3 grid: up to ~190 cells along a diagonal
    seg_o, cnt_o = orc.majorant_segments(m, rays, 200)
    assert np.array_equal(cnt_g, cnt_o)
    assert cnt_o.max() > 60 and (cnt_o == 0).any()
    for i in range(len(cnt_o)):
        c = min(int(cnt_o[i]), 200)
        assert np.array_equal(seg_g[i, :c].view(np.uint32), seg_o[i, :c].view(np.uint32)), f"ray {i}"


@pytest.mark.gpu
@pytest.mark.parametrize("nee", [0, 1])
def test_gpu_render_of_the_nanovdb_scene_equals_the_oracle(ctx, orc, nee):
    sc = nvdb_scene()
    _setup(ctx, sc)
    ip = api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0, use_nee=nee)
    ctx.set_integrator(ip)
    ctx.film_clear(); ctx.counters_reset()
    ctx.render(0, 8, seed=3)
    film, cnt = ctx.film_download(), ctx.counters()
    ref, cnt_o = orc.render(sc, None, ip, 0, 8, seed=3, nthreads=8)
    assert cnt["paths"] == cnt_o["paths"] and cnt["density_queries"] == cnt_o["density_queries"] and cnt["dda_steps"] == cnt_o["dda_steps"]
    assert np.array_equal(film[..., 3], ref[..., 3])
    assert np.allclose(film[..., :3], ref[..., :3], rtol=1e-12, atol=0)
