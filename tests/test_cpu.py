"""CPU-only tests (`-m "not gpu"`): the oracle against golden vectors / known answers, the host-side logic, and the
C-ABI library surface.  No kernel is launched here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from pyvspg import _abi as A
from pyvspg import api

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


# ------------------------------------------------------------------ C ABI surface
def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "vspg.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(vspg_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(A.SYMBOLS), declared ^ set(A.SYMBOLS)
    lib = api.lib()                                  # binds every symbol; raises if one is missing
    assert lib.vspg_version().decode().startswith("vspg-b200")


def test_struct_sizes_match_header():
    # sizes the C++ side static_asserts / relies on (include/vspg.h)
    assert C.sizeof(A.Mixture) == 10 * 32 * 4 + 3 * 4 + 5 * 4
    assert C.sizeof(A.Neighbours) == 8 * 4 * 4 + 8 * 4
    assert C.sizeof(A.KdNode) == 8
    assert C.sizeof(A.Counters) == 64


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.VspgError) as e:
        api.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_integrator_defaults_follow_the_reference_create():
    ip = api.integrator_defaults()                   # guidedvolpathvspgintegrator.cpp:1263-1319
    assert (ip.max_depth, ip.min_rr_depth, ip.use_nee) == (5, 1, 1)
    assert (ip.surface_guiding, ip.volume_guiding, ip.surface_guiding_ris, ip.volume_guiding_ris) == (1, 1, 1, 0)
    assert (ip.vsp_guiding, ip.vsp_primary_guiding, ip.vsp_secondary_guiding) == (1, 1, 1)
    assert ip.vsp_mis_ratio == 0.5 and ip.vsp_criterion_variance == 1 and ip.vsp_sampling_nds == 0
    assert ip.collision_probability_bias == 0 and ip.light_sampler == A.LIGHTSAMPLER_BVH and ip.regularize == 0
    c_side = A.IntegratorParams()
    api.lib().vspg_integrator_defaults(C.byref(c_side))          # the C ABI's defaults are the same values
    assert bytes(c_side) == bytes(ip)


# ------------------------------------------------------------------ known answers for the integer / polynomial pieces
def test_pcg32_known_answer(orc):
    """PCG32 reference vector (pcg32-demo, seed 42 / sequence 54) -- util/rng.h:29-135 is that generator."""
    out = np.zeros(6, np.uint32)
    orc.lib().orc_pcg32_sequence(54, 42, 1, 6, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert [hex(int(v)) for v in out] == ["0xa15c02b7", "0x7b47f409", "0xba1d3330", "0x83d2f293", "0xbfa4784b", "0xcbed606e"]


def test_pcg32_advance_matches_stepping(orc):
    """util/rng_test.cpp:52-92: Advance(n) == n draws."""
    a, b = np.zeros(40, np.float32), np.zeros(8, np.float32)
    orc.lib().orc_pcg32_floats(7, 1234, 0, 40, a.ctypes.data_as(A.PF))
    orc.lib().orc_pcg32_floats(7, 1234, 32, 8, b.ctypes.data_as(A.PF))      # Advance(32) first
    assert np.array_equal(a[32:], b)
    orc.lib().orc_pcg32_floats(7, 1234, 5 * 65536 + 3, 8, b.ctypes.data_as(A.PF))   # sampleIndex*65536 + dim
    c = np.zeros(5 * 65536 + 3 + 8, np.float32)
    orc.lib().orc_pcg32_floats(7, 1234, 0, len(c), c.ctypes.data_as(A.PF))
    assert np.array_equal(c[-8:], b)


def test_fastexp_tolerance_of_the_reference_test(orc):
    """util/math_test.cpp:365-378: |FastExp(x) - exp(x)| / exp(x) < 3e-4 on [-20, 20] (so __expf is NOT this function)."""
    xs = np.linspace(-20, 20, 4001, dtype=np.float32)
    got = np.array([orc.lib().orc_fastexp(float(x)) for x in xs], dtype=np.float64)
    want = np.exp(xs.astype(np.float64))
    assert (np.abs(got - want) / want).max() < 3e-4
    assert orc.lib().orc_fastexp(0.0) == 1.0


def test_pgl_fastapprox_exp_log_accuracy(orc):
    """embree::fastapprox exp/log (transcendental.h:333-472) are degree-7/9 polynomials: ~1e-7 / 1e-6 relative."""
    xs = np.linspace(-40, 20, 3001, dtype=np.float32)
    got = np.array([orc.lib().orc_pgl_exp(float(x)) for x in xs], dtype=np.float64)
    assert (np.abs(got - np.exp(xs.astype(np.float64))) / np.exp(xs.astype(np.float64))).max() < 5e-7
    ys = np.exp(np.linspace(-10, 10, 2001)).astype(np.float32)
    gl = np.array([orc.lib().orc_pgl_log(float(y)) for y in ys], dtype=np.float64)
    assert np.abs(gl - np.log(ys.astype(np.float64))).max() < 2e-6


def test_henyey_greenstein_normalised_and_sampling_matches_pdf(orc):
    """media_test.cpp:15-97."""
    rng = np.random.default_rng(1)
    for g in (-0.75, -0.3, 0.0, 0.4, 0.9):
        u = rng.random((20000, 2))
        d = np.stack([np.sqrt(1 - (1 - 2 * u[:, 0]) ** 2) * np.cos(2 * np.pi * u[:, 1]),
                      np.sqrt(1 - (1 - 2 * u[:, 0]) ** 2) * np.sin(2 * np.pi * u[:, 1]), 1 - 2 * u[:, 0]], axis=1)
        vals = np.array([orc.lib().orc_hg(float(z), g) for z in d[:, 2]])
        assert abs(vals.mean() * 4 * np.pi - 1) < 0.03            # integrates to one
    wo = np.float32([0, 0, 1])
    wi, pdf = np.zeros(3, np.float32), C.c_float()
    u2 = np.float32([0.3, 0.6])
    orc.lib().orc_sample_hg(wo.ctypes.data_as(A.PF), 0.6, u2.ctypes.data_as(A.PF), wi.ctypes.data_as(A.PF), C.byref(pdf))
    assert abs(np.linalg.norm(wi) - 1) < 1e-5
    assert abs(pdf.value - orc.lib().orc_hg(float(np.dot(wo, wi)), 0.6)) < 1e-6 * pdf.value + 1e-9


# ------------------------------------------------------------------ majorant grid (media.cpp:218-270, 370-409)
@pytest.mark.parametrize("name", ["C2", "C5"])
def test_majorant_grid_host_builder_matches_oracle(orc, name):
    cfg = H.small(name)                              # keeps the density arrays alive
    med = cfg.scene.desc.media[0]
    got, want = api.build_majorant_grid(med), orc.build_majorant_grid(med)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert got.max() > 0 and (got >= 0).all()


def test_majorant_bounds_the_density(orc):
    cfg = H.small("C2")
    med = cfg.scene.desc.media[0]
    maj = api.build_majorant_grid(med).reshape(16, 16, 16)
    rng = np.random.default_rng(2)
    pts = rng.random((4000, 3)).astype(np.float32)
    eye = cfg.scene.eye.astype(np.float32)
    for p in pts[:1500]:
        d = orc.density_lookup(med, p - eye)         # sigma_t at p (render space)
        v = np.minimum((p * 16).astype(int), 15)
        sigma_t = med.sigma_a[0] + med.sigma_s[0]
        assert d <= sigma_t * maj[v[2], v[1], v[0]] * (1 + 1e-6) + 1e-7


# ------------------------------------------------------------------ OpenPGL field loader + oracle pinned to the reference library
@pytest.fixture(scope="module")
def golden_field():
    lib = api.lib()
    h, err = A.P(), C.create_string_buffer(512)
    rc = lib.vspg_host_field_load(os.path.join(GOLD, "field_small.pgl").encode(), C.byref(h), err, 512)
    assert rc == 0, err.value
    yield {"surface": lib.vspg_host_field_surface(h).contents, "volume": lib.vspg_host_field_volume(h).contents}
    lib.vspg_host_field_free(h)


def test_pgl_loader_reads_a_field_written_by_the_reference(golden_field):
    """tests/golden/field_small.pgl was trained and stored by the reference's own OpenPGL (make_golden.py)."""
    s, v = golden_field["surface"], golden_field["volume"]
    assert (s.n_regions, s.n_nodes, v.n_regions, v.n_nodes) == (16, 31, 32, 63)
    for f in (s, v):
        assert f.usable == 1 and f.stochastic_nn == 1 and f.is_nn == 0
        assert list(f.bounds_min) == [-1.0, -1.0, -1.0] and list(f.bounds_max) == [2.0, 2.0, 2.0]
        for i in range(f.n_regions):
            m = f.regions[i]
            n = m.num_components
            assert m.valid == 1 and 1 <= n <= 32
            w = np.array(m.weight[:])
            assert abs(w.sum() - 1) < 1e-4 and (w[n:] == 0).all()
            mu = np.stack([m.mu_x[:n], m.mu_y[:n], m.mu_z[:n]], axis=1)
            assert np.allclose(np.linalg.norm(mu, axis=1), 1, atol=1e-4)
            nb = f.neighbours[i]
            assert nb.ids[0] == i and nb.size == 8     # nearest region point is the region's own


def test_pgl_loader_rejects_foreign_and_truncated_files(tmp_path):
    lib = api.lib()
    h, err = A.P(), C.create_string_buffer(512)
    bad = tmp_path / "x.pgl"
    bad.write_bytes(b"OPENPGL_0.6.0_FIELD\0" + b"\0" * 64)
    assert lib.vspg_host_field_load(str(bad).encode(), C.byref(h), err, 512) == A.ERR_IO and b"0.7.0" in err.value
    data = open(os.path.join(GOLD, "field_small.pgl"), "rb").read()
    bad.write_bytes(data[: len(data) // 2])
    assert lib.vspg_host_field_load(str(bad).encode(), C.byref(h), err, 512) == A.ERR_IO
    assert lib.vspg_host_field_load(b"/nonexistent.pgl", C.byref(h), err, 512) == A.ERR_IO


PROBES = [("surf_cos", 0, "surface", "pos_surface"), ("vol_hg", 1, "volume", "pos"), ("surf_plain", 2, "surface", "pos_surface"),
          ("vol_plain", 2, "volume", "pos")]


@pytest.mark.parametrize("name,mode,which,pos", PROBES)
def test_oracle_guiding_math_is_bit_exact_against_the_reference_library(orc, golden_field, name, mode, which, pos):
    """pgl_probe.npz holds the answers of the reference's OpenPGL (pglFieldInit*SamplingDistribution, ApplyCosineProduct /
    ApplySingleLobeHenyeyGreensteinProduct, PDF, VolumeScatterProbability, IncomingRadiancePDF, SamplePDF) on the golden
    field.  The oracle -- through the product's loader and its rebuilt neighbour tables -- must reproduce them exactly:
    KD leaf, u after draw(), and (because the stochastically selected region feeds them) every pdf bit for bit."""
    G = np.load(os.path.join(GOLD, "pgl_probe.npz"))
    f = golden_field[which]
    leaf, region, ua = orc.cache_lookup(f, G[pos], G["u"])
    ref_id = G[name + "_id"]                             # GetId() = KD leaf data index (Field.h:399), -1 outside the bounds
    assert np.array_equal(leaf, ref_id)
    ok = ref_id >= 0
    assert ok.sum() > 400 and (~ok).sum() > 20
    assert np.array_equal(ua[ok].view(np.uint32), G[name + "_u_after"][ok].view(np.uint32))
    assert (region[ok] != leaf[ok]).any()                # the stochastic neighbour pick was exercised
    o = orc.distribution_probe(f, mode, G[pos], G["u"], G["aux"], G["g"], G["s2"], G["wq"])
    for k, gk in (("pdf", "pdf_wq"), ("vsp", "vsp_wq"), ("li_pdf", "li_pdf_wq"), ("sampled_pdf", "sampled_pdf")):
        assert np.array_equal(o[k][ok].view(np.uint32), G[name + "_" + gk][ok].view(np.uint32)), k
    a, b = o["sampled"][ok].astype(np.float64), G[name + "_sampled"][ok].astype(np.float64)
    # directions: the reference normalises frame axes with rsqrtss + one Newton step (embree), the oracle exactly
    assert np.arccos(np.clip((a * b).sum(1), -1, 1)).max() < 1e-3
    assert ((o["vsp"][ok] >= 0) & (o["vsp"][ok] <= 1)).all()


# ------------------------------------------------------------------ EXR I/O (ISG buffer in, film out)
def _read_channel(path, ch):
    lib = api.lib()
    out, w, h, err = A.PF(), C.c_int(), C.c_int(), C.create_string_buffer(256)
    rc = lib.vspg_exr_read_channel(path.encode(), ch.encode(), C.byref(out), C.byref(w), C.byref(h), err, 256)
    if rc != 0:
        raise api.VspgError(rc, err.value.decode())
    a = np.ctypeslib.as_array(out, shape=(h.value, w.value)).copy()
    lib.vspg_free(out)
    return a


def test_isg_buffer_reader_matches_the_reference_estimate():
    """isg_small.exr was stored by the reference's ImageSpaceGuidingBuffer; isg_small_vsp.npy is what its
    GetVolumeScatterProbabilityEstimate returned per pixel (ImageSpaceGuidingBuffer.h:802-809).  Layer VSP, channel Y."""
    got = _read_channel(os.path.join(GOLD, "isg_small.exr"), "VSP.Y")
    want = np.load(os.path.join(GOLD, "isg_small_vsp.npy"))
    assert got.shape == want.shape == (24, 40)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    with pytest.raises(api.VspgError) as e:
        _read_channel(os.path.join(GOLD, "isg_small.exr"), "Nope.Y")
    assert "not found" in str(e.value)


@pytest.mark.parametrize("half", [0, 1])
def test_exr_film_writer_round_trip(tmp_path, half):
    """R,G,B scanline ZIP file, HALF by default like film.cpp's WriteImage; half conversion rounds to nearest even
    (util/float_test.cpp); cross-checked with the reference's tinyexr when oracle/_ref is present."""
    rgb = (np.random.default_rng(0).random((37, 53, 3)) * 4).astype(np.float32)
    rgb[0, 0] = [0.0, 1e-6, 70000.0]                     # zero, half-subnormal, half-overflow
    path = str(tmp_path / "film.exr")
    assert api.lib().vspg_exr_write_rgb(path.encode(), rgb.ctypes.data_as(A.PF), 53, 37, half) == 0
    with np.errstate(over="ignore"):
        want = rgb.astype(np.float16).astype(np.float32) if half else rgb
    for i, ch in enumerate("RGB"):
        assert np.array_equal(_read_channel(path, ch), want[..., i])
    ref = os.path.join(ROOT, "oracle", "_ref", "libopenpgl_ref.so")
    if os.path.exists(ref):
        L = C.CDLL(ref)
        L.LoadEXR.argtypes = [C.POINTER(A.PF), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, C.POINTER(C.c_char_p)]
        out, w, h, err = A.PF(), C.c_int(), C.c_int(), C.c_char_p()
        assert L.LoadEXR(C.byref(out), C.byref(w), C.byref(h), path.encode(), C.byref(err)) == 0
        m = np.ctypeslib.as_array(out, shape=(37, 53, 4))
        assert np.array_equal(m[..., :3], want)


# ------------------------------------------------------------------ oracle invariants the reference's tests use
def test_sampler_stream_is_a_function_of_pixel_seed_and_sample_index(orc):
    """samplers.h:455-458: the stream depends only on (pixel, seed, sampleIndex) -- what makes sharding exact."""
    a, b, c = (np.zeros(16, np.float32) for _ in range(3))
    orc.lib().orc_sampler_stream(3, 5, 0, 9, 16, a.ctypes.data_as(A.PF))
    orc.lib().orc_sampler_stream(3, 5, 0, 9, 16, b.ctypes.data_as(A.PF))
    orc.lib().orc_sampler_stream(3, 5, 0, 10, 16, c.ctypes.data_as(A.PF))
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert (a >= 0).all() and (a < 1).all()


def test_oracle_render_sharding_composes(orc):
    """Windows and interleaved sample passes sum to the whole film exactly (fp64 accumulation of identical paths)."""
    cfg = H.small("C2", res=24, spp=4)
    whole, cnt = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, 4, seed=3, nthreads=4)
    film = np.zeros_like(whole)
    orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, 4, spp_stride=2, seed=3, nthreads=4, film=film)
    orc.render(cfg.scene, cfg.guiding, cfg.ip, 1, 4, spp_stride=2, seed=3, nthreads=4, film=film)
    assert np.allclose(film, whole, rtol=1e-12, atol=0)
    assert cnt["paths"] == 24 * 24 * 4


def test_vsp_guiding_is_unbiased_in_the_oracle(orc):
    """VSPG on/off and resampling/NDS must converge to the same image (same estimator expectation)."""
    means = {}
    for tag, over in (("off", dict(vsp_guiding=0)), ("resampling", {}), ("nds", dict(vsp_sampling_nds=1))):
        cfg = H.small("C2", res=20, spp=96)
        for k, v in over.items():
            setattr(cfg.ip, k, v)
        film, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, seed=11, nthreads=8)
        px = film[..., :3].mean(axis=2) / cfg.spp
        means[tag] = (px.mean(), px.std(ddof=1) / np.sqrt(px.size) * 3)   # pixels are correlated: generous 3x
    ref, ref_se = means["off"]
    for tag in ("resampling", "nds"):
        m, se = means[tag]
        assert abs(m - ref) <= 4 * np.hypot(se, ref_se) + 0.02 * ref, (tag, means)


def test_gaussian_filter_sampling_is_filter_importance_sampling(orc):
    """pbrt's default pixel filter (gaussian r=1.5 sigma=0.5) is sampled through the tabulated FilterSampler
    (filters.h:26-45): every sample carries weight f/pdf, which for a non-negative filter is the constant
    integral of the tabulated function, and offsets stay inside the radius.  Rendered through the oracle: weight sums per pixel
    are spp * that constant, and the image agrees in expectation with the box-filtered one away from edges."""
    g = H.small("C1G", res=24, spp=32)
    b = H.small("C1", res=24, spp=32)
    fg, _ = orc.render(g.scene, None, g.ip, 0, g.spp, seed=2, nthreads=8)
    fb, _ = orc.render(b.scene, None, b.ip, 0, b.spp, seed=2, nthreads=8)
    w = fg[..., 3] / g.spp
    assert np.allclose(w, w.mean(), rtol=1e-3) and 0.5 < w.mean() < 1.5     # constant weight = table integral
    assert np.array_equal(fb[..., 3], np.full_like(fb[..., 3], b.spp))
    mg = (fg[..., :3] / fg[..., 3:4]).mean(); mb = (fb[..., :3] / fb[..., 3:4]).mean()
    assert abs(mg - mb) < 0.08 * mb, (mg, mb)
