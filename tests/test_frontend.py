"""The .pbrt front end (csrc/host_scene.cpp): the reference's scene-description surface for the guidedvolpath render
pass.  CPU tests compare the POD descs the C++ parser produces with the descs the Python harness builds for the same
workload (written out by pyvspg.pbrt_writer), and the reference's error behaviour (unknown / unused parameters,
unsupported features, missing files => hard errors).  The GPU test renders through both and compares films."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

import helpers as H
from pyvspg import _abi as A
from pyvspg import api, configs, pbrt_writer

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def raw(x):
    return bytes(C.string_at(C.addressof(x), C.sizeof(x)))


def tf(t):
    return (np.array([[t.m[i][j] for j in range(4)] for i in range(4)]), np.array([[t.mInv[i][j] for j in range(4)] for i in range(4)]))


def assert_descs_equal(d, e):
    """d: front end, e: harness.  Geometry, materials, lights, media parameters bit-equal; transforms to 1e-6 relative
    (the two sides invert the same double matrices with different eliminations)."""
    for n in ("n_media", "n_quads", "n_triangles", "n_spheres", "n_materials", "n_lights"):
        assert getattr(d, n) == getattr(e, n), n
    for n, _ in A.Camera._fields_:
        x, y = getattr(d.camera, n), getattr(e.camera, n)
        if isinstance(x, A.Transform):
            for a, b in zip(tf(x), tf(y)):
                np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-12, err_msg=n)
        elif hasattr(x, "__len__"):
            assert list(x) == list(y), n
        else:
            assert x == y, n
    for i in range(d.n_quads):
        assert raw(d.quads[i]) == raw(e.quads[i]), f"quad {i}"
    for i in range(d.n_triangles):
        assert raw(d.triangles[i]) == raw(e.triangles[i]), f"triangle {i}"
    for i in range(d.n_spheres):
        assert raw(d.spheres[i]) == raw(e.spheres[i]), f"sphere {i}"
    for i in range(d.n_materials):
        assert raw(d.materials[i]) == raw(e.materials[i]), f"material {i}"
    for i in range(d.n_lights):
        a, b = d.lights[i], e.lights[i]
        assert (a.type, a.shape_kind, a.shape_index, a.two_sided, a.medium) == (b.type, b.shape_kind, b.shape_index, b.two_sided, b.medium)
        assert list(a.L) == list(b.L)
        if a.type == A.LIGHT_DISTANT:   # the harness passes the direction un-normalised, the front end normalises (lights.cpp:292)
            np.testing.assert_allclose(np.array(a.pos), np.array(b.pos) / np.linalg.norm(b.pos), rtol=2e-7)
        else:
            assert list(a.pos) == list(b.pos)
    for i in range(d.n_media):
        a, b = d.media[i], e.media[i]
        for n in ("type", "sigma_scale", "g", "nx", "ny", "nz"):
            assert getattr(a, n) == getattr(b, n), n
        for n in ("bounds_min", "bounds_max", "sigma_a", "sigma_s", "maj_res"):
            assert list(getattr(a, n)) == list(getattr(b, n)), n
        for x, y in zip(tf(a.render_from_medium), tf(b.render_from_medium)):
            np.testing.assert_allclose(x, y, rtol=1e-6, atol=1e-12)
        nvox = a.nx * a.ny * a.nz
        for n, k in (("density", 1), ("sigma_a_rgb", 3), ("sigma_s_rgb", 3)):
            pa, pb = getattr(a, n), getattr(b, n)
            assert bool(pa) == bool(pb), n
            if pa:
                assert np.array_equal(np.ctypeslib.as_array(pa, shape=(nvox * k,)), np.ctypeslib.as_array(pb, shape=(nvox * k,))), n


def write_exr(path, channels):
    """Minimal uncompressed float scanline OpenEXR writer (test fixture for the Tr-buffer reader)."""
    names = sorted(channels)
    h, w = channels[names[0]].shape
    hdr = struct.pack("<II", 20000630, 2)

    def attr(name, typ, data):
        return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<I", len(data)) + data
    chl = b"".join(n.encode() + b"\0" + struct.pack("<IBBBBII", 2, 0, 0, 0, 0, 1, 1) for n in names) + b"\0"
    hdr += attr("channels", "chlist", chl) + attr("compression", "compression", b"\0")
    box = struct.pack("<iiii", 0, 0, w - 1, h - 1)
    hdr += attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box) + attr("lineOrder", "lineOrder", b"\0")
    hdr += attr("pixelAspectRatio", "float", struct.pack("<f", 1)) + attr("screenWindowCenter", "v2f", struct.pack("<ff", 0, 0))
    hdr += attr("screenWindowWidth", "float", struct.pack("<f", 1)) + b"\0"
    row = 4 * w * len(names)
    off0 = len(hdr) + 8 * h
    body = b"".join(struct.pack("<Q", off0 + y * (8 + row)) for y in range(h))
    for y in range(h):
        body += struct.pack("<iI", y, row) + b"".join(channels[n][y].astype("<f4").tobytes() for n in names)
    with open(path, "wb") as f:
        f.write(hdr + body)


# ------------------------------------------------------------------ descs
@pytest.mark.parametrize("name", ["C1", "C1T", "C1S", "C1E", "C1G", "C2-unguided", "C5-unguided"])
def test_front_end_reproduces_the_harness_descs(name, tmp_path):
    base = name.split("-")[0]
    cfg = H.small(base, grid_n=8)
    if not base.startswith("C1"):   # guiding inputs come from files in the scene language: covered by the next test
        cfg.ip.surface_guiding = cfg.ip.volume_guiding = cfg.ip.vsp_guiding = 0
    path = pbrt_writer.write_scene(cfg, str(tmp_path / "scene.pbrt"), seed=3, raw_density=(base == "C2"))
    hs = api.HostScene(path=path)
    assert_descs_equal(hs.desc, cfg.scene.desc)
    assert raw(hs.ip) == raw(cfg.ip)
    assert (hs.spp, hs.seed, hs.outfile) == (cfg.spp, 3, "out.exr")
    assert hs.guiding is None
    hs.close()


def _guided_scene(tmp_path, nds_plus=False):
    from pyvspg.scenes import Scene, perspective_camera
    cam, eye = perspective_camera([0.5, 0.6, -2.2], [0.5, 0.45, 0.5], [0, 1, 0], 40.0, 40, 24)
    sc = Scene(cam, eye)
    white = sc.add_material([.6, .6, .6])
    sc.add_quad([-2, -0.05, -2], [5, 0, 0], [0, 0, 5], material=white)
    rng = np.random.default_rng(5)
    med = sc.add_grid_medium(rng.random((6, 5, 4), dtype=np.float32), (0, 0, 0), (1, 1, 1), [0.1, 0.1, 0.1], [6.0, 6.0, 6.0], 8.0, 0.6)
    sc.add_box([0, 0, 0], [1, 1, 1], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    sc.add_point_light([0.2, 1.8, -0.4], [9.0, 8.0, 7.0])
    ip = api.integrator_defaults(light_sampler=A.LIGHTSAMPLER_POWER, vsp_sampling_nds=int(nds_plus), collision_probability_bias=int(nds_plus))
    cfg = configs.Config("frontend", sc.finalize(), None, ip, 3, "front-end fixture")
    tr = None
    if nds_plus:
        tr = rng.random((24, 40, 3), dtype=np.float32)
        write_exr(str(tmp_path / "tr.exr"), {"Transmittance.R": tr[..., 0], "Transmittance.G": tr[..., 1], "Transmittance.B": tr[..., 2]})
    return cfg, tr


@pytest.mark.parametrize("nds_plus", [False, True])
def test_front_end_loads_cache_isg_and_tr_buffers_named_by_the_integrator(tmp_path, nds_plus):
    cfg, tr = _guided_scene(tmp_path, nds_plus)
    path = pbrt_writer.write_scene(cfg, str(tmp_path / "scene.pbrt"), integrator_name="guidedvolpath",   # the name README / BASELINE use
                                   cache_file=os.path.join(GOLD, "field_small.pgl"), isg_file=os.path.join(GOLD, "isg_small.exr"),
                                   tr_file="tr.exr" if nds_plus else None)
    hs = api.HostScene(path=path)
    assert_descs_equal(hs.desc, cfg.scene.desc)
    assert raw(hs.ip) == raw(cfg.ip)
    g = hs.guiding.desc
    # the same field through the stand-alone loader
    lib = api.lib()
    h, err = A.P(), C.create_string_buffer(512)
    assert lib.vspg_host_field_load(os.path.join(GOLD, "field_small.pgl").encode(), C.byref(h), err, 512) == 0
    for got, want in ((g.surface.contents, lib.vspg_host_field_surface(h).contents), (g.volume.contents, lib.vspg_host_field_volume(h).contents)):
        assert (got.n_nodes, got.n_regions, got.usable, got.stochastic_nn) == (want.n_nodes, want.n_regions, want.usable, want.stochastic_nn)
        assert raw(got.regions[got.n_regions - 1]) == raw(want.regions[want.n_regions - 1])
        assert raw(got.nodes[got.n_nodes - 1]) == raw(want.nodes[want.n_nodes - 1])
    lib.vspg_host_field_free(h)
    assert (g.isg_width, g.isg_height) == (40, 24)
    isg = np.ctypeslib.as_array(g.isg_vsp, shape=(24, 40))
    assert np.array_equal(isg.view(np.uint32), np.load(os.path.join(GOLD, "isg_small_vsp.npy")).view(np.uint32))
    if nds_plus:
        assert (g.tr_width, g.tr_height) == (40, 24)
        assert np.array_equal(np.ctypeslib.as_array(g.tr_rgb, shape=(24, 40, 3)), tr)
    else:
        assert not g.tr_rgb
    hs.close()


# ------------------------------------------------------------------ the language itself
HEAD = """LookAt 0 0 -3  0 0 0  0 1 0
Camera "perspective" "float fov" 40
Film "rgb" "integer xresolution" [16] "integer yresolution" [8] "string filename" "a.exr"
Sampler "independent" "integer pixelsamples" 2
PixelFilter "box"
Integrator "guidedvolpath" "bool vspguiding" false "bool surfaceguiding" false "bool volumeguiding" false "string lightsampler" "power"
"""


def test_transforms_includes_comments_named_objects(tmp_path):
    (tmp_path / "geo.pbrt").write_text('Shape "trianglemesh" "point3 P" [0 0 0  1 0 0  0 1 0  1 1 0] "integer indices" [0 1 2  2 1 3]  # two triangles\n')
    text = HEAD + """WorldBegin   # comment after a directive
MakeNamedMedium "fog" "string type" "homogeneous" "rgb sigma_a" [1 2 3] "rgb sigma_s" [4 5 6] "float scale" 0.5 "float g" [0.25]
MakeNamedMaterial "red" "string type" "diffuse" "rgb reflectance" [0.9 0.1 0.1]
AttributeBegin
  Translate 1 2 3
  Scale 2 2 2
  NamedMaterial "red"
  Include "geo.pbrt"
AttributeEnd
AttributeBegin
  Rotate 90 0 0 1
  Material "diffuse" "float reflectance" 0.25
  ReverseOrientation
  AreaLightSource "diffuse" "rgb L" [2 3 4] "float scale" 2 "bool twosided" true
  Shape "sphere" "float radius" 0.5
AttributeEnd
AttributeBegin
  MediumInterface "fog" ""
  Material "interface"
  Transform [1 0 0 0  0 1 0 0  0 0 1 0  5 6 7 1]
  Shape "bilinearmesh" "point3 P" [0 0 0  1 0 0  0 1 0  1 1 0]
AttributeEnd
LightSource "distant" "point3 from" [0 10 0] "point3 to" [0 0 0] "rgb L" [1 1 1] "float scale" 3
LightSource "infinite"
"""
    hs = api.HostScene(text=text, base_dir=str(tmp_path))
    d = hs.desc
    assert (d.n_triangles, d.n_spheres, d.n_quads, d.n_media, d.n_materials, d.n_lights) == (2, 1, 1, 1, 2, 3)
    eye = np.array([0, 0, -3.0])
    t0, t1 = d.triangles[0], d.triangles[1]
    assert np.allclose(t0.p1, np.array([3, 2, 3]) - eye) and np.allclose(t1.p2, np.array([3, 4, 3]) - eye) and np.allclose(t1.p0, np.array([1, 4, 3]) - eye)
    assert t0.material == 0 and list(d.materials[0].reflectance) == [np.float32(0.9), np.float32(0.1), np.float32(0.1)] and t0.has_interface == 0
    s = d.spheres[0]
    assert np.allclose(s.center, -eye) and s.radius == 0.5 and s.flip_normal == 1 and s.material == 1 and s.area_light == 0
    assert list(d.materials[1].reflectance) == [0.25, 0.25, 0.25]
    al = d.lights[0]
    assert (al.type, al.shape_kind, al.shape_index, al.two_sided) == (A.LIGHT_DIFFUSE_AREA, 2, 0, 1) and list(al.L) == [2, 3, 4] and al.scale == 2   # spectrum and scale stay apart (lights.h LightBase)
    q = d.quads[0]
    assert (q.material, q.medium_inside, q.medium_outside, q.has_interface) == (-1, 0, -1, 1)
    assert np.allclose(q.p11, np.array([6, 7, 7]) - eye)   # column-major Transform: translation in the last four numbers
    m = d.media[0]
    assert list(m.sigma_a) == [0.5, 1.0, 1.5] and list(m.sigma_s) == [2.0, 2.5, 3.0] and m.g == 0.25 and m.type == A.MEDIUM_HOMOGENEOUS
    dl, il = d.lights[1], d.lights[2]
    assert dl.type == A.LIGHT_DISTANT and np.allclose(dl.pos, [0, 1, 0]) and list(dl.L) == [1, 1, 1] and dl.scale == 3
    assert il.type == A.LIGHT_UNIFORM_INFINITE and list(il.L) == [1, 1, 1] and il.scale == 1
    assert (d.camera.width, d.camera.height) == (16, 8) and d.camera.filter_type == 0 and hs.spp == 2 and hs.outfile == "a.exr"
    hs.close()


def test_spot_light_frame_cone_and_power():
    """LightSource "spot" (lights.cpp:1505-1545): position and axis through the CTM, the render -> light rotation the inverse of the
    light -> render one, cosines of coneangle / coneangle - conedeltaangle, and "power" turned into a scale by
    2 Pi ((1 - cosStart) + (cosStart - cosEnd) / 2).  (Bit-level agreement with the reference: golden scenes cs / csp.)"""
    text = HEAD + """WorldBegin
AttributeBegin
  Translate 1 0 0
  Rotate 90 0 1 0
  LightSource "spot" "point3 from" [0 2 0] "point3 to" [0 1 1] "rgb I" [1 2 3] "float coneangle" 40 "float conedeltaangle" 10 "float power" 50 "float scale" 2
AttributeEnd
LightSource "spot"
"""
    hs = api.HostScene(text=text)
    d = hs.desc
    eye = np.array([0, 0, -3.0])
    l = d.lights[0]
    assert l.type == A.LIGHT_SPOT and list(l.L) == [1, 2, 3] and l.medium == -1
    assert np.allclose(l.pos, np.array([1, 2, 0]) - eye, atol=1e-6)            # Rotate 90 about y leaves (0, 2, 0) in place; then the translation
    ax = np.array([0, -1, 1]) / np.sqrt(2)
    assert np.allclose(l.axis, [ax[2], ax[1], -ax[0]], atol=1e-6)              # the rotation takes +z to +x
    R = np.array(l.light_from_render, np.float64).reshape(3, 3)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-6) and np.allclose(R @ np.array(l.axis), [0, 0, 1], atol=1e-6)
    cs, ce = np.cos(np.radians(30.0)), np.cos(np.radians(40.0))
    assert np.isclose(l.cos_falloff_start, cs, rtol=1e-6) and np.isclose(l.cos_falloff_end, ce, rtol=1e-6)
    assert np.isclose(l.scale, 2 * 50 / (2 * np.pi * ((1 - cs) + (cs - ce) / 2)), rtol=1e-5)
    m = d.lights[1]                                                             # defaults: at the origin, along +z, 30 / 5 degrees
    assert np.allclose(m.pos, -eye) and np.allclose(m.axis, [0, 0, 1]) and m.scale == 1 and list(m.L) == [1, 1, 1]
    assert np.isclose(m.cos_falloff_end, np.cos(np.radians(30.0)), rtol=1e-6) and np.isclose(m.cos_falloff_start, np.cos(np.radians(25.0)), rtol=1e-6)
    hs.close()


def test_medium_emission_parameters():
    """"Le" / "Lescale" (media.cpp:188-194, 304-329): homogeneous media fold the float "Lescale" into Le; uniformgrid media keep Le
    and take "Lescale" as a grid (without one: the reference's 1x1x1 grid of 1); an all-zero Le is no emission."""
    text = HEAD + """WorldBegin
MakeNamedMedium "a" "string type" "homogeneous" "rgb Le" [1 2 3] "float Lescale" 0.5
MakeNamedMedium "b" "string type" "uniformgrid" "float density" [1 2 3 4] "integer nx" 2 "integer ny" 2 "rgb Le" [4 5 6] "float Lescale" [0 1 2 3]
MakeNamedMedium "c" "string type" "uniformgrid" "float density" [1 2] "integer nx" 2 "rgb Le" [7 8 9]
MakeNamedMedium "d" "string type" "homogeneous" "rgb Le" [0 0 0] "float Lescale" 5
"""
    hs = api.HostScene(text=text)
    a, b, c, d = (hs.desc.media[i] for i in range(4))
    assert list(a.Le) == [0.5, 1.0, 1.5] and not a.le_scale
    assert list(b.Le) == [4, 5, 6] and (b.le_nx, b.le_ny, b.le_nz) == (2, 2, 1) and [b.le_scale[i] for i in range(4)] == [0, 1, 2, 3]
    assert list(c.Le) == [7, 8, 9] and (c.le_nx, c.le_ny, c.le_nz) == (1, 1, 1) and c.le_scale[0] == 1
    assert list(d.Le) == [0, 0, 0] and not d.le_scale
    hs.close()


@pytest.mark.parametrize("body,code,needle", [
    ("WorldBegin\nMakeNamedMedium \"m\" \"string type\" \"uniformgrid\" \"float density\" [1 2] \"integer nx\" 2 \"rgb Le\" [1 1 1] \"float Lescale\" [1 2 3]\n", A.ERR_INVALID, "values for \"Lescale\""),
    ("WorldBegin\nMakeNamedMedium \"m\" \"string type\" \"rgbgrid\" \"rgb sigma_a\" [1 1 1] \"rgb Le\" [1 1 1]\n", A.ERR_UNSUPPORTED, "homogeneous and uniformgrid"),
    ("WorldBegin\nLightSource \"spot\" \"point3 from\" [1 1 1] \"point3 to\" [1 1 1]\n", A.ERR_INVALID, "from == to"),
    ("WorldBegin\nLightSource \"spot\" \"float conedeltaangle\" -1\n", A.ERR_INVALID, "conedeltaangle"),
    ("WorldBegin\nLightSource \"goniometric\"\n", A.ERR_UNSUPPORTED, "goniometric"),
    ("WorldBegin\nShape \"sphere\" \"float radious\" 1\n", A.ERR_INVALID, "unused"),                       # paramdict.cpp:642-664
    ("WorldBegin\nShape \"cylinder\"\n", A.ERR_UNSUPPORTED, "cylinder"),
    ("WorldBegin\nMaterial \"dielectric\"\n", A.ERR_UNSUPPORTED, "dielectric"),
    ("WorldBegin\nFrobnicate 1 2 3\n", A.ERR_INVALID, "unknown directive"),
    ("WorldBegin\nAttributeEnd\n", A.ERR_INVALID, "unmatched"),
    ("WorldBegin\nNamedMaterial \"nope\"\n", A.ERR_INVALID, "not defined"),
    ("WorldBegin\nAttributeBegin\nMediumInterface \"nope\" \"\"\nShape \"sphere\"\nAttributeEnd\n", A.ERR_INVALID, "not defined"),
    ("WorldBegin\nMakeNamedMedium \"m\" \"string type\" \"uniformgrid\" \"float density\" [1 2 3] \"integer nx\" 2\n", A.ERR_INVALID, "expected nx*ny*nz"),
    ("WorldBegin\nMakeNamedMedium \"m\" \"string type\" \"nanovdb\" \"string filename\" \"a.nvdb\"\n", A.ERR_IO, "cannot open"),
    ("WorldBegin\nMakeNamedMedium \"m\" \"string type\" \"nanovdb\"\n", A.ERR_INVALID, "Must supply \"filename\""),
    ("WorldBegin\nMakeNamedMedium \"m\" \"string type\" \"cloud\"\n", A.ERR_UNSUPPORTED, "outside the hot-path scope"),
    ("WorldBegin\nShape \"sphere\" \"float radius\" [1\n", A.ERR_INVALID, "missing ]"),
    ("Camera \"perspective\"\nWorldBegin\n", A.ERR_INVALID, "not allowed after WorldBegin"),
])
def test_errors_mirror_the_reference_error_exits(body, code, needle):
    text = HEAD + body if not body.startswith("Camera") else HEAD + "WorldBegin\n" + body
    with pytest.raises(api.VspgError) as e:
        api.HostScene(text=text)
    assert e.value.code == code and needle in str(e.value), str(e.value)
    assert "<string>:" in str(e.value)   # file:line like the reference's FileLoc


@pytest.mark.parametrize("integrator,code,needle", [
    ('Integrator "guidedvolpath" "bool vspguidin" true', A.ERR_INVALID, "unused"),
    ('Integrator "volpath"', A.ERR_UNSUPPORTED, "guidedvolpath"),
    ('Integrator "guidedvolpath" "bool storeGuidingCache" true "string guidingCacheFileName" "x.pgl"', A.ERR_UNSUPPORTED, "training"),
    ('Integrator "guidedvolpath"', A.ERR_UNSUPPORTED, "loadGuidingCache"),                       # guiding on, nothing to load: no training here
    ('Integrator "guidedvolpath" "bool loadGuidingCache" true "string guidingCacheFileName" "missing.pgl"', A.ERR_IO, "missing.pgl"),
    ('Integrator "guidedvolpath" "string vspsamplingmethod" "magic"', A.ERR_INVALID, "vspsamplingmethod"),
    ('Integrator "guidedvolpath" "bool storeISGBuffer" true "bool vspguiding" false "string isgBufferFileName" "x.exr"', A.ERR_UNSUPPORTED, "stores no VSP layer"),
    ('Integrator "guidedvolpath" "bool storeISGBuffer" true', A.ERR_INVALID, "without isgBufferFileName"),
    ('Integrator "guidedvolpath" "bool rrguiding" true', A.ERR_UNSUPPORTED, "rrguiding"),
])
def test_integrator_parameter_errors(integrator, code, needle):
    text = HEAD.replace(HEAD.splitlines()[5], integrator) + "WorldBegin\n"
    with pytest.raises(api.VspgError) as e:
        api.HostScene(text=text)
    assert e.value.code == code and needle in str(e.value), str(e.value)


def test_missing_scene_file_is_an_io_error():
    with pytest.raises(api.VspgError) as e:
        api.HostScene(path="/nonexistent/scene.pbrt")
    assert e.value.code == A.ERR_IO


# ------------------------------------------------------------------ GPU: the two routes render the same film
@pytest.mark.gpu
@pytest.mark.parametrize("nds_plus", [False, True])
def test_render_through_the_front_end_equals_render_through_the_harness(tmp_path, nds_plus):
    from pyvspg import synthetic as syn
    cfg, tr = _guided_scene(tmp_path, nds_plus)
    path = pbrt_writer.write_scene(cfg, str(tmp_path / "scene.pbrt"), seed=11, cache_file=os.path.join(GOLD, "field_small.pgl"),
                                   isg_file=os.path.join(GOLD, "isg_small.exr"), tr_file="tr.exr" if nds_plus else None)
    hs = api.HostScene(path=path)
    films = []
    for scene, guiding, ip in ((hs, hs.guiding, hs.ip), (cfg.scene, None, cfg.ip)):
        ctx = api.Context(0)
        ctx.set_scene(scene)
        if guiding is None:   # harness route: the same inputs through the stand-alone loaders
            lib = api.lib()
            h, err = A.P(), C.create_string_buffer(512)
            assert lib.vspg_host_field_load(os.path.join(GOLD, "field_small.pgl").encode(), C.byref(h), err, 512) == 0
            guiding = syn.Guiding()
            guiding.set_field("surface", lib.vspg_host_field_surface(h).contents, h)
            guiding.set_field("volume", lib.vspg_host_field_volume(h).contents, h)
            guiding.set_isg(api.exr_read_channel(os.path.join(GOLD, "isg_small.exr"), "VSP.Y"))
            if nds_plus:
                guiding.set_tr(tr)
        ctx.set_guiding(guiding)
        ctx.set_integrator(ip)
        ctx.render(0, hs.spp, seed=hs.seed)
        films.append(ctx.film_download())
        assert ctx.counters()["paths"] == 40 * 24 * hs.spp
        ctx.close()
    assert np.array_equal(films[0][..., 3], films[1][..., 3])
    assert films[0][..., :3].sum() > 0
    np.testing.assert_allclose(films[0], films[1], rtol=1e-6, atol=1e-9)
    hs.close()


def test_plymesh_reader_errors(tmp_path):
    """Shape "plymesh" (util/mesh.cpp:322-420): the reference's error exits, plus what this path leaves out by name."""
    def parse(body):
        return api.HostScene(text=HEAD + "WorldBegin\n" + body, base_dir=str(tmp_path))
    with pytest.raises(api.VspgError) as e:
        parse('Shape "plymesh" "string filename" "nope.ply"\n')
    assert e.value.code == A.ERR_IO and "Couldn't open PLY file" in str(e.value)
    (tmp_path / "n.ply").write_text("ply\nformat ascii 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\nproperty float nx\n"
                                    "property float ny\nproperty float nz\nelement face 1\nproperty list uchar int vertex_indices\nend_header\n"
                                    "0 0 0 0 0 1\n1 0 0 0 0 1\n0 1 0 0 0 1\n3 0 1 2\n")
    with pytest.raises(api.VspgError) as e:
        parse('Shape "plymesh" "string filename" "n.ply"\n')
    assert e.value.code == A.ERR_UNSUPPORTED and "per-vertex normals" in str(e.value)
    (tmp_path / "p.ply").write_text("ply\nformat ascii 1.0\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\nelement face 1\n"
                                    "property list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n1 1 0\n0 1 0\n.5 .5 1\n5 0 1 2 3 4\n")
    with pytest.raises(api.VspgError) as e:
        parse('Shape "plymesh" "string filename" "p.ply"\n')
    assert "Only triangles and quads" in str(e.value)
    (tmp_path / "i.ply").write_text("ply\nformat ascii 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\nelement face 1\n"
                                    "property list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n0 1 0\n3 0 1 7\n")
    with pytest.raises(api.VspgError) as e:
        parse('Shape "plymesh" "string filename" "i.ply"\n')
    assert "out of bounds" in str(e.value)
    (tmp_path / "ok.ply").write_text("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\nelement face 2\n"
                                     "property list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n1 1 0\n0 1 0\n3 0 1 2\n4 0 1 2 3\n")
    hs = parse('Shape "plymesh" "string filename" "ok.ply"\n')
    assert (hs.desc.n_triangles, hs.desc.n_quads) == (1, 1)
    hs.close()
