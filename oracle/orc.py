"""TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
ctypes binding of liboracle.so (oracle/vspg_oracle.cpp), the CPU restatement of the reference's render pass.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The POD struct definitions are shared with the product harness (pyvspg._abi) because both sides of the parity
check consume the same include/vspg.h descriptions.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
from pyvspg import _abi as A  # noqa: E402

LIB_PATH = os.path.join(HERE, "liboracle.so")
_lib = None
P, PF, i32 = C.c_void_p, A.PF, C.c_int
PI = C.POINTER(C.c_int)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `make -C oracle liboracle.so`")
        L = C.CDLL(LIB_PATH)
        L.orc_set_hg_table.argtypes = [PF, i32]
        L.orc_pcg32_sequence.argtypes = [C.c_uint64, C.c_uint64, i32, i32, C.POINTER(C.c_uint32)]
        L.orc_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, i32, PF]
        L.orc_murmur.argtypes = [C.c_char_p, C.c_size_t, C.c_uint64]
        L.orc_murmur.restype = C.c_uint64
        L.orc_mixbits.argtypes = [C.c_uint64]
        L.orc_mixbits.restype = C.c_uint64
        for n in ("orc_fastexp", "orc_pgl_exp", "orc_pgl_log"):
            getattr(L, n).argtypes = [C.c_float]
            getattr(L, n).restype = C.c_float
        L.orc_hg.argtypes = [C.c_float, C.c_float]
        L.orc_hg.restype = C.c_float
        L.orc_sample_hg.argtypes = [PF, C.c_float, PF, PF, PF]
        L.orc_sampler_stream.argtypes = [i32, i32, i32, i32, i32, PF]
        L.orc_build_majorant_grid.argtypes = [C.POINTER(A.Medium), PF]
        L.orc_density_lookup.argtypes = [C.POINTER(A.Medium), PF]
        L.orc_density_lookup.restype = C.c_float
        L.orc_majorant_segments.argtypes = [C.POINTER(A.Medium), i32, PF, i32, C.POINTER(A.Segment), PI]
        L.orc_cache_lookup.argtypes = [C.POINTER(A.Field), i32, PF, PF, PI, PI, PF]
        L.orc_distribution_probe.argtypes = [C.POINTER(A.Field), i32, i32, PF, PF, PF, PF, PF, PF, PI, PF, PF, PF, PF, PF]
        L.orc_guided_sample.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(A.GuidingDesc), C.POINTER(A.IntegratorParams), i32, i32,
                                        C.POINTER(A.GsInput), C.POINTER(A.GsOutput)]
        L.orc_sample_distance.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(A.GuidingDesc), C.POINTER(A.IntegratorParams), i32,
                                          C.POINTER(A.SdInput), C.POINTER(A.SdOutput)]
        L.orc_render.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(A.GuidingDesc), C.POINTER(A.IntegratorParams), i32, i32, i32,
                                 i32, i32, i32, i32, i32, i32, P, C.POINTER(C.c_longlong)]
        L.orc_render_pixel_sample.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(A.GuidingDesc), C.POINTER(A.IntegratorParams), i32, i32,
                                              i32, i32, PF]
        L.orc_camera_ray.argtypes = [C.POINTER(A.SceneDesc), i32, i32, i32, i32, PF, PF, PI]
        tbl = np.fromfile(os.path.join(ROOT, "tests", "golden", "hg_vmm_table.bin"), dtype=np.float32)
        assert tbl.size == 128 * 18
        rc = L.orc_set_hg_table(tbl.ctypes.data_as(PF), 128)
        assert rc == 0
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(PF)


def _gp(guiding):
    return C.byref(guiding.desc) if guiding is not None else None


def majorant_segments(medium, rays7, max_segments=64):
    rays7 = np.ascontiguousarray(rays7, dtype=np.float32).reshape(-1, 7)
    n = rays7.shape[0]
    seg = (A.Segment * (n * max_segments))()
    cnt = np.zeros(n, dtype=np.int32)
    lib().orc_majorant_segments(C.byref(medium), n, _fp(rays7), max_segments, seg, cnt.ctypes.data_as(PI))
    return np.frombuffer(seg, dtype=np.float32).reshape(n, max_segments, 5).copy(), cnt


def build_majorant_grid(medium):
    out = np.zeros(medium.maj_res[0] * medium.maj_res[1] * medium.maj_res[2], dtype=np.float32)
    lib().orc_build_majorant_grid(C.byref(medium), _fp(out))
    return out


def density_lookup(medium, p):
    p = np.ascontiguousarray(p, dtype=np.float32)
    return float(lib().orc_density_lookup(C.byref(medium), _fp(p)))


def cache_lookup(field, pos3, u):
    pos3 = np.ascontiguousarray(pos3, dtype=np.float32).reshape(-1, 3)
    u = np.ascontiguousarray(u, dtype=np.float32)
    n = pos3.shape[0]
    leaf, region, uo = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
    lib().orc_cache_lookup(C.byref(field), n, _fp(pos3), _fp(u), leaf.ctypes.data_as(PI), region.ctypes.data_as(PI), _fp(uo))
    return leaf, region, uo


def distribution_probe(field, mode, pos3, u, aux3, g, s2, wq3):
    pos3, aux3, wq3 = (np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 3) for a in (pos3, aux3, wq3))
    u, g = np.ascontiguousarray(u, dtype=np.float32), np.ascontiguousarray(g, dtype=np.float32)
    s2 = np.ascontiguousarray(s2, dtype=np.float32).reshape(-1, 2)
    n = pos3.shape[0]
    region = np.zeros(n, np.int32)
    pdf, vsp, lipdf, spdf = (np.zeros(n, np.float32) for _ in range(4))
    sampled = np.zeros((n, 3), np.float32)
    rc = lib().orc_distribution_probe(C.byref(field), mode, n, _fp(pos3), _fp(u), _fp(aux3), _fp(g), _fp(s2), _fp(wq3),
                                      region.ctypes.data_as(PI), _fp(pdf), _fp(vsp), _fp(lipdf), _fp(sampled), _fp(spdf))
    assert rc == 0, rc
    return dict(region=region, pdf=pdf, vsp=vsp, li_pdf=lipdf, sampled=sampled, sampled_pdf=spdf)


def light_sample(scene, ip, pos3, ns3, u, light_of):
    pos3 = np.ascontiguousarray(pos3, dtype=np.float32).reshape(-1, 3)
    ns3 = np.ascontiguousarray(ns3, dtype=np.float32).reshape(-1, 3)
    u = np.ascontiguousarray(u, dtype=np.float32)
    light_of = np.ascontiguousarray(light_of, dtype=np.int32)
    n = pos3.shape[0]
    light, pmf, pmf_of = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros(n, np.float32)
    rc = lib().orc_light_sample(C.byref(scene.desc), C.byref(ip), n, _fp(pos3), _fp(ns3), _fp(u), light_of.ctypes.data_as(PI),
                                light.ctypes.data_as(PI), _fp(pmf), _fp(pmf_of))
    assert rc == 0, rc
    return light, pmf, pmf_of


def intersect(scene, ip, rays7):
    rays7 = np.ascontiguousarray(rays7, dtype=np.float32).reshape(-1, 7)
    n = rays7.shape[0]
    prim, out = np.zeros(n, np.int32), np.zeros((n, 10), np.float32)
    rc = lib().orc_intersect(C.byref(scene.desc), C.byref(ip), n, _fp(rays7), prim.ctypes.data_as(PI), _fp(out))
    assert rc == 0, rc
    return prim, out


def sample_li(scene, ip, ctx9, light, u2):
    ctx9 = np.ascontiguousarray(ctx9, dtype=np.float32).reshape(-1, 9)
    light = np.ascontiguousarray(light, dtype=np.int32)
    u2 = np.ascontiguousarray(u2, dtype=np.float32).reshape(-1, 2)
    n = ctx9.shape[0]
    out = np.zeros((n, 12), np.float32)
    rc = lib().orc_sample_li(C.byref(scene.desc), C.byref(ip), n, _fp(ctx9), light.ctypes.data_as(PI), _fp(u2), _fp(out))
    assert rc == 0, rc
    return out


def guided_sample(scene, guiding, ip, is_surface, inputs):
    n = len(inputs)
    out = (A.GsOutput * n)()
    rc = lib().orc_guided_sample(C.byref(scene.desc), _gp(guiding), C.byref(ip), int(is_surface), n, inputs, out)
    assert rc == 0, rc
    return out


def sample_distance(scene, guiding, ip, inputs):
    n = len(inputs)
    out = (A.SdOutput * n)()
    rc = lib().orc_sample_distance(C.byref(scene.desc), _gp(guiding), C.byref(ip), n, inputs, out)
    assert rc == 0, rc
    return out


def render(scene, guiding, ip, spp_begin, spp_end, spp_stride=1, window=None, seed=0, nthreads=1, film=None):
    cam = scene.desc.camera
    x0, y0, x1, y1 = window if window is not None else (0, 0, cam.width, cam.height)
    if film is None:
        film = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
    cnt = (C.c_longlong * 5)()
    rc = lib().orc_render(C.byref(scene.desc), _gp(guiding), C.byref(ip), spp_begin, spp_end, spp_stride, x0, y0, x1, y1, int(seed),
                          int(nthreads), film.ctypes.data_as(P), cnt)
    assert rc == 0, rc
    return film, dict(paths=cnt[0], density_queries=cnt[1], dda_steps=cnt[2], cache_lookups=cnt[3], shadow_rays=cnt[4])


def render_isg(scene, guiding, ip, spp, seed=0, nthreads=1):
    """The wave loop with the ISG buffer trained on the fly -> film, vsp [h, w], records [h, w, 8]."""
    cam = scene.desc.camera
    film = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
    vsp = np.zeros((cam.height, cam.width), dtype=np.float32)
    rec = np.zeros((cam.height, cam.width, 8), dtype=np.float32)
    rc = lib().orc_render_isg(C.byref(scene.desc), _gp(guiding), C.byref(ip), int(spp), int(seed), int(nthreads), film.ctypes.data_as(P), _fp(vsp), _fp(rec))
    assert rc == 0, rc
    return film, vsp, rec


def render_tr(scene, guiding, ip, spp_begin, spp_end, seed=0, nthreads=1):
    """Render with the storeTrBuffer pre-pass on: returns film, counters, tr_rgb [h, w, 3] float32, tr_spp [h, w] int32."""
    cam = scene.desc.camera
    film = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
    tr = np.zeros((cam.height, cam.width, 3), dtype=np.float32)
    n = np.zeros((cam.height, cam.width), dtype=np.int32)
    cnt = (C.c_longlong * 5)()
    rc = lib().orc_render_ex(C.byref(scene.desc), _gp(guiding), C.byref(ip), spp_begin, spp_end, 1, 0, 0, cam.width, cam.height, int(seed),
                             int(nthreads), film.ctypes.data_as(P), cnt, _fp(tr), n.ctypes.data_as(PI))
    assert rc == 0, rc
    return film, dict(paths=cnt[0], density_queries=cnt[1]), tr, n


def camera_ray(scene, px, py, sample_index, seed=0):
    o, d, ch = np.zeros(3, np.float32), np.zeros(3, np.float32), C.c_int()
    lib().orc_camera_ray(C.byref(scene.desc), px, py, sample_index, seed, _fp(o), _fp(d), C.byref(ch))
    return o, d, ch.value
