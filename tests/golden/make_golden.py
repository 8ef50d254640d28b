"""Generates the committed golden fixtures from the REFERENCE's own OpenPGL build (oracle/_ref, built by
oracle/build_ref.sh from /root/reference; this script only runs where /root/reference exists):

  tests/golden/field_small.pgl    a real OpenPGL 0.7.0 field trained through the reference's public API
                                   (SampleStorage::AddSample + Field::Update) and stored with Field::Store
  tests/golden/isg_small.exr      an image-space guiding buffer accumulated, updated and stored by the reference's
                                   ImageSpaceGuidingBuffer (multi-layer float EXR via its tinyexr path)
  tests/golden/isg_small_vsp.npy  GetVolumeScatterProbabilityEstimate() of that buffer for every pixel
  tests/golden/pgl_probe.npz      reference answers on seeded probes of that field: region id, u after draw(),
                                   PDF / VSP / IncomingRadiancePDF at a query direction, SamplePDF (direction + pdf)
                                   for surface (+cosine product) and volume (+HG product) distributions

Run:  python tests/golden/make_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "libpglharness.so")
PF, PI = C.POINTER(C.c_float), C.POINTER(C.c_int)


def harness():
    if not os.path.exists(REF):
        raise SystemExit(f"{REF} missing: run oracle/build_ref.sh (needs /root/reference)")
    L = C.CDLL(REF)
    L.ref_field_train.argtypes = [C.c_char_p, PF, PF, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    L.ref_field_load.argtypes = [C.c_char_p]
    L.ref_field_load.restype = C.c_void_p
    L.ref_field_free.argtypes = [C.c_void_p]
    L.ref_field_iteration.argtypes = [C.c_void_p]
    L.ref_isgb_make.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, PF]
    L.ref_distribution_probe.argtypes = [C.c_void_p, C.c_int, C.c_int, PF, PF, PF, PF, PF, PF, PI, PF, PF, PF, PF, PF, PF]
    return L


def fp(a):
    return a.ctypes.data_as(PF)


def unit(rng, n):
    v = rng.normal(size=(n, 3))
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)


def probe(L, h, mode, pos, u, aux, g, s2, wq):
    n = len(pos)
    rid = np.zeros(n, np.int32)
    ua, pdf, vsp, li, spdf = (np.zeros(n, np.float32) for _ in range(5))
    sdir = np.zeros((n, 3), np.float32)
    L.ref_distribution_probe(h, mode, n, fp(pos), fp(u), fp(aux), fp(g), fp(s2), fp(wq), rid.ctypes.data_as(PI), fp(ua), fp(pdf),
                             fp(vsp), fp(li), fp(sdir), fp(spdf))
    return dict(id=rid, u_after=ua, pdf_wq=pdf, vsp_wq=vsp, li_pdf_wq=li, sampled=sdir, sampled_pdf=spdf)


def main():
    # The probes are answered by the reference library running with the path's portable libm (include/vspg_libm.h through
    # oracle/libportm.so), like the pbrt-side vectors (make_golden_pbrt.py): sincosf in the vMF sampler is then the same
    # function on every machine.  `--probe-only` keeps the committed field / ISG files and re-answers the probes.
    shim = os.path.join(ROOT, "oracle", "libportm.so")
    if shim not in os.environ.get("LD_PRELOAD", ""):
        os.environ["LD_PRELOAD"] = shim
        os.execv(sys.executable, [sys.executable] + sys.argv)
    probe_only = "--probe-only" in sys.argv
    L = harness()
    path = os.path.join(HERE, "field_small.pgl")
    bmin, bmax = np.float32([-1, -1, -1]), np.float32([2, 2, 2])
    if not probe_only:
        it = L.ref_field_train(path.encode(), fp(bmin), fp(bmax), 6000, 14000, 3, 7, 1, 1500)
        assert it == 3, it
    h = L.ref_field_load(path.encode())
    assert L.ref_field_iteration(h) == 3
    rng = np.random.default_rng(99)
    n = 600
    pos = rng.uniform(bmin - 0.1, bmax + 0.1, size=(n, 3)).astype(np.float32)
    pos_s = pos.copy()
    pos_s[:, 1] = rng.uniform(-1.0, -0.9, size=n).astype(np.float32)      # surface samples were trained near the floor
    u = rng.random(n).astype(np.float32)
    u[:20] = -1.0                                                          # deterministic leaf lookup
    aux = unit(rng, n)
    g = rng.choice(np.float32([0.0, 0.3, 0.6, 0.85, -0.4, 0.97]), size=n).astype(np.float32)
    s2 = rng.random((n, 2)).astype(np.float32)
    wq = unit(rng, n)
    out = dict(pos=pos, pos_surface=pos_s, u=u, aux=aux, g=g, s2=s2, wq=wq)
    for mode, name, p in ((0, "surf_cos", pos_s), (1, "vol_hg", pos), (2, "surf_plain", pos_s), (3, "vol_plain", pos)):
        for k, v in probe(L, h, mode, p, u, aux, g, s2, wq).items():
            out[f"{name}_{k}"] = v
    L.ref_field_free(h)
    np.savez_compressed(os.path.join(HERE, "pgl_probe.npz"), **out)
    if probe_only:
        print("re-answered the probes on the committed field")
        return
    w, hgt = 40, 24
    vsp = np.zeros((hgt, w), np.float32)
    assert L.ref_isgb_make(os.path.join(HERE, "isg_small.exr").encode(), w, hgt, 8, 5, fp(vsp)) == 1
    np.save(os.path.join(HERE, "isg_small_vsp.npy"), vsp)
    print("wrote", path, os.path.getsize(path), "bytes;", {k: int((out[k] >= 0).sum()) for k in out if k.endswith("_id")})


if __name__ == "__main__":
    main()
