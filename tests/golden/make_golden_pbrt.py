"""Generates the committed golden vectors of the pbrt half of the path from the REFERENCE RENDERER ITSELF: the trimmed
CPU build of /root/reference that oracle/build_ref_pbrt.sh puts in oracle/_ref (libpbrt_ref.so + libpbrtharness.so).
Runs only where /root/reference exists; the fixtures travel, the reference does not.

  tests/golden/pbrt/<scene>.pbrt   small scenes in the reference's own scene language, one per BASELINE config family
                                   (C1 + triangle / sphere / emitter variants, C2 resampling VSPG with a loaded cache +
                                   ISG buffer, C3 NDS / NDS+ with a loaded Tr buffer, C4-like mis/ris with distant +
                                   infinite lights, C5-like rgbgrid + triangle meshes + BVH light sampler, homogeneous
                                   guided, gaussian pixel filter)
  tests/golden/pbrt/<scene>.npz    what the reference computes on that file:
        cam_*        ProjectiveCamera matrices (cameraFromRaster, renderFromCamera)
        samples      (px, py, sampleIndex) triples
        rays         RayIntegrator::EvaluatePixelSample's camera ray per sample: o, d, channelIdx, filterWeight
        L            GuidedVolPathVSPGIntegrator::Li for that sample (after the NaN/Inf guard)  -- the whole render pass
        li_*         Light::SampleLi + Light::PDF_Li on seeded contexts (medium points, axis-aligned and random normals)
        ls_*         the integrator's LightSampler::Sample / PMF on seeded contexts
        tm_*         SampleT_maj on seeded rays per medium: every callback (p, sigma_a, sigma_s, sigma_maj, T_maj), the
                     returned T_maj and the RNG's next draw
        ld_*         GuidedVolPathVSPGIntegrator::SampleLd at seeded surface / volume vertices
        hit_*        Integrator::Intersect on seeded rays: tHit, p(), n, shading.n
  tests/golden/isg_min.exr, tests/golden/tr_small.exr  inputs the scenes name (written with the product's EXR writer:
        the reference's own ISG store writes layer sets its own loader trips over -- tinyexr::ChannelsInLayer matches
        "Spp." inside "surfSpp." -- so the buffer here carries exactly the layers the loader needs: VSP.Y, surfContrib.RGB)

Run:  python tests/golden/make_golden_pbrt.py        (from the repo root, after `python -c "import __graft_entry__ as g; g.build()"`)
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
OUT = os.path.join(HERE, "pbrt")
PF, PI = C.POINTER(C.c_float), C.POINTER(C.c_int)

from ref_scenes import SCENES, build_scene, write_inputs  # noqa: E402  (tests/ref_scenes.py: the scene builders, shared with the tests)


def write_nvdb(golden_dir):
    """tests/golden/nvdb_small.nvdb through the reference's vendored NanoVDB (oracle/_ref/nvdb_writer, built by
    oracle/build_ref_pbrt.sh from oracle/ref_harness/nvdb_writer.cpp): a sparse blob of 20 x 18 x 16 seeded values at index
    offset (-5, 3, 100), voxel size 0.05, ZIP codec (12 KB; the same grid uncompressed is 660 KB of mostly empty nodes)."""
    import subprocess
    import tempfile
    rng = np.random.default_rng(77)
    nx, ny, nz = 20, 18, 16
    g = rng.random((nz, ny, nx), dtype=np.float32)
    z, y, x = np.mgrid[0:nz, 0:ny, 0:nx]
    mask = ((x - 10) ** 2 / 90 + (y - 9) ** 2 / 70 + (z - 8) ** 2 / 50) < 1.0
    g = np.where(mask, g * 2.0 + 0.05, 0).astype(np.float32)
    with tempfile.NamedTemporaryFile(suffix=".f32") as t:
        g.tofile(t.name)
        subprocess.run([os.path.join(ROOT, "oracle", "_ref", "nvdb_writer"), os.path.join(golden_dir, "nvdb_small.nvdb"), t.name, str(nx), str(ny), str(nz),
                        "-5", "3", "100", "0.05", "0.22", "-0.15", "-4.95", "zip"], check=True)


def harness():
    path = os.path.join(ROOT, "oracle", "_ref", "libpbrtharness.so")
    if not os.path.exists(path):
        raise SystemExit(f"{path} missing: run oracle/build_ref_pbrt.sh (needs /root/reference)")
    L = C.CDLL(path)
    L.refp_open.restype = C.c_void_p
    L.refp_open.argtypes = [C.c_char_p]
    L.refp_close.argtypes = [C.c_void_p]
    L.refp_camera_ray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, PF]
    L.refp_pixel_sample.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, PF]
    L.refp_camera_matrices.argtypes = [C.c_void_p, PF]
    L.refp_sample_li.argtypes = [C.c_void_p, C.c_int, PF, PI, PF, PF]
    L.refp_light_sample.argtypes = [C.c_void_p, C.c_int, PF, PF, PI, PI, PF, PF]
    L.refp_sample_tmaj.argtypes = [C.c_void_p, C.c_char_p, PF, C.c_float, C.c_float, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, PF, PF]
    L.refp_sample_ld.argtypes = [C.c_void_p, C.c_int, PF, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, PF, C.c_int, PF]
    L.refp_intersect.argtypes = [C.c_void_p, C.c_int, PF, PI, PF]
    L.refp_init(1, 0)   # one thread, rendering space "cameraworld" (the reference's default)
    return L


def fp(a):
    return a.ctypes.data_as(PF)


def ip_(a):
    return a.ctypes.data_as(PI)


def probes(spec, rng, n_lights, n_media):
    """Seeded probe inputs shared by the generator and the tests (the tests re-derive nothing: everything is stored)."""
    eye = np.float32(spec["eye"])
    n = 600
    p = (rng.random((n, 3)).astype(np.float32) * 1.6 - 0.3) - eye
    nn = rng.normal(size=(n, 3)).astype(np.float32)
    nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    ax = np.eye(3, dtype=np.float32)[rng.integers(3, size=n)] * np.where(rng.random(n) < 0.5, -1, 1)[:, None].astype(np.float32)
    nn[1::3] = ax[1::3]
    nn[::3] = 0
    ctx = np.concatenate([p, nn, nn], axis=1).astype(np.float32)
    out = dict(ctx=ctx, u2=rng.random((n, 2)).astype(np.float32), u1=rng.random(n).astype(np.float32),
               light=rng.integers(max(n_lights, 1), size=n).astype(np.int32))
    # rays through the medium box (world [0,1]^3)
    m = 300
    o = (rng.random((m, 3)) * 1.6 - 0.3 - eye).astype(np.float32)
    tgt = (rng.random((m, 3)) - eye).astype(np.float32)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    d[::7] *= np.float32(0.37)   # unnormalised directions (shadow rays are)
    out.update(ray=np.concatenate([o, d], axis=1).astype(np.float32), tmax=(rng.random(m) * 2.5 + 0.1).astype(np.float32),
               ru=rng.random(m).astype(np.float32), ch=rng.integers(3, size=m).astype(np.int32))
    # interior points / directions for SampleLd at volume vertices, incoming rays for surface vertices
    k = 300
    pv = (rng.random((k, 3)) * 0.9 + 0.05 - eye).astype(np.float32)
    dv = rng.normal(size=(k, 3))
    dv = (dv / np.linalg.norm(dv, axis=1, keepdims=True)).astype(np.float32)
    ps = (rng.random((k, 3)) * 2.0 - 0.5 - eye).astype(np.float32)
    ps[:, 1] = np.abs(ps[:, 1] + eye[1]) - eye[1] + np.float32(0.05)   # above the ground plane
    out.update(ld_vol=np.concatenate([pv, dv], axis=1).astype(np.float32), ld_surf=np.concatenate([ps, dv], axis=1).astype(np.float32))
    return out


def main():
    # The reference runs with the path's portable libm (include/vspg_libm.h, oracle/libportm.so) instead of whichever
    # variant the host's libm selects for this CPU: the golden vectors are then the same on every machine, and they are
    # what the CUDA kernels -- which compute with the same functions -- must reproduce bit for bit.
    # One process per scene: the reference keeps process-wide state (parser, transform cache, guiding device).
    shim = os.path.join(ROOT, "oracle", "libportm.so")
    if not os.path.exists(shim):
        raise SystemExit(f"{shim} missing: run `make -C oracle libportm.so`")
    if len(sys.argv) < 2:
        import subprocess
        os.makedirs(OUT, exist_ok=True)
        write_inputs(HERE)
        write_nvdb(HERE)
        env = dict(os.environ, LD_PRELOAD=shim)
        for name in SCENES:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), name], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            tail = [l for l in r.stdout.splitlines() if l.startswith(name + ":")]
            print(tail[-1] if tail else f"{name}: FAILED rc={r.returncode}\n" + r.stdout[-2000:])
        return
    assert C.CDLL(None).vspg_portm_loaded() == 1
    L = harness()
    os.chdir(OUT)   # the reference resolves the cache / buffer file names against the working directory
    total = 0
    for name in sys.argv[1:]:
        spec = build_scene(name, OUT)
        h = L.refp_open(spec["path"].encode())
        assert h, name
        W, H, spp = spec["width"], spec["height"], spec["spp"]
        rng = np.random.default_rng(sum(map(ord, name)) * 7919)
        out = {}
        mats = np.zeros(66, np.float32)
        assert L.refp_camera_matrices(h, fp(mats)) == 0
        out["cam_camera_from_raster"], out["cam_render_from_camera"] = mats[:16].reshape(4, 4), mats[32:48].reshape(4, 4)
        n = spec.get("n_samples", 1500)
        smp = np.stack([rng.integers(W, size=n), rng.integers(H, size=n), rng.integers(spp, size=n)], axis=1).astype(np.int32)
        rays, Ls = np.zeros((n, 8), np.float32), np.zeros((n, 3), np.float32)
        o14, o8 = np.zeros(14, np.float32), np.zeros(8, np.float32)
        for i, (px, py, s) in enumerate(smp):
            assert L.refp_camera_ray(h, int(px), int(py), int(s), fp(o14)) == 0
            rays[i] = np.concatenate([o14[:6], o14[11:12], o14[10:11]])
            assert L.refp_pixel_sample(h, int(px), int(py), int(s), fp(o8)) == 0
            Ls[i] = o8[:3]
        out.update(samples=smp, rays=rays, L=Ls)
        pr = probes(spec, rng, spec["n_lights"], len(spec["media"]))
        m = len(pr["ctx"])
        if spec["n_lights"]:
            li = np.zeros((m, 12), np.float32)
            assert L.refp_sample_li(h, m, fp(pr["ctx"]), ip_(pr["light"]), fp(pr["u2"]), fp(li)) == 0
            pick, pmf, pmf_of = np.zeros(m, np.int32), np.zeros(m, np.float32), np.zeros(m, np.float32)
            assert L.refp_light_sample(h, m, fp(pr["ctx"]), fp(pr["u1"]), ip_(pr["light"]), ip_(pick), fp(pmf), fp(pmf_of)) == 0
            out.update(li_ctx=pr["ctx"], li_light=pr["light"], li_u2=pr["u2"], li_out=li, ls_u=pr["u1"], ls_pick=pick, ls_pmf=pmf, ls_pmf_of=pmf_of)
        for mi, mname in enumerate(spec["media"]):
            k = len(pr["ray"])
            rec, fin, cnt = np.zeros((k, 48, 15), np.float32), np.zeros((k, 4), np.float32), np.zeros(k, np.int32)
            for i in range(k):
                cnt[i] = L.refp_sample_tmaj(h, mname.encode(), fp(pr["ray"][i]), float(pr["tmax"][i]), float(pr["ru"][i]), 1000 + i, 17, int(pr["ch"][i]),
                                            -1, 48, fp(rec[i]), fp(fin[i]))
                assert cnt[i] >= 0
            out.update({f"tm{mi}_ray": pr["ray"], f"tm{mi}_tmax": pr["tmax"], f"tm{mi}_u": pr["ru"], f"tm{mi}_ch": pr["ch"], f"tm{mi}_n": cnt,
                        f"tm{mi}_rec": rec[:, :int(min(48, cnt.max()))], f"tm{mi}_final": fin})
        # SampleLd at seeded vertices
        rp = np.ones(3, np.float32)
        for kind, rays6, mname in (("vol", pr["ld_vol"], spec["media"][0] if spec["media"] else None), ("surf", pr["ld_surf"], "")):
            if mname is None or not spec["n_lights"]:
                continue
            k = len(rays6)
            ld, rc = np.zeros((k, 3), np.float32), np.zeros(k, np.int32)
            for i in range(k):
                rc[i] = L.refp_sample_ld(h, 1 if kind == "surf" else 0, fp(rays6[i]), mname.encode(), 3, 4, i, 9, fp(rp), int(pr["ch"][i % len(pr["ch"])]), fp(ld[i]))
            out.update({f"ld_{kind}_in": rays6, f"ld_{kind}_rc": rc, f"ld_{kind}_out": ld, f"ld_{kind}_ch": pr["ch"][np.arange(k) % len(pr["ch"])]})
        # closest hits
        hr = np.concatenate([pr["ld_surf"], np.full((len(pr["ld_surf"]), 1), np.inf, np.float32)], axis=1).astype(np.float32)
        hit, h10 = np.zeros(len(hr), np.int32), np.zeros((len(hr), 10), np.float32)
        assert L.refp_intersect(h, len(hr), fp(hr), ip_(hit), fp(h10)) == 0
        out.update(hit_rays=hr, hit_valid=hit, hit_out=h10)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        sz = os.path.getsize(os.path.join(OUT, name + ".npz"))
        total += sz
        print(f"{name}: {n} samples, mean L {Ls.mean():.5f}, {int((Ls != 0).any(axis=1).sum())} non-black, npz {sz // 1024} KiB")
        L.refp_close(h)
    print("total", total // 1024, "KiB")
    sys.stdout.flush()
    os._exit(0)   # the reference's static destructors are not exercised


if __name__ == "__main__":
    main()
