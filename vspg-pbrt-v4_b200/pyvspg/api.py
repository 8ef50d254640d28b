"""Python harness over the C ABI of libvspg.so (include/vspg.h): the call a user of the reference's
`Integrator::Create(...)->Render()` seam makes (cpu/integrators.cpp:3711-3768, cpu/render.cpp:163), as a context
object.  Host arrays are numpy; device memory is owned by the context.  No CPU fallback exists.
"""
import ctypes as C

import numpy as np

from . import _abi as A

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = A.load_library()
    return _lib


class VspgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vspg error {A.ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


_DEFAULTS = dict(max_depth=5, min_rr_depth=1, use_nee=1, surface_guiding=1, volume_guiding=1, surface_guiding_ris=1, volume_guiding_ris=0,
                 vsp_guiding=1, vsp_primary_guiding=1, vsp_secondary_guiding=1, vsp_mis_ratio=0.5, vsp_criterion_variance=1,
                 vsp_sampling_nds=0, collision_probability_bias=0, light_sampler=A.LIGHTSAMPLER_BVH, regularize=0)


def integrator_defaults(**over):
    """guidedvolpathvspgintegrator.cpp:1260-1322 defaults, overridable by keyword (field names of vspg_integrator_params)."""
    ip = A.IntegratorParams()
    # the same values vspg_integrator_defaults() fills in (tests/test_cpu.py checks the two against each other); kept in
    # Python so that building a configuration -- e.g. to write its scene file for the reference arm of bench.py -- does
    # not need the CUDA library
    for k, v in _DEFAULTS.items():
        setattr(ip, k, v)
    for k, v in over.items():
        if not hasattr(ip, k):
            raise KeyError(f"unknown integrator parameter {k!r}")  # the reference ErrorExits on unused parameters
        setattr(ip, k, v)
    return ip


def _fp(a):
    return a.ctypes.data_as(A.PF)


class Context:
    """One vspg_ctx (= one GPU).  Thread-compatible, not thread-safe."""

    def __init__(self, device=0):
        self._h = A.P()
        rc = lib().vspg_create(C.byref(self._h), int(device))
        if rc != 0:
            raise VspgError(rc, lib().vspg_last_error(None).decode())
        self.width = self.height = 0

    def close(self):
        if self._h:
            lib().vspg_destroy(self._h)
            self._h = A.P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise VspgError(rc, lib().vspg_last_error(self._h).decode())

    # ---- inputs
    def set_scene(self, scene):
        """scene: pyvspg.scenes.Scene (holds the vspg_scene_desc and keeps its arrays alive)."""
        self._ck(lib().vspg_set_scene(self._h, C.byref(scene.desc)))
        self.width, self.height = scene.desc.camera.width, scene.desc.camera.height

    def set_guiding(self, guiding):
        self._ck(lib().vspg_set_guiding(self._h, C.byref(guiding.desc) if guiding is not None else None))

    def set_integrator(self, ip):
        self._ck(lib().vspg_set_integrator(self._h, C.byref(ip)))

    def set_capacity(self, n):
        self._ck(lib().vspg_set_capacity(self._h, int(n)))

    # ---- rendering
    def render(self, spp_begin, spp_end, spp_stride=1, window=None, seed=0, stream=None):
        x0, y0, x1, y1 = window if window is not None else (0, 0, self.width, self.height)
        self._ck(lib().vspg_render(self._h, spp_begin, spp_end, spp_stride, x0, y0, x1, y1, int(seed), stream))

    def film_clear(self):
        self._ck(lib().vspg_film_clear(self._h))

    def film_device_ptr(self):
        p, n = A.P(), C.c_size_t()
        self._ck(lib().vspg_film_device_ptr(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def film_download(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.float64)
        self._ck(lib().vspg_film_download(self._h, out.ctypes.data_as(A.P)))
        return out

    def film_download_to(self, host_ptr):
        """Download into a caller-owned (e.g. pinned) host buffer given by address."""
        self._ck(lib().vspg_film_download(self._h, A.P(host_ptr)))

    def film_resolve(self):
        out = np.empty((self.height, self.width, 3), dtype=np.float32)
        self._ck(lib().vspg_film_resolve(self._h, out.ctypes.data_as(A.P)))
        return out

    def counters(self):
        c = A.Counters()
        self._ck(lib().vspg_counters_get(self._h, C.byref(c)))
        return {k: int(getattr(c, k)) for k, _ in A.Counters._fields_}

    def counters_reset(self):
        self._ck(lib().vspg_counters_reset(self._h))

    def kernel_stats(self, enable_timing=True):
        """Per-kernel device ms / launches / grid-traffic counters since the last counters_reset()."""
        st = A.KernelStats()
        self._ck(lib().vspg_kernel_stats_get(self._h, int(enable_timing), C.byref(st)))
        out = {"ms_total": st.ms_total}
        for i, k in enumerate(A.KERNEL_NAMES):
            out[k] = {"ms": st.ms[i], "launches": int(st.launches[i]), "density_queries": int(st.density_queries[i]),
                      "dda_steps": int(st.dda_steps[i])}
        return out

    # ---- per-kernel parity hooks
    def debug_majorant_segments(self, medium, rays7, max_segments=64):
        rays7 = np.ascontiguousarray(rays7, dtype=np.float32).reshape(-1, 7)
        n = rays7.shape[0]
        seg = (A.Segment * (n * max_segments))()
        cnt = np.zeros(n, dtype=np.int32)
        self._ck(lib().vspg_debug_majorant_segments(self._h, medium, n, _fp(rays7), max_segments, seg, cnt.ctypes.data_as(C.POINTER(A.i32))))
        arr = np.frombuffer(seg, dtype=np.float32).reshape(n, max_segments, 5).copy()
        return arr, cnt

    def debug_majorant_segments_noalloc(self, medium, rays7, max_segments=64):
        """vspg_debug_majorant_segments_noalloc: same results, staged in the library's fixed device scratch (no cudaMalloc)."""
        rays7 = np.ascontiguousarray(rays7, dtype=np.float32).reshape(-1, 7)
        n = rays7.shape[0]
        seg = (A.Segment * max(1, n * max_segments))()
        cnt = np.zeros(n, dtype=np.int32)
        self._ck(lib().vspg_debug_majorant_segments_noalloc(self._h, medium, n, _fp(rays7), max_segments, seg, cnt.ctypes.data_as(C.POINTER(A.i32))))
        arr = np.frombuffer(seg, dtype=np.float32)[:n * max_segments * 5].reshape(n, max_segments, 5).copy()
        return arr, cnt

    def debug_density(self, medium, pos3):
        pos3 = np.ascontiguousarray(pos3, dtype=np.float32).reshape(-1, 3)
        out = np.zeros((pos3.shape[0], 7), dtype=np.float32)
        self._ck(lib().vspg_debug_density(self._h, medium, pos3.shape[0], _fp(pos3), _fp(out)))
        return out

    def debug_sample_distance(self, inputs):
        n = len(inputs)
        out = (A.SdOutput * n)()
        self._ck(lib().vspg_debug_sample_distance(self._h, n, inputs, out))
        return out

    def debug_cache_lookup(self, is_surface, pos3, u):
        pos3 = np.ascontiguousarray(pos3, dtype=np.float32).reshape(-1, 3)
        u = np.ascontiguousarray(u, dtype=np.float32)
        n = pos3.shape[0]
        leaf, region, uo = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
        self._ck(lib().vspg_debug_cache_lookup(self._h, int(is_surface), n, _fp(pos3), _fp(u), leaf.ctypes.data_as(C.POINTER(A.i32)),
                                               region.ctypes.data_as(C.POINTER(A.i32)), _fp(uo)))
        return leaf, region, uo

    def debug_light_sample(self, pos3, ns3, u, light_of):
        pos3 = np.ascontiguousarray(pos3, dtype=np.float32).reshape(-1, 3)
        ns3 = np.ascontiguousarray(ns3, dtype=np.float32).reshape(-1, 3)
        u = np.ascontiguousarray(u, dtype=np.float32)
        light_of = np.ascontiguousarray(light_of, dtype=np.int32)
        n = pos3.shape[0]
        light, pmf, pmf_of = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros(n, np.float32)
        ip = C.POINTER(A.i32)
        self._ck(lib().vspg_debug_light_sample(self._h, n, _fp(pos3), _fp(ns3), _fp(u), light_of.ctypes.data_as(ip), light.ctypes.data_as(ip),
                                               _fp(pmf), _fp(pmf_of)))
        return light, pmf, pmf_of

    def tr_store_begin(self):
        self._ck(lib().vspg_tr_store_begin(self._h))

    def tr_store_download(self):
        """(mean transmittance [h, w, 3] float32, sample counts [h, w] int32) of the storeTrBuffer pre-pass."""
        rgb = np.zeros((self.height, self.width, 3), np.float32)
        cnt = np.zeros((self.height, self.width), np.int32)
        self._ck(lib().vspg_tr_store_download(self._h, _fp(rgb), cnt.ctypes.data_as(C.POINTER(A.i32))))
        return rgb, cnt

    def tr_store_end(self):
        self._ck(lib().vspg_tr_store_end(self._h))

    def isg_store_begin(self):
        self._ck(lib().vspg_isg_store_begin(self._h))

    def isg_store_update(self):
        self._ck(lib().vspg_isg_store_update(self._h))

    def isg_store_download(self, width, height):
        """-> vsp [h, w] float32, records [h, w, 8] float32 (surfSpp, volSpp, surfContrib rgb, volContrib rgb)."""
        vsp = np.zeros((height, width), dtype=np.float32)
        rec = np.zeros((height, width, 8), dtype=np.float32)
        self._ck(lib().vspg_isg_store_download(self._h, _fp(vsp), _fp(rec)))
        return vsp, rec

    def isg_store_end(self):
        self._ck(lib().vspg_isg_store_end(self._h))

    def render_training_isg(self, spp, seed=0, stream=None):
        """The reference's wave loop with the ISG buffer trained on the fly (cpu/integrators.cpp:123-239 with nextWaveSize = 1,
        vspg.cpp:230-258): waves of one sample index, ImageSpaceGuidingBuffer::Update after waves 1, 2, 4, 8, ..."""
        self.isg_store_begin()
        next_update = 1
        for wave in range(1, spp + 1):
            self.render(wave - 1, wave, seed=seed, stream=stream)
            if wave == next_update:
                self.isg_store_update()
                next_update *= 2

    def debug_intersect(self, rays7):
        rays7 = np.ascontiguousarray(rays7, dtype=np.float32).reshape(-1, 7)
        n = rays7.shape[0]
        prim, out = np.zeros(n, np.int32), np.zeros((n, 10), np.float32)
        self._ck(lib().vspg_debug_intersect(self._h, n, _fp(rays7), prim.ctypes.data_as(C.POINTER(A.i32)), _fp(out)))
        return prim, out

    def debug_sample_li(self, ctx9, light, u2):
        ctx9 = np.ascontiguousarray(ctx9, dtype=np.float32).reshape(-1, 9)
        light = np.ascontiguousarray(light, dtype=np.int32)
        u2 = np.ascontiguousarray(u2, dtype=np.float32).reshape(-1, 2)
        n = ctx9.shape[0]
        out = np.zeros((n, 12), np.float32)
        self._ck(lib().vspg_debug_sample_li(self._h, n, _fp(ctx9), light.ctypes.data_as(C.POINTER(A.i32)), _fp(u2), _fp(out)))
        return out

    def debug_sample_ld(self, inputs):
        """SampleLd at explicit vertices (A.LdInput array) -> (status[n], Ld[n,3])."""
        n = len(inputs)
        out = (A.LdOutput * n)()
        self._ck(lib().vspg_debug_sample_ld(self._h, n, inputs, out))
        return (np.array([o.status for o in out], dtype=np.int32),
                np.array([[o.Ld[0], o.Ld[1], o.Ld[2]] for o in out], dtype=np.float32).reshape(n, 3))

    def debug_libm(self, fn, a, b=None):
        """vspg_debug_libm: the device's value of libm function `fn` at a (and b)."""
        a = np.ascontiguousarray(a, dtype=np.float32)
        out = np.zeros_like(a)
        bb = None if b is None else np.ascontiguousarray(b, dtype=np.float32)
        self._ck(lib().vspg_debug_libm(self._h, int(fn), a.size, _fp(a), None if bb is None else _fp(bb), _fp(out)))
        return out

    def debug_guided_sample(self, is_surface, inputs):
        n = len(inputs)
        out = (A.GsOutput * n)()
        self._ck(lib().vspg_debug_guided_sample(self._h, int(is_surface), n, inputs, out))
        return out


class HostScene:
    """A .pbrt scene parsed by the C++ front end (vspg_host_scene_parse): exposes .desc / .guiding / .ip with the
    same attribute names Context.set_scene / set_guiding / set_integrator consume, plus spp / seed / outfile.
    Raises VspgError with the front end's file:line message (the reference ErrorExits there)."""

    def __init__(self, path=None, text=None, base_dir="."):
        self._h = A.P()
        err = C.create_string_buffer(1024)
        if path is not None:
            rc = lib().vspg_host_scene_parse(str(path).encode(), C.byref(self._h), err, len(err))
        else:
            rc = lib().vspg_host_scene_parse_string(text.encode(), str(base_dir).encode(), C.byref(self._h), err, len(err))
        if rc != 0:
            self._h = A.P()
            raise VspgError(rc, err.value.decode())
        self.desc = lib().vspg_host_scene_desc(self._h).contents
        g = lib().vspg_host_scene_guiding(self._h)
        self.guiding = _Holder(g.contents) if g else None
        self.ip = lib().vspg_host_scene_integrator(self._h).contents
        self.spp = lib().vspg_host_scene_spp(self._h)
        self.seed = lib().vspg_host_scene_seed(self._h)
        self.outfile = lib().vspg_host_scene_outfile(self._h).decode()
        self.store_tr_file = lib().vspg_host_scene_store_tr_file(self._h).decode()   # "" unless "storeTrBuffer" true

    def close(self):
        if self._h:
            self.desc = self.guiding = self.ip = None
            lib().vspg_host_scene_free(self._h)
            self._h = A.P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _Holder:
    def __init__(self, desc):
        self.desc = desc


def exr_read_channel(path, channel):
    """vspg_exr_read_channel -> float32 [h, w]."""
    out, w, h = A.PF(), A.i32(), A.i32()
    err = C.create_string_buffer(512)
    rc = lib().vspg_exr_read_channel(str(path).encode(), channel.encode(), C.byref(out), C.byref(w), C.byref(h), err, len(err))
    if rc != 0:
        raise VspgError(rc, err.value.decode())
    a = np.ctypeslib.as_array(out, shape=(h.value, w.value)).copy()
    lib().vspg_free(out)
    return a


def exr_write_rgb(path, rgb, as_half=False):
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    rc = lib().vspg_exr_write_rgb(str(path).encode(), _fp(rgb), rgb.shape[1], rgb.shape[0], int(as_half))
    if rc != 0:
        raise VspgError(rc, f"cannot write {path}")


def exr_write_channels(path, channels):
    """channels: {name: [h, w] float array}; float ZIP EXR (the layout TrBuffer::Store writes)."""
    names = list(channels)
    planes = [np.ascontiguousarray(channels[n], dtype=np.float32) for n in names]
    h, w = planes[0].shape
    cn = (C.c_char_p * len(names))(*[n.encode() for n in names])
    cp = (A.PF * len(names))(*[_fp(p) for p in planes])
    rc = lib().vspg_exr_write_channels(str(path).encode(), w, h, len(names), cn, cp)
    if rc != 0:
        raise VspgError(rc, f"cannot write {path}")


def build_majorant_grid(medium):
    n = medium.maj_res[0] * medium.maj_res[1] * medium.maj_res[2]
    out = np.zeros(n, dtype=np.float32)
    rc = lib().vspg_build_majorant_grid(C.byref(medium), _fp(out))
    if rc != 0:
        raise VspgError(rc, "build_majorant_grid")
    return out
