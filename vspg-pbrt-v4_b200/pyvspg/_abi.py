"""ctypes mirror of include/vspg.h (the C ABI of libvspg.so).  Plumbing only: no arithmetic lives here.

The library is loaded from the package directory (built in-tree by csrc/Makefile).  There is no fallback: if the
CUDA library is missing, importing `pyvspg.api` raises.
"""
import ctypes as C
import os

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("VSPG_LIB") or os.path.join(PKG_DIR, "libvspg.so")   # VSPG_LIB: A/B builds of the same ABI (tools/ab_run.py)

VSPG_OK = 0
ERR_INVALID, ERR_CUDA, ERR_IO, ERR_UNSUPPORTED, ERR_STATE = -1, -2, -3, -4, -5
ERR_NAMES = {0: "OK", -1: "INVALID", -2: "CUDA", -3: "IO", -4: "UNSUPPORTED", -5: "STATE"}

NOALLOC_MAX_RAYS, NOALLOC_MAX_SEGMENTS = 16384, 262144
MEDIUM_HOMOGENEOUS, MEDIUM_GRID, MEDIUM_RGBGRID, MEDIUM_NANOVDB = 0, 1, 2, 3
LIGHT_DIFFUSE_AREA, LIGHT_POINT, LIGHT_DISTANT, LIGHT_UNIFORM_INFINITE, LIGHT_SPOT = 0, 1, 2, 3, 4
LIGHTSAMPLER_UNIFORM, LIGHTSAMPLER_POWER, LIGHTSAMPLER_BVH = 0, 1, 2
MAX_LOBES = 32

f32, i32, u32, u64 = C.c_float, C.c_int, C.c_uint32, C.c_uint64
F3 = f32 * 3
PF = C.POINTER(f32)


class Transform(C.Structure):
    _fields_ = [("m", (f32 * 4) * 4), ("mInv", (f32 * 4) * 4)]


class Camera(C.Structure):
    _fields_ = [("camera_from_raster", Transform), ("render_from_camera", Transform), ("lens_radius", f32),
                ("focal_distance", f32), ("medium", i32), ("width", i32), ("height", i32), ("filter_radius", f32 * 2),
                ("filter_type", i32), ("filter_sigma", f32), ("imaging_ratio", f32), ("max_component_value", f32)]


class Medium(C.Structure):
    _fields_ = [("type", i32), ("bounds_min", F3), ("bounds_max", F3), ("render_from_medium", Transform),
                ("sigma_a", F3), ("sigma_s", F3), ("sigma_scale", f32), ("g", f32), ("nx", i32), ("ny", i32), ("nz", i32),
                ("density", PF), ("sigma_a_rgb", PF), ("sigma_s_rgb", PF), ("maj_res", i32 * 3), ("majorant", PF),
                ("index_min", i32 * 3), ("index_from_medium", f32 * 9), ("index_origin", F3), ("density_offset", f32), ("background", f32),
                ("majorant_scale", f32), ("Le", F3), ("le_scale", PF), ("le_nx", i32), ("le_ny", i32), ("le_nz", i32)]


class Quad(C.Structure):
    _fields_ = [("p00", F3), ("p10", F3), ("p01", F3), ("p11", F3), ("material", i32), ("area_light", i32),
                ("medium_inside", i32), ("medium_outside", i32), ("has_interface", i32), ("flip_normal", i32)]


class Triangle(C.Structure):
    _fields_ = [("p0", F3), ("p1", F3), ("p2", F3), ("material", i32), ("area_light", i32), ("medium_inside", i32),
                ("medium_outside", i32), ("has_interface", i32), ("flip_normal", i32)]


class Sphere(C.Structure):
    _fields_ = [("center", F3), ("radius", f32), ("material", i32), ("area_light", i32), ("medium_inside", i32),
                ("medium_outside", i32), ("has_interface", i32), ("flip_normal", i32)]


class Material(C.Structure):
    _fields_ = [("type", i32), ("reflectance", F3)]


class Light(C.Structure):
    _fields_ = [("type", i32), ("L", F3), ("scale", C.c_float), ("pos", F3), ("shape_kind", i32), ("shape_index", i32), ("two_sided", i32),
                ("medium", i32), ("light_from_render", C.c_float * 9), ("axis", F3), ("cos_falloff_start", C.c_float),
                ("cos_falloff_end", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("camera", Camera), ("n_media", i32), ("media", C.POINTER(Medium)), ("n_quads", i32),
                ("quads", C.POINTER(Quad)), ("n_triangles", i32), ("triangles", C.POINTER(Triangle)), ("n_spheres", i32),
                ("spheres", C.POINTER(Sphere)), ("n_materials", i32), ("materials", C.POINTER(Material)),
                ("n_lights", i32), ("lights", C.POINTER(Light))]


class KdNode(C.Structure):
    _fields_ = [("split", f32), ("dim_idx", u32)]


L32 = f32 * MAX_LOBES


class Mixture(C.Structure):
    _fields_ = [("weight", L32), ("kappa", L32), ("mu_x", L32), ("mu_y", L32), ("mu_z", L32), ("norm", L32),
                ("e_minus_2k", L32), ("mean_cos", L32), ("distance", L32), ("vsp_weight", L32), ("pivot", F3),
                ("num_components", i32), ("valid", i32), ("pad_", i32 * 3)]


class Neighbours(C.Structure):
    _fields_ = [("ids", u32 * 8), ("px", f32 * 8), ("py", f32 * 8), ("pz", f32 * 8), ("size", u32), ("pad_", u32 * 7)]


class Field(C.Structure):
    _fields_ = [("usable", i32), ("bounds_min", F3), ("bounds_max", F3), ("n_nodes", i32), ("nodes", C.POINTER(KdNode)),
                ("n_regions", i32), ("regions", C.POINTER(Mixture)), ("neighbours", C.POINTER(Neighbours)),
                ("stochastic_nn", i32), ("is_nn", i32)]


class GuidingDesc(C.Structure):
    _fields_ = [("surface", C.POINTER(Field)), ("volume", C.POINTER(Field)), ("isg_vsp", PF), ("isg_width", i32),
                ("isg_height", i32), ("tr_rgb", PF), ("tr_width", i32), ("tr_height", i32)]


class IntegratorParams(C.Structure):
    _fields_ = [("max_depth", i32), ("min_rr_depth", i32), ("use_nee", i32), ("surface_guiding", i32),
                ("volume_guiding", i32), ("surface_guiding_ris", i32), ("volume_guiding_ris", i32), ("vsp_guiding", i32),
                ("vsp_primary_guiding", i32), ("vsp_secondary_guiding", i32), ("vsp_mis_ratio", f32),
                ("vsp_criterion_variance", i32), ("vsp_sampling_nds", i32), ("collision_probability_bias", i32),
                ("light_sampler", i32), ("regularize", i32)]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_ulonglong), ("density_queries", C.c_ulonglong), ("dda_steps", C.c_ulonglong),
                ("cache_lookups", C.c_ulonglong), ("bvh_nodes", C.c_ulonglong), ("shadow_rays", C.c_ulonglong),
                ("kernel_launches", C.c_ulonglong), ("wave_iterations", C.c_ulonglong)]


class KernelStats(C.Structure):
    _fields_ = [("ms", f32 * 8), ("launches", C.c_ulonglong * 8), ("density_queries", C.c_ulonglong * 8),
                ("dda_steps", C.c_ulonglong * 8), ("ms_total", f32)]


KERNEL_NAMES = ("generate", "intersect", "medium", "volscatter", "surface", "shade", "shadow", "finish")


class Segment(C.Structure):
    _fields_ = [("t_min", f32), ("t_max", f32), ("sigma_maj", F3)]


class SdInput(C.Structure):
    _fields_ = [("o", F3), ("d", F3), ("t_max", f32), ("medium", i32), ("pixel_x", i32), ("pixel_y", i32), ("depth", i32),
                ("channel", i32), ("vsp", f32), ("beta", F3), ("r_u", F3), ("r_l", F3), ("rng_state", u64), ("rng_inc", u64)]


class SdOutput(C.Structure):
    _fields_ = [("scattered", i32), ("terminated", i32), ("depth", i32), ("p", F3), ("sigma_a", F3), ("sigma_s", F3),
                ("beta", F3), ("r_u", F3), ("r_l", F3), ("rng_state", u64), ("density_queries", i32), ("dda_steps", i32)]


class LdInput(C.Structure):
    _fields_ = [("is_surface", C.c_int), ("ray_o", C.c_float * 3), ("ray_d", C.c_float * 3), ("medium", C.c_int), ("px", C.c_int), ("py", C.c_int),
                ("sample_index", C.c_int), ("dim", C.c_int), ("seed", C.c_int), ("r_p", C.c_float * 3), ("channel", C.c_int)]


class LdOutput(C.Structure):
    _fields_ = [("status", C.c_int), ("Ld", C.c_float * 3)]


class GsInput(C.Structure):
    _fields_ = [("p", F3), ("wo", F3), ("n", F3), ("g", f32), ("reflectance", F3), ("v", f32), ("u1", f32), ("u2", f32 * 2),
                ("wq", F3), ("rng_state", u64), ("rng_inc", u64)]


class GsOutput(C.Structure):
    _fields_ = [("ok", i32), ("region", i32), ("valid", i32), ("wi", F3), ("f", F3), ("pdf", f32), ("scatter_pdf", f32),
                ("mis_pdf", f32), ("pdf_wq", f32), ("vsp_wq", f32), ("rng_state", u64)]


# every symbol include/vspg.h declares: name -> (restype, argtypes)
P = C.c_void_p
SYMBOLS = {
    "vspg_create": (i32, [C.POINTER(P), i32]),
    "vspg_destroy": (None, [P]),
    "vspg_last_error": (C.c_char_p, [P]),
    "vspg_version": (C.c_char_p, []),
    "vspg_integrator_defaults": (None, [C.POINTER(IntegratorParams)]),
    "vspg_set_scene": (i32, [P, C.POINTER(SceneDesc)]),
    "vspg_set_guiding": (i32, [P, C.POINTER(GuidingDesc)]),
    "vspg_set_integrator": (i32, [P, C.POINTER(IntegratorParams)]),
    "vspg_set_capacity": (i32, [P, i32]),
    "vspg_render": (i32, [P, i32, i32, i32, i32, i32, i32, i32, i32, P]),
    "vspg_film_clear": (i32, [P]),
    "vspg_film_device_ptr": (i32, [P, C.POINTER(P), C.POINTER(C.c_size_t)]),
    "vspg_film_download": (i32, [P, P]),
    "vspg_film_resolve": (i32, [P, P]),
    "vspg_counters_get": (i32, [P, C.POINTER(Counters)]),
    "vspg_counters_reset": (i32, [P]),
    "vspg_kernel_stats_get": (i32, [P, i32, C.POINTER(KernelStats)]),
    "vspg_host_scene_parse": (i32, [C.c_char_p, C.POINTER(P), C.c_char_p, C.c_size_t]),
    "vspg_host_scene_parse_string": (i32, [C.c_char_p, C.c_char_p, C.POINTER(P), C.c_char_p, C.c_size_t]),
    "vspg_host_scene_free": (None, [P]),
    "vspg_host_scene_desc": (C.POINTER(SceneDesc), [P]),
    "vspg_host_scene_guiding": (C.POINTER(GuidingDesc), [P]),
    "vspg_host_scene_integrator": (C.POINTER(IntegratorParams), [P]),
    "vspg_host_scene_spp": (i32, [P]),
    "vspg_host_scene_seed": (i32, [P]),
    "vspg_host_scene_outfile": (C.c_char_p, [P]),
    "vspg_host_field_load": (i32, [C.c_char_p, C.POINTER(P), C.c_char_p, C.c_size_t]),
    "vspg_host_field_free": (None, [P]),
    "vspg_host_field_surface": (C.POINTER(Field), [P]),
    "vspg_host_field_volume": (C.POINTER(Field), [P]),
    "vspg_exr_read_channel": (i32, [C.c_char_p, C.c_char_p, C.POINTER(PF), C.POINTER(i32), C.POINTER(i32), C.c_char_p, C.c_size_t]),
    "vspg_exr_write_rgb": (i32, [C.c_char_p, PF, i32, i32, i32]),
    "vspg_free": (None, [P]),
    "vspg_build_majorant_grid": (i32, [C.POINTER(Medium), PF]),
    "vspg_debug_majorant_segments": (i32, [P, i32, i32, PF, i32, C.POINTER(Segment), C.POINTER(i32)]),
    "vspg_debug_majorant_segments_noalloc": (i32, [P, i32, i32, PF, i32, C.POINTER(Segment), C.POINTER(i32)]),
    "vspg_debug_density": (i32, [P, i32, i32, PF, PF]),
    "vspg_debug_sample_distance": (i32, [P, i32, C.POINTER(SdInput), C.POINTER(SdOutput)]),
    "vspg_debug_cache_lookup": (i32, [P, i32, i32, PF, PF, C.POINTER(i32), C.POINTER(i32), PF]),
    "vspg_host_scene_store_tr_file": (C.c_char_p, [P]),
    "vspg_host_scene_store_isg_file": (C.c_char_p, [P]),
    "vspg_host_scene_train_isg": (i32, [P]),
    "vspg_tr_store_begin": (i32, [P]),
    "vspg_tr_store_download": (i32, [P, PF, C.POINTER(i32)]),
    "vspg_tr_store_end": (i32, [P]),
    "vspg_isg_store_begin": (i32, [P]),
    "vspg_isg_store_update": (i32, [P]),
    "vspg_isg_store_download": (i32, [P, PF, PF]),
    "vspg_isg_store_end": (i32, [P]),
    "vspg_exr_write_channels": (i32, [C.c_char_p, i32, i32, i32, C.POINTER(C.c_char_p), C.POINTER(PF)]),
    "vspg_debug_intersect": (i32, [P, i32, PF, C.POINTER(i32), PF]),
    "vspg_debug_sample_li": (i32, [P, i32, PF, C.POINTER(i32), PF, PF]),
    "vspg_debug_light_sample": (i32, [P, i32, PF, PF, PF, C.POINTER(i32), C.POINTER(i32), PF, PF]),
    "vspg_debug_guided_sample": (i32, [P, i32, i32, C.POINTER(GsInput), C.POINTER(GsOutput)]),
    "vspg_debug_sample_ld": (i32, [P, i32, C.POINTER(LdInput), C.POINTER(LdOutput)]),
    "vspg_debug_libm": (i32, [P, i32, i32, PF, PF, PF]),
}


def load_library(path=LIB_PATH):
    """Load libvspg.so and bind every declared symbol.  Raises (never falls back) when the library is missing
    or does not export a symbol include/vspg.h declares."""
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not found: build it with `make -C {os.path.join(PKG_DIR, 'csrc')}` "
            "(or __graft_entry__.build()).  There is no CPU fallback for the render pass.")
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib
