"""The small scenes the reference-pinning fixtures are made of (tests/golden/pbrt/*.pbrt, written by
tests/golden/make_golden_pbrt.py with the reference renderer; read by tests/test_ref_pbrt.py).  One scene per BASELINE
config family, all 40x24 (the size of the golden ISG buffer) or 32x32, 4 spp, with 6x5x4 grids so the .pbrt text stays
a few KiB.  File names inside the scenes are relative to tests/golden/pbrt/."""
import os

import numpy as np

from pyvspg import _abi as A
from pyvspg import api, configs, pbrt_writer
from pyvspg.scenes import Scene, perspective_camera

SCENES = ["c1", "c1t", "c1e", "c1s", "c2", "c2u", "c3", "c3nds", "c3contrib", "c4", "c4r", "c5", "c5u", "ch", "cg", "cn", "cnu", "cp", "cs", "csp", "ce", "ceg", "cet"]


def write_inputs(golden_dir):
    """isg_min.exr (VSP.Y of the reference-made golden buffer + the surfContrib layer the reference's loader keys on)
    and tr_small.exr (Transmittance.R/G/B), both through the product's EXR writer."""
    vsp = np.load(os.path.join(golden_dir, "isg_small_vsp.npy"))
    z = np.zeros_like(vsp)
    api.exr_write_channels(os.path.join(golden_dir, "isg_min.exr"), {"VSP.Y": vsp, "surfContrib.R": z, "surfContrib.G": z, "surfContrib.B": z})
    tr = np.random.default_rng(5).random((24, 40, 3), dtype=np.float32)
    api.exr_write_channels(os.path.join(golden_dir, "tr_small.exr"), {"Transmittance.R": tr[..., 0], "Transmittance.G": tr[..., 1], "Transmittance.B": tr[..., 2]})


def _guided(name, kind="grid", filt=None, lights="pa", emit=None, **ipkw):
    cam, eye = perspective_camera([0.5, 0.6, -2.2], [0.5, 0.45, 0.5], [0, 1, 0], 40.0, 40, 24)
    sc = Scene(cam, eye)
    white, red = sc.add_material([.6, .6, .6]), sc.add_material([.7, .2, .1])
    sc.add_quad([-2, -0.05, -2], [5, 0, 0], [0, 0, 5], material=white)
    rng = np.random.default_rng(5)
    if kind == "grid":
        dens = rng.random((6, 5, 4), dtype=np.float32)
        ekw = {}
        if emit == "grid":     # "Le" scaled by a "Lescale" grid with some zero voxels (media.cpp:317-329)
            les = np.random.default_rng(9).random((6, 5, 4), dtype=np.float32)
            ekw = dict(Le=[1.5, 0.9, 0.4], Le_scale=np.where(les > 0.3, les * 2, 0).astype(np.float32))
        elif emit == "tent":   # "Le" alone: the reference's 1x1x1 scale grid
            ekw = dict(Le=[0.8, 1.1, 1.6])
        med = sc.add_grid_medium(dens, (0, 0, 0), (1, 1, 1), [0.1, 0.1, 0.1], [6.0, 6.0, 6.0], 8.0, 0.6, **ekw)
    elif kind == "rgb":
        sa = rng.random((6, 5, 4, 3), dtype=np.float32) * np.float32([0.02, 0.05, 0.2])
        ss = rng.random((6, 5, 4, 3), dtype=np.float32) * np.float32([4, 3, 2])
        med = sc.add_rgbgrid_medium(sa, ss, (0, 0, 0), (1, 1, 1), 3.0, 0.4)
    elif kind == "nvdb":
        # tests/golden/nvdb_small.nvdb: written by the reference's vendored NanoVDB (oracle/ref_harness/nvdb_writer.cpp, see
        # make_golden_pbrt.py): a sparse 19x17x15 blob at index offset (-4, 4, 101), voxel 0.05, world box (.02,.05,.1)-(.97,.9,.85)
        med = sc.add_nanovdb_medium(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nvdb_small.nvdb"), [0.1, 0.1, 0.1], [3.0, 3.0, 3.0],
                                    4.0, 0.5, density_offset=0.02, majorant_scale=1.25, written_name="../nvdb_small.nvdb")
    else:
        med = sc.add_homogeneous_medium([0.2, 0.3, 0.5], [3.0, 2.0, 1.5], 1.5, 0.3, **(dict(Le=[0.6, 0.5, 0.9], Le_scale=1.5) if emit else {}))
    sc.add_box([0, 0, 0], [1, 1, 1], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    if "p" in lights:
        sc.add_point_light([0.2, 1.8, -0.4], [9.0, 8.0, 7.0])
    if "a" in lights:
        sc.add_quad([1.4, 1.5, 0.2], [0.4, 0, 0], [0, 0, 0.4], material=white, emission=[40.0, 36.0, 30.0])
    if "s" in lights:   # LightSource "spot": one with the default smooth-step falloff, one with conedeltaangle 0 (hard edge)
        sc.add_spot_light([0.3, 1.9, -0.5], [0.5, 0.4, 0.5], [30.0, 26.0, 22.0], coneangle=25.0, conedeltaangle=8.0)
        sc.add_spot_light([1.6, 1.2, 0.3], [0.6, 0.2, 0.6], [20.0, 25.0, 30.0], coneangle=18.0, conedeltaangle=0.0)
    if "d" in lights:
        sc.add_distant_light([0.4, 1.0, -0.3], [3.0, 2.5, 2.0])
    if "i" in lights:
        sc.add_uniform_infinite_light([0.3, 0.4, 0.6])
    if "t" in lights:
        sc.add_triangles([[-0.6, 0.9, 0.1], [-0.3, 1.0, 0.2], [-0.5, 0.95, 0.5], [-0.2, 1.2, 0.6]], [[0, 1, 2], [1, 3, 2]], material=red,
                         emission=[8.0, 9.0, 12.0], two_sided=True)
    if "m" in lights:   # Shape "plymesh": tests/golden/mesh_small.ply (4 triangles + 2 quads, binary), emissive, and its ascii twin as a diffuse object
        gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
        sc.add_plymesh(os.path.join(gold, "mesh_small.ply"), 4, 2, material=red, emission=[14.0, 12.0, 9.0], two_sided=True, translate=[-0.7, 0.7, 0.1],
                       written_name="../mesh_small.ply")
        sc.add_plymesh(os.path.join(gold, "mesh_small_ascii.ply"), 4, 2, material=white, translate=[1.25, -0.05, -0.3], written_name="../mesh_small_ascii.ply")
    if "b" in lights:
        sc.add_triangle_box([1.2, -0.05, 0.6], [1.5, 0.4, 0.9], material=red)
    ip = api.integrator_defaults(**ipkw)
    cfg = configs.Config(name, sc.finalize(), None, ip, 4, "")
    if filt == "gauss":   # pbrt's default pixel filter
        for c in (cfg.scene.desc.camera, cfg.scene.camera):
            c.filter_type, c.filter_sigma = 1, 0.5
            c.filter_radius[0] = c.filter_radius[1] = 1.5
    return cfg, dict(cache_file="../field_small.pgl", isg_file="../isg_min.exr",
                     tr_file="../tr_small.exr" if ipkw.get("collision_probability_bias", 0) else None)


def _config(name):
    U, P, B = A.LIGHTSAMPLER_UNIFORM, A.LIGHTSAMPLER_POWER, A.LIGHTSAMPLER_BVH
    if name == "c1":
        return configs.config1(res=32, spp=4), {}
    if name == "c1t":
        return configs.config1_triangles(res=32, spp=4), {}
    if name == "c1e":
        return configs.config1_emitters(res=32, spp=4), {}
    if name == "c1s":
        return configs.config1_spheres(res=32, spp=4), {}
    if name == "c2":
        return _guided(name, light_sampler=P)
    if name == "c2u":   # guiding off, NEE on: the unguided VolPath core over the grid medium
        return _guided(name, light_sampler=P, surface_guiding=0, volume_guiding=0, vsp_guiding=0)
    if name == "c3":
        return _guided(name, light_sampler=P, vsp_sampling_nds=1, collision_probability_bias=1)
    if name == "c3nds":
        return _guided(name, light_sampler=P, vsp_sampling_nds=1)
    if name == "c3contrib":
        return _guided(name, light_sampler=U, vsp_criterion_variance=0)
    if name == "c4":
        return _guided(name, lights="di", light_sampler=B, volume_guiding_ris=0, surface_guiding_ris=1)
    if name == "c4r":
        return _guided(name, lights="dia", light_sampler=B, volume_guiding_ris=1, surface_guiding_ris=0)
    if name == "c5":
        return _guided(name, kind="rgb", lights="patdb", light_sampler=B)
    if name == "c5u":
        return _guided(name, kind="rgb", lights="patdb", light_sampler=U)
    if name == "cn":     # NanoVDB medium, full VSPG
        return _guided(name, kind="nvdb", light_sampler=B)
    if name == "cnu":    # NanoVDB medium, guiding off
        return _guided(name, kind="nvdb", light_sampler=P, surface_guiding=0, volume_guiding=0, vsp_guiding=0)
    if name == "cp":     # PLY meshes (triangles + bilinear patches from one file), emissive and diffuse
        return _guided(name, lights="pm", light_sampler=B)
    if name == "ce":     # emissive homogeneous medium, guided
        return _guided(name, kind="hom", emit="hom", light_sampler=B)
    if name == "ceg":    # emissive grid medium with a "Lescale" grid, NDS distance sampling (the delta-tracking branch adds the emission)
        return _guided(name, emit="grid", light_sampler=P, vsp_sampling_nds=1)
    if name == "cet":    # emissive grid medium without "Lescale", unguided NDS
        return _guided(name, emit="tent", light_sampler=U, vsp_sampling_nds=1, surface_guiding=0, volume_guiding=0, vsp_guiding=0)
    if name == "cs":     # spot lights, light BVH
        return _guided(name, lights="sa", light_sampler=B)
    if name == "csp":    # spot lights, power light sampler, guiding off
        return _guided(name, lights="sat", light_sampler=P, surface_guiding=0, volume_guiding=0, vsp_guiding=0)
    if name == "ch":
        return _guided(name, kind="hom", light_sampler=B, vsp_mis_ratio=0.3)
    if name == "cg":
        return _guided(name, filt="gauss", light_sampler=B)
    raise KeyError(name)


def build_scene(name, out_dir):
    """Writes <out_dir>/<name>.pbrt and returns what the generator needs to know about it."""
    cfg, files = _config(name)
    path = pbrt_writer.write_scene(cfg, os.path.join(out_dir, name + ".pbrt"), outfile=name + ".pfm", seed=0, **{k: v for k, v in files.items() if v})
    d = cfg.scene.desc
    return dict(path=path, width=d.camera.width, height=d.camera.height, spp=cfg.spp, n_lights=d.n_lights,
                media=[f"medium{i}" for i in range(d.n_media)], eye=[float(x) for x in cfg.scene.eye])
