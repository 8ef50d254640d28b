"""The five BASELINE.json configurations as seeded synthetic scenes.  Each builder returns a Config
holding the scene, guiding inputs (or None), integrator parameters, spp and a name.  `res_scale` / `grid_n` let the
tests run the same scene at a size the CPU oracle finishes in seconds.
"""
from dataclasses import dataclass

import numpy as np

from . import _abi as A
from . import api
from .scenes import Scene, perspective_camera
from . import synthetic as syn


@dataclass
class Config:
    name: str
    scene: Scene
    guiding: object
    ip: A.IntegratorParams
    spp: int
    description: str

    @property
    def width(self):
        return self.scene.camera.width

    @property
    def height(self):
        return self.scene.camera.height


def _cornell_room(sc, white, red, green, light_L=(17.0, 12.0, 4.0)):
    # unit Cornell box, open towards -z (camera side)
    sc.add_quad([0, 0, 0], [1, 0, 0], [0, 0, 1], material=white)            # floor
    sc.add_quad([0, 1, 0], [0, 0, 1], [1, 0, 0], material=white)            # ceiling
    sc.add_quad([0, 0, 1], [1, 0, 0], [0, 1, 0], material=white)            # back wall
    sc.add_quad([0, 0, 0], [0, 0, 1], [0, 1, 0], material=red)              # left
    sc.add_quad([1, 0, 0], [0, 1, 0], [0, 0, 1], material=green)            # right
    # emissive quad just below the ceiling, normal pointing down (one-sided)
    sc.add_quad([0.35, 0.998, 0.35], [0.3, 0, 0], [0, 0, 0.3], material=white, emission=light_L)


def config1(res=256, spp=16):
    """C1: Cornell box + homogeneous-medium cube, vspguiding=false, 256x256, 16 spp."""
    cam, eye = perspective_camera([0.5, 0.5, -1.4], [0.5, 0.5, 0.5], [0, 1, 0], 40.0, res, res)
    sc = Scene(cam, eye)
    white, red, green = sc.add_material([.73, .73, .73]), sc.add_material([.65, .05, .05]), sc.add_material([.12, .45, .15])
    _cornell_room(sc, white, red, green)
    med = sc.add_homogeneous_medium([0.5, 0.5, 0.5], [2.0, 2.0, 2.0], 1.0, 0.0)
    sc.add_box([0.3, 0.3, 0.3], [0.7, 0.7, 0.7], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    ip = api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0, light_sampler=A.LIGHTSAMPLER_UNIFORM)
    return Config("C1-cornell-homogeneous", sc.finalize(), None, ip, spp,
                  "Cornell box with homogeneous-medium cube, guidedvolpath vspguiding=false")


def config1_triangles(res=256, spp=16):
    """C1 with the medium cube bounded by a triangle mesh and a diffuse triangle pyramid in the room: the same path
    over Shape "trianglemesh" (watertight ray-triangle test, shapes.cpp:168-263)."""
    cam, eye = perspective_camera([0.5, 0.5, -1.4], [0.5, 0.5, 0.5], [0, 1, 0], 40.0, res, res)
    sc = Scene(cam, eye)
    white, red, green = sc.add_material([.73, .73, .73]), sc.add_material([.65, .05, .05]), sc.add_material([.12, .45, .15])
    _cornell_room(sc, white, red, green)
    med = sc.add_homogeneous_medium([0.5, 0.5, 0.5], [2.0, 2.0, 2.0], 1.0, 0.0)
    sc.add_triangle_box([0.3, 0.3, 0.3], [0.7, 0.7, 0.7], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    apex, b = [0.2, 0.35, 0.75], [[0.05, 0.0, 0.6], [0.35, 0.0, 0.6], [0.35, 0.0, 0.9], [0.05, 0.0, 0.9]]
    sc.add_triangles([apex] + b, [(0, 2, 1), (0, 3, 2), (0, 4, 3), (0, 1, 4)], material=green)
    ip = api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0, light_sampler=A.LIGHTSAMPLER_UNIFORM)
    return Config("C1-cornell-triangles", sc.finalize(), None, ip, spp, "Cornell box, triangle-mesh medium boundary and pyramid, vspguiding=false")


def config1_spheres(res=256, spp=16):
    """C1 with the homogeneous medium bounded by a sphere (interface material) and a diffuse sphere in the room: the
    same path over Shape "sphere" (Sphere::BasicIntersect with interval arithmetic, shapes.h:147-233)."""
    cam, eye = perspective_camera([0.5, 0.5, -1.4], [0.5, 0.5, 0.5], [0, 1, 0], 40.0, res, res)
    sc = Scene(cam, eye)
    white, red, green = sc.add_material([.73, .73, .73]), sc.add_material([.65, .05, .05]), sc.add_material([.12, .45, .15])
    _cornell_room(sc, white, red, green)
    med = sc.add_homogeneous_medium([0.5, 0.5, 0.5], [2.0, 2.0, 2.0], 1.0, 0.0)
    sc.add_sphere([0.55, 0.5, 0.5], 0.28, material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    sc.add_sphere([0.2, 0.15, 0.75], 0.15, material=green)
    ip = api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0, light_sampler=A.LIGHTSAMPLER_UNIFORM)
    return Config("C1-cornell-spheres", sc.finalize(), None, ip, spp, "Cornell box, sphere medium boundary and diffuse sphere, vspguiding=false")


def config1_emitters(res=256, spp=16):
    """C1 lit by an emissive triangle pair (the usual way .pbrt scenes model a ceiling light) and a small emissive
    sphere instead of the bilinear patch: Triangle::Sample(ctx) / Sphere::Sample(ctx) (shapes.h:1053-1173, 300-392)."""
    cam, eye = perspective_camera([0.5, 0.5, -1.4], [0.5, 0.5, 0.5], [0, 1, 0], 40.0, res, res)
    sc = Scene(cam, eye)
    white, red, green = sc.add_material([.73, .73, .73]), sc.add_material([.65, .05, .05]), sc.add_material([.12, .45, .15])
    sc.add_quad([0, 0, 0], [1, 0, 0], [0, 0, 1], material=white)
    sc.add_quad([0, 1, 0], [0, 0, 1], [1, 0, 0], material=white)
    sc.add_quad([0, 0, 1], [1, 0, 0], [0, 1, 0], material=white)
    sc.add_quad([0, 0, 0], [0, 0, 1], [0, 1, 0], material=red)
    sc.add_quad([1, 0, 0], [0, 1, 0], [0, 0, 1], material=green)
    # ceiling light as two triangles, normals pointing down: n = normalize(cross(p1 - p0, p2 - p0))
    P = [[0.35, 0.998, 0.35], [0.65, 0.998, 0.35], [0.65, 0.998, 0.65], [0.35, 0.998, 0.65]]
    sc.add_triangles(P, [(0, 1, 2), (0, 2, 3)], material=white, emission=[14.0, 10.0, 4.0])
    sc.add_sphere([0.82, 0.2, 0.25], 0.06, material=white, emission=[6.0, 9.0, 12.0])
    med = sc.add_homogeneous_medium([0.5, 0.5, 0.5], [2.0, 2.0, 2.0], 1.0, 0.0)
    sc.add_box([0.3, 0.3, 0.3], [0.7, 0.7, 0.7], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    ip = api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0, light_sampler=A.LIGHTSAMPLER_BVH)
    return Config("C1-cornell-emitters", sc.finalize(), None, ip, spp, "Cornell box lit by emissive triangles and an emissive sphere, bvh light sampler, vspguiding=false")


def _guiding_for(sc, rng, width, height, vol_leaves, surf_leaves, with_tr=False):
    g = syn.Guiding()
    e = sc.eye.astype(np.float32)
    lo, hi = np.float32([-0.6, -0.6, -0.6]) - e, np.float32([1.6, 1.6, 1.6]) - e   # render-space scene bounds
    fv, keep_v = syn.synthetic_field(lo, hi, vol_leaves, rng, lobes=8)
    fs, keep_s = syn.synthetic_field(lo, hi, surf_leaves, rng, lobes=8)
    g.set_field("volume", fv, keep_v)
    g.set_field("surface", fs, keep_s)
    g.set_isg(syn.synthetic_isg(width, height, rng))
    if with_tr:
        g.set_tr(syn.synthetic_tr(width, height))
    return g


def _fbm_scene(res, grid_n, density=None, black_patches=False):
    cam, eye = perspective_camera([0.5, 0.6, -2.2], [0.5, 0.45, 0.5], [0, 1, 0], 40.0, res, res)
    sc = Scene(cam, eye)
    white = sc.add_material([.6, .6, .6])
    sc.add_quad([-2, -0.05, -2], [5, 0, 0], [0, 0, 5], material=white)       # ground
    if black_patches:   # reflectance 0: a diffuse BSDF without the DiffuseReflection flag (bxdfs.h:82-84) -- no NEE, no scattering
        black = sc.add_material([0.0, 0.0, 0.0])
        sc.add_quad([-0.6, -0.04, -0.4], [0.9, 0, 0], [0, 0, 0.9], material=black)
        sc.add_quad([1.05, -0.04, 0.1], [0.5, 0, 0], [0, 0.6, 0.3], material=black)
    if density is None:
        density = syn.fbm_grid(grid_n, seed=1337)
    med = sc.add_grid_medium(density, (0, 0, 0), (1, 1, 1), [0.1, 0.1, 0.1], [6.0, 6.0, 6.0], 8.0, 0.6)
    sc.add_box([0, 0, 0], [1, 1, 1], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    sc.add_distant_light([0.4, 1.0, -0.3], [3.0, 3.0, 3.0])
    sc.add_quad([1.4, 1.5, 0.2], [0.4, 0, 0], [0, 0, 0.4], material=white, emission=[40.0, 36.0, 30.0])  # small area light (faces down)
    return sc


def config2(res=512, spp=64, grid_n=128, vol_leaves=4096, surf_leaves=1024, density=None, black_patches=False):
    """C2: fBm 128^3 grid medium, VSPG resampling + variance criterion, loaded cache + ISG buffer, 512x512, 64 spp."""
    sc = _fbm_scene(res, grid_n, density, black_patches)
    rng = np.random.default_rng(2024)
    g = _guiding_for(sc, rng, res, res, vol_leaves, surf_leaves)
    ip = api.integrator_defaults(vsp_sampling_nds=0, vsp_criterion_variance=1, light_sampler=A.LIGHTSAMPLER_POWER)
    return Config("C2-fbm128-vspg-resampling", sc.finalize(), g, ip, spp,
                  "procedural fBm grid medium, guidedvolpath VSPG resampling + variance criterion, loaded guiding cache + ISG buffer")


def config3(res=512, spp=64, grid_n=128, vol_leaves=4096, surf_leaves=1024, density=None):
    """C3: same medium, vspsamplingmethod=nds + collisionProbabilityBias (NDS+) with a loaded transmittance buffer."""
    sc = _fbm_scene(res, grid_n, density)
    rng = np.random.default_rng(2024)
    g = _guiding_for(sc, rng, res, res, vol_leaves, surf_leaves, with_tr=True)
    ip = api.integrator_defaults(vsp_sampling_nds=1, collision_probability_bias=1, light_sampler=A.LIGHTSAMPLER_POWER)
    return Config("C3-fbm128-nds-plus", sc.finalize(), g, ip, spp,
                  "same medium with vspsamplingmethod=nds + collisionProbabilityBias (NDS+) using a loaded transmittance buffer")


def config4(width=1920, height=1080, spp=256, grid_n=512, vol_leaves=8192, surf_leaves=1024, density=None):
    """C4: thick cloud-like 512^3 grid, volume guiding mis + surface guiding ris, 1920x1080, 256 spp."""
    cam, eye = perspective_camera([0.5, 0.35, -2.4], [0.5, 0.45, 0.5], [0, 1, 0], 40.0, width, height)
    sc = Scene(cam, eye)
    white = sc.add_material([.5, .5, .5])
    sc.add_quad([-4, -0.02, -4], [9, 0, 0], [0, 0, 9], material=white)
    if density is None:
        density = syn.cloud_grid(grid_n)
    med = sc.add_grid_medium(density, (0, 0, 0), (1, 1, 1), [0.4, 0.4, 0.4], [39.6, 39.6, 39.6], 1.0, 0.85)
    sc.add_box([0, 0, 0], [1, 1, 1], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    sc.add_distant_light([0.5, 0.8, -0.4], [6.0, 5.6, 5.0])
    sc.add_uniform_infinite_light([0.25, 0.35, 0.5])
    rng = np.random.default_rng(4004)
    g = _guiding_for(sc, rng, width, height, vol_leaves, surf_leaves)
    ip = api.integrator_defaults(volume_guiding_ris=0, surface_guiding_ris=1, light_sampler=A.LIGHTSAMPLER_POWER)
    return Config("C4-cloud512", sc.finalize(), g, ip, spp,
                  "thick procedural cloud-like grid, volume guiding mis + surface guiding ris")


def config5(width=1920, height=1080, spp=1024, grid_n=256, vol_leaves=8192, surf_leaves=4096, grids=None):
    """C5: emissive quads + diffuse boxes (BVH) + chromatic heterogeneous medium (rgbgrid), full VSPG."""
    cam, eye = perspective_camera([0.5, 0.55, -2.3], [0.5, 0.4, 0.5], [0, 1, 0], 40.0, width, height)
    sc = Scene(cam, eye)
    white, grey = sc.add_material([.7, .7, .7]), sc.add_material([.35, .35, .4])
    sc.add_quad([-3, -0.3, -3], [7, 0, 0], [0, 0, 7], material=white)
    rng = np.random.default_rng(5005)
    for i in range(12):   # diffuse boxes scattered around the medium
        c = np.array([rng.uniform(-1.0, 2.0), -0.3, rng.uniform(-0.6, 2.2)])
        if -0.15 < c[0] < 1.15 and -0.15 < c[2] < 1.15:
            c[0] += 1.4
        s = rng.uniform(0.12, 0.3, size=3)
        sc.add_box(c, c + s, material=grey if i % 2 else white)
    for i, (o, L) in enumerate((([-0.6, 1.6, 0.2], [30, 6, 4]), ([0.35, 1.7, 0.35], [6, 28, 8]), ([1.3, 1.6, 0.5], [5, 8, 30]))):
        sc.add_quad(o, [0.3, 0, 0], [0, 0, 0.3], material=white, emission=L)      # facing down
    if grids is None:
        grids = syn.rgb_grids(grid_n)
    med = sc.add_rgbgrid_medium(grids[0], grids[1], (0, 0, 0), (1, 1, 1), 1.0, 0.3)
    sc.add_box([0, 0, 0], [1, 1, 1], material=-1, medium_inside=med, medium_outside=-1, has_interface=True)
    g = _guiding_for(sc, rng, width, height, vol_leaves, surf_leaves)
    ip = api.integrator_defaults(light_sampler=A.LIGHTSAMPLER_POWER)
    return Config("C5-mixed-rgbgrid", sc.finalize(), g, ip, spp,
                  "mixed scene: emissive quads + diffuse surfaces (BVH) + chromatic heterogeneous medium, full VSPG")


BUILDERS = {"C1T": config1_triangles, "C1S": config1_spheres, "C1E": config1_emitters, "C1": config1, "C2": config2, "C3": config3, "C4": config4, "C5": config5}
