"""Scene / guiding-input containers and the seeded synthetic inputs of the BASELINE.json configs.

Everything here only fills the POD structs of include/vspg.h with numpy data (host plumbing); the arithmetic of the
render pass lives in the CUDA library.  The same structs are handed to the CPU oracle by the tests.
"""
import ctypes as C
import math

import numpy as np

from . import _abi as A


# --------------------------------------------------------------------------- transforms
def _to_transform(m):
    m = np.asarray(m, dtype=np.float64)
    t = A.Transform()
    mi = np.linalg.inv(m)
    for r in range(4):
        for c in range(4):
            t.m[r][c] = float(np.float32(m[r, c]))
            t.mInv[r][c] = float(np.float32(mi[r, c]))
    return t


def translate(v):
    m = np.eye(4)
    m[:3, 3] = v
    return m


def scale(x, y, z):
    return np.diag([x, y, z, 1.0])


def perspective_camera(eye, look, up, fov_deg, width, height, lens_radius=0.0, focal_distance=1e6, medium=-1):
    """cameras.h ProjectiveCamera/PerspectiveCamera ctor + CameraTransform with the default "cameraworld" render
    space (render = world translated by -eye).  Returns (vspg_camera, eye) -- scene geometry must be given in render
    space, i.e. world - eye (Scene does that)."""
    eye, look, up = (np.asarray(v, dtype=np.float64) for v in (eye, look, up))
    d = (look - eye) / np.linalg.norm(look - eye)
    right = np.cross(up / np.linalg.norm(up), d)
    right /= np.linalg.norm(right)
    new_up = np.cross(d, right)
    world_from_camera = np.eye(4)
    world_from_camera[:3, 0], world_from_camera[:3, 1], world_from_camera[:3, 2] = right, new_up, d
    render_from_camera = world_from_camera.copy()  # translation removed: camera sits at the render-space origin
    n, f = 1e-2, 1000.0
    persp = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, f / (f - n), -f * n / (f - n)], [0, 0, 1, 0]], dtype=np.float64)
    inv_tan = 1.0 / math.tan(math.radians(fov_deg) / 2)
    screen_from_camera = scale(inv_tan, inv_tan, 1) @ persp
    aspect = width / height
    if aspect > 1:
        smin, smax = (-aspect, -1.0), (aspect, 1.0)
    else:
        smin, smax = (-1.0, -1 / aspect), (1.0, 1 / aspect)
    ndc_from_screen = scale(1 / (smax[0] - smin[0]), 1 / (smax[1] - smin[1]), 1) @ translate([-smin[0], -smax[1], 0])
    raster_from_ndc = scale(width, -height, 1)
    raster_from_screen = raster_from_ndc @ ndc_from_screen
    camera_from_raster = np.linalg.inv(screen_from_camera) @ np.linalg.inv(raster_from_screen)
    cam = A.Camera()
    cam.camera_from_raster = _to_transform(camera_from_raster)
    cam.render_from_camera = _to_transform(render_from_camera)
    cam.lens_radius, cam.focal_distance, cam.medium = lens_radius, focal_distance, medium
    cam.width, cam.height = width, height
    cam.filter_radius[0] = cam.filter_radius[1] = 0.5
    cam.filter_type, cam.filter_sigma = 0, 0.0
    cam.imaging_ratio = 1.0                      # film.cpp:233 exposureTime * ISO / 100 with the defaults
    cam.max_component_value = float("inf")
    cam._params = dict(eye=eye.tolist(), look=look.tolist(), up=up.tolist(), fov=float(fov_deg), lens_radius=float(lens_radius),
                       focal_distance=float(focal_distance))   # for pbrt_writer
    return cam, eye


# --------------------------------------------------------------------------- scene container
class Scene:
    """Collects shapes / materials / lights / media in WORLD space and exposes a vspg_scene_desc in render space."""

    def __init__(self, camera, eye):
        self.camera, self.eye = camera, np.asarray(eye, dtype=np.float64)
        self.quads, self.triangles, self.spheres, self.materials, self.lights, self.media = [], [], [], [], [], []
        self._keep = []
        self.desc = None
        self.log = []   # world-space record of every add_* call, in order (pbrt_writer turns it into a .pbrt file)

    def add_material(self, reflectance):
        self.log.append(("material", dict(reflectance=[float(x) for x in reflectance], index=len(self.materials))))
        m = A.Material()
        m.type = 0
        m.reflectance[:] = [float(x) for x in reflectance]
        self.materials.append(m)
        return len(self.materials) - 1

    def add_quad(self, origin, eu, ev, material=-1, emission=None, two_sided=False, medium_inside=-1, medium_outside=-1,
                 has_interface=False, corners=None):
        """Bilinear patch p00=o, p10=o+eu, p01=o+ev, p11=o+eu+ev (geometric normal = normalize(eu x ev))."""
        if corners is None:
            o, eu, ev = (np.asarray(v, dtype=np.float64) for v in (origin, eu, ev))
            corners = [o, o + eu, o + ev, o + eu + ev]
        self.log.append(("quad", dict(corners=[[float(x) for x in np.asarray(p, dtype=np.float64)] for p in corners], material=material,
                                      emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided),
                                      medium_inside=medium_inside, medium_outside=medium_outside, has_interface=bool(has_interface))))
        q = A.Quad()
        for name, p in zip(("p00", "p10", "p01", "p11"), corners):
            getattr(q, name)[:] = [float(np.float32(x)) for x in (np.asarray(p, dtype=np.float64) - self.eye)]
        q.material, q.area_light = material, -1
        q.medium_inside, q.medium_outside, q.has_interface, q.flip_normal = medium_inside, medium_outside, int(has_interface), 0
        self.quads.append(q)
        qi = len(self.quads) - 1
        if emission is not None:
            l = A.Light()
            l.type = A.LIGHT_DIFFUSE_AREA
            l.L[:] = [float(x) for x in emission]
            l.shape_kind, l.shape_index, l.two_sided, l.medium = 0, qi, int(two_sided), -1
            self.lights.append(l)
            q.area_light = len(self.lights) - 1
        return qi

    def _area_light(self, kind, index, emission, two_sided):
        l = A.Light()
        l.type = A.LIGHT_DIFFUSE_AREA
        l.L[:] = [float(x) for x in emission]
        l.shape_kind, l.shape_index, l.two_sided, l.medium = kind, index, int(two_sided), -1
        self.lights.append(l)
        return len(self.lights) - 1

    def add_triangles(self, points, indices, material=-1, medium_inside=-1, medium_outside=-1, has_interface=False, emission=None,
                      two_sided=False):
        """Shape "trianglemesh": points [n][3] (world space), indices [m][3].  With `emission` every triangle becomes
        its own DiffuseAreaLight, as the reference creates one light per shape."""
        pts = np.asarray(points, dtype=np.float64).reshape(-1, 3)
        idx = np.asarray(indices, dtype=np.int64).reshape(-1, 3)
        self.log.append(("trimesh", dict(points=pts.tolist(), indices=idx.tolist(), material=material, medium_inside=medium_inside,
                                         medium_outside=medium_outside, has_interface=bool(has_interface),
                                         emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided))))
        for tri in idx:
            t = A.Triangle()
            for name, vi in zip(("p0", "p1", "p2"), tri):
                getattr(t, name)[:] = [float(np.float32(x)) for x in (pts[vi] - self.eye)]
            t.material, t.area_light = material, -1
            t.medium_inside, t.medium_outside, t.has_interface, t.flip_normal = medium_inside, medium_outside, int(has_interface), 0
            self.triangles.append(t)
            if emission is not None:
                t.area_light = self._area_light(1, len(self.triangles) - 1, emission, two_sided)

    def add_sphere(self, center, radius, material=-1, medium_inside=-1, medium_outside=-1, has_interface=False, emission=None, two_sided=False):
        """Shape "sphere" (full sphere, shapes.h:107) under Translate(center); outward normal."""
        c = np.asarray(center, dtype=np.float64)
        self.log.append(("sphere", dict(center=[float(x) for x in c], radius=float(radius), material=material, medium_inside=medium_inside,
                                        medium_outside=medium_outside, has_interface=bool(has_interface),
                                        emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided))))
        sp = A.Sphere()
        sp.center[:] = [float(np.float32(x)) for x in (c - self.eye)]
        sp.radius = float(np.float32(radius))
        sp.material, sp.area_light = material, -1
        sp.medium_inside, sp.medium_outside, sp.has_interface, sp.flip_normal = medium_inside, medium_outside, int(has_interface), 0
        self.spheres.append(sp)
        if emission is not None:
            sp.area_light = self._area_light(2, len(self.spheres) - 1, emission, two_sided)
        return len(self.spheres) - 1

    def add_triangle_box(self, lo, hi, **kw):
        """Axis-aligned box as 12 outward-facing triangles (n = normalize(cross(p0 - p2, p1 - p2)))."""
        a, b = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
        P = [[a[0], a[1], a[2]], [b[0], a[1], a[2]], [b[0], b[1], a[2]], [a[0], b[1], a[2]],
             [a[0], a[1], b[2]], [b[0], a[1], b[2]], [b[0], b[1], b[2]], [a[0], b[1], b[2]]]
        quads = [(0, 3, 2, 1), (4, 5, 6, 7), (0, 1, 5, 4), (3, 7, 6, 2), (0, 4, 7, 3), (1, 2, 6, 5)]   # outward (counter-clockwise from outside)
        idx = []
        for q in quads:
            idx += [(q[0], q[1], q[2]), (q[0], q[2], q[3])]
        self.add_triangles(P, idx, **kw)

    def add_box(self, lo, hi, material=-1, medium_inside=-1, medium_outside=-1, has_interface=False):
        """Axis-aligned box with outward-facing normals."""
        a, b = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
        s = b - a
        X, Y, Z = np.array([s[0], 0, 0]), np.array([0, s[1], 0]), np.array([0, 0, s[2]])
        kw = dict(material=material, medium_inside=medium_inside, medium_outside=medium_outside, has_interface=has_interface)
        self.add_quad(a, Z, Y, **kw)                       # -x
        self.add_quad([b[0], a[1], a[2]], Y, Z, **kw)      # +x
        self.add_quad(a, X, Z, **kw)                       # -y
        self.add_quad([a[0], b[1], a[2]], Z, X, **kw)      # +y
        self.add_quad(a, Y, X, **kw)                       # -z
        self.add_quad([a[0], a[1], b[2]], X, Y, **kw)      # +z

    def add_point_light(self, pos, intensity, medium=-1):
        self.log.append(("point", dict(pos=[float(x) for x in pos], I=[float(x) for x in intensity], medium=medium)))
        l = A.Light()
        l.type = A.LIGHT_POINT
        l.L[:] = [float(x) for x in intensity]
        l.pos[:] = [float(np.float32(x)) for x in (np.asarray(pos, dtype=np.float64) - self.eye)]
        l.shape_kind, l.shape_index, l.two_sided, l.medium = 0, -1, 0, medium
        self.lights.append(l)

    def add_distant_light(self, to_light, radiance):
        self.log.append(("distant", dict(to_light=[float(x) for x in to_light], L=[float(x) for x in radiance])))
        l = A.Light()
        l.type = A.LIGHT_DISTANT
        l.L[:] = [float(x) for x in radiance]
        l.pos[:] = [float(x) for x in to_light]
        l.shape_kind, l.shape_index, l.two_sided, l.medium = 0, -1, 0, -1
        self.lights.append(l)

    def add_uniform_infinite_light(self, radiance):
        self.log.append(("infinite", dict(L=[float(x) for x in radiance])))
        l = A.Light()
        l.type = A.LIGHT_UNIFORM_INFINITE
        l.L[:] = [float(x) for x in radiance]
        l.shape_kind, l.shape_index, l.two_sided, l.medium = 0, -1, 0, -1
        self.lights.append(l)

    def _medium_common(self, mtype, lo, hi, sigma_a, sigma_s, sigma_scale, g, world_from_medium):
        m = A.Medium()
        m.type = mtype
        m.bounds_min[:] = [float(x) for x in lo]
        m.bounds_max[:] = [float(x) for x in hi]
        wfm = np.eye(4) if world_from_medium is None else np.asarray(world_from_medium, dtype=np.float64)
        m.render_from_medium = _to_transform(translate(-self.eye) @ wfm)
        m.sigma_a[:] = [float(x) for x in sigma_a]
        m.sigma_s[:] = [float(x) for x in sigma_s]
        m.sigma_scale, m.g = float(sigma_scale), float(g)
        m.maj_res[:] = [16, 16, 16]
        return m

    def add_homogeneous_medium(self, sigma_a, sigma_s, scale_=1.0, g=0.0):
        # media.h:240-242: sigma_* spectra scaled by "scale" in the ctor
        self.log.append(("medium", dict(type="homogeneous", sigma_a=[float(x) for x in sigma_a], sigma_s=[float(x) for x in sigma_s],
                                        scale=float(scale_), g=float(g), index=len(self.media))))
        m = self._medium_common(A.MEDIUM_HOMOGENEOUS, (0, 0, 0), (1, 1, 1), np.float32(sigma_a) * np.float32(scale_),
                                np.float32(sigma_s) * np.float32(scale_), 1.0, g, None)
        self.media.append(m)
        return len(self.media) - 1

    def add_grid_medium(self, density, lo, hi, sigma_a, sigma_s, scale_=1.0, g=0.0, world_from_medium=None):
        """uniformgrid (media.cpp:272-361): density[nz][ny][nx] float32, medium-space bounds lo..hi."""
        d = np.ascontiguousarray(density, dtype=np.float32)
        self._keep.append(d)
        self.log.append(("medium", dict(type="uniformgrid", density=d, lo=[float(x) for x in lo], hi=[float(x) for x in hi],
                                        sigma_a=[float(x) for x in sigma_a], sigma_s=[float(x) for x in sigma_s], scale=float(scale_), g=float(g),
                                        world_from_medium=world_from_medium, index=len(self.media))))
        m = self._medium_common(A.MEDIUM_GRID, lo, hi, np.float32(sigma_a) * np.float32(scale_),
                                np.float32(sigma_s) * np.float32(scale_), 1.0, g, world_from_medium)
        m.nz, m.ny, m.nx = d.shape
        m.density = d.ctypes.data_as(A.PF)
        self.media.append(m)
        return len(self.media) - 1

    def add_rgbgrid_medium(self, sigma_a_rgb, sigma_s_rgb, lo, hi, scale_=1.0, g=0.0, world_from_medium=None):
        """rgbgrid (media.cpp:411-484): two [nz][ny][nx][3] float32 grids, sigma_scale applied at lookup."""
        sa = np.ascontiguousarray(sigma_a_rgb, dtype=np.float32)
        ss = np.ascontiguousarray(sigma_s_rgb, dtype=np.float32)
        self._keep += [sa, ss]
        self.log.append(("medium", dict(type="rgbgrid", sigma_a_rgb=sa, sigma_s_rgb=ss, lo=[float(x) for x in lo], hi=[float(x) for x in hi],
                                        scale=float(scale_), g=float(g), world_from_medium=world_from_medium, index=len(self.media))))
        m = self._medium_common(A.MEDIUM_RGBGRID, lo, hi, (1, 1, 1), (1, 1, 1), scale_, g, world_from_medium)
        m.nz, m.ny, m.nx = sa.shape[:3]
        m.sigma_a_rgb = sa.ctypes.data_as(A.PF)
        m.sigma_s_rgb = ss.ctypes.data_as(A.PF)
        self.media.append(m)
        return len(self.media) - 1

    def finalize(self):
        d = A.SceneDesc()
        d.camera = self.camera
        arrs = {}
        for name, items, T in (("media", self.media, A.Medium), ("quads", self.quads, A.Quad), ("triangles", self.triangles, A.Triangle),
                               ("spheres", self.spheres, A.Sphere),
                               ("materials", self.materials, A.Material), ("lights", self.lights, A.Light)):
            arr = (T * max(1, len(items)))(*items)
            arrs[name] = arr
            setattr(d, "n_" + name, len(items))
            setattr(d, name, arr)
        self._arrs = arrs
        self.desc = d
        return self

    @property
    def n_pixels(self):
        return self.camera.width * self.camera.height
