"""Scene / guiding-input containers and the seeded synthetic inputs of the BASELINE.json configs.

Everything here only fills the POD structs of include/vspg.h with numpy data (host plumbing); the arithmetic of the
render pass lives in the CUDA library.  The same structs are handed to the CPU oracle by the tests.
"""
import numpy as np

from . import _abi as A


# --------------------------------------------------------------------------- camera
class CameraSpec:
    """The Camera / Film / PixelFilter directives of a scene, in world space (what pbrt_writer prints).  The POD
    vspg_camera -- cameraFromRaster, renderFromCamera in the reference's float arithmetic -- is produced by the C++
    front end when the scene is finalised (csrc/host_scene.cpp, csrc/host_xform.h)."""

    def __init__(self, eye, look, up, fov_deg, width, height, lens_radius, focal_distance, medium):
        self.width, self.height, self.medium = int(width), int(height), medium
        self.filter_type, self.filter_sigma, self.filter_radius = 0, 0.0, [0.5, 0.5]   # PixelFilter "box"
        self._params = dict(eye=[float(x) for x in eye], look=[float(x) for x in look], up=[float(x) for x in up], fov=float(fov_deg),
                            lens_radius=float(lens_radius), focal_distance=float(focal_distance))


def perspective_camera(eye, look, up, fov_deg, width, height, lens_radius=0.0, focal_distance=1e6, medium=-1):
    """LookAt + Camera "perspective" with the default "cameraworld" render space (render = world translated by
    -eye, cameras.cpp:26-57).  Returns (CameraSpec, eye)."""
    return CameraSpec(eye, look, up, fov_deg, width, height, lens_radius, focal_distance, medium), np.asarray(eye, dtype=np.float64)


# --------------------------------------------------------------------------- scene container
class Scene:
    """Collects shapes / materials / lights / media in WORLD space (self.log) and, on finalize(), hands the scene to
    the C++ front end as .pbrt text: self.desc is the vspg_scene_desc the front end built (render space, float
    transforms composed exactly as the reference composes them).  The tests hand the same desc to the CPU oracle."""

    def __init__(self, camera, eye):
        self.camera, self.eye = camera, np.asarray(eye, dtype=np.float64)
        self.n_quads = self.n_triangles = self.n_spheres = self.n_materials = self.n_lights = self.n_media = 0
        self._keep = []
        self._desc = None
        self._hs = None
        self.log = []   # world-space record of every add_* call, in order (pbrt_writer turns it into a .pbrt file)

    def add_material(self, reflectance):
        self.log.append(("material", dict(reflectance=[float(x) for x in reflectance], index=self.n_materials)))
        self.n_materials += 1
        return self.n_materials - 1

    def add_quad(self, origin, eu, ev, material=-1, emission=None, two_sided=False, medium_inside=-1, medium_outside=-1,
                 has_interface=False, corners=None):
        """Bilinear patch p00=o, p10=o+eu, p01=o+ev, p11=o+eu+ev (geometric normal = normalize(eu x ev))."""
        if corners is None:
            o, eu, ev = (np.asarray(v, dtype=np.float64) for v in (origin, eu, ev))
            corners = [o, o + eu, o + ev, o + eu + ev]
        self.log.append(("quad", dict(corners=[[float(x) for x in np.asarray(p, dtype=np.float64)] for p in corners], material=material,
                                      emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided),
                                      medium_inside=medium_inside, medium_outside=medium_outside, has_interface=bool(has_interface))))
        self.n_quads += 1
        if emission is not None:
            self.n_lights += 1
        return self.n_quads - 1

    def add_triangles(self, points, indices, material=-1, medium_inside=-1, medium_outside=-1, has_interface=False, emission=None,
                      two_sided=False):
        """Shape "trianglemesh": points [n][3] (world space), indices [m][3].  With `emission` every triangle becomes
        its own DiffuseAreaLight, as the reference creates one light per shape."""
        pts = np.asarray(points, dtype=np.float64).reshape(-1, 3)
        idx = np.asarray(indices, dtype=np.int64).reshape(-1, 3)
        self.log.append(("trimesh", dict(points=pts.tolist(), indices=idx.tolist(), material=material, medium_inside=medium_inside,
                                         medium_outside=medium_outside, has_interface=bool(has_interface),
                                         emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided))))
        self.n_triangles += len(idx)
        if emission is not None:
            self.n_lights += len(idx)

    def add_sphere(self, center, radius, material=-1, medium_inside=-1, medium_outside=-1, has_interface=False, emission=None, two_sided=False):
        """Shape "sphere" (full sphere, shapes.h:107) under Translate(center); outward normal."""
        c = np.asarray(center, dtype=np.float64)
        self.log.append(("sphere", dict(center=[float(x) for x in c], radius=float(radius), material=material, medium_inside=medium_inside,
                                        medium_outside=medium_outside, has_interface=bool(has_interface),
                                        emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided))))
        self.n_spheres += 1
        if emission is not None:
            self.n_lights += 1
        return self.n_spheres - 1

    def add_plymesh(self, filename, n_triangles, n_quads, material=-1, emission=None, two_sided=False, translate=None, written_name=None):
        """Shape "plymesh" (shapes.cpp:1415-1472): the front end reads the file; triangles come first, then the quads as
        bilinear patches.  The face counts are the caller's (who wrote the file); finalize checks them."""
        import os
        self.log.append(("plymesh", dict(filename=os.path.abspath(filename), written_name=written_name, material=material,
                                         emission=None if emission is None else [float(x) for x in emission], two_sided=bool(two_sided),
                                         translate=None if translate is None else [float(x) for x in translate])))
        self.n_triangles += n_triangles
        self.n_quads += n_quads
        if emission is not None:
            self.n_lights += n_triangles + n_quads

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
        self.n_lights += 1

    def add_spot_light(self, pos, to, intensity, coneangle=30.0, conedeltaangle=5.0, medium=-1):
        """LightSource "spot" (lights.cpp:1505-1545): at `pos`, looking at `to`, full intensity inside
        coneangle - conedeltaangle degrees of the axis and a smooth-step falloff to zero at coneangle."""
        self.log.append(("spot", dict(pos=[float(x) for x in pos], to=[float(x) for x in to], I=[float(x) for x in intensity],
                                      coneangle=float(coneangle), conedeltaangle=float(conedeltaangle), medium=medium)))
        self.n_lights += 1

    def add_distant_light(self, to_light, radiance):
        self.log.append(("distant", dict(to_light=[float(x) for x in to_light], L=[float(x) for x in radiance])))
        self.n_lights += 1

    def add_uniform_infinite_light(self, radiance):
        self.log.append(("infinite", dict(L=[float(x) for x in radiance])))
        self.n_lights += 1

    def add_homogeneous_medium(self, sigma_a, sigma_s, scale_=1.0, g=0.0, Le=None, Le_scale=1.0):
        """homogeneous (media.cpp:164-206); Le (rgb) x Le_scale = emitted radiance (media.h:243, 259)."""
        self.log.append(("medium", dict(type="homogeneous", sigma_a=[float(x) for x in sigma_a], sigma_s=[float(x) for x in sigma_s],
                                        scale=float(scale_), g=float(g), index=self.n_media,
                                        Le=None if Le is None else [float(x) for x in Le], Le_scale=float(Le_scale))))
        self.n_media += 1
        return self.n_media - 1

    def add_grid_medium(self, density, lo, hi, sigma_a, sigma_s, scale_=1.0, g=0.0, world_from_medium=None, Le=None, Le_scale=None):
        """uniformgrid (media.cpp:272-361): density[nz][ny][nx] float32, medium-space bounds lo..hi.  Le (rgb): emission
        spectrum; Le_scale: optional [nz][ny][nx] float32 grid scaling it (without one the reference uses a 1x1x1 grid of 1,
        i.e. a trilinear tent around the centre of the box)."""
        d = np.ascontiguousarray(density, dtype=np.float32)
        self._keep.append(d)
        if Le_scale is not None:
            Le_scale = np.ascontiguousarray(Le_scale, dtype=np.float32)
            assert Le is not None and Le_scale.shape == d.shape
            self._keep.append(Le_scale)
        self.log.append(("medium", dict(type="uniformgrid", density=d, lo=[float(x) for x in lo], hi=[float(x) for x in hi],
                                        sigma_a=[float(x) for x in sigma_a], sigma_s=[float(x) for x in sigma_s], scale=float(scale_), g=float(g),
                                        world_from_medium=world_from_medium, index=self.n_media,
                                        Le=None if Le is None else [float(x) for x in Le], Le_scale=Le_scale)))
        self.n_media += 1
        return self.n_media - 1

    def add_rgbgrid_medium(self, sigma_a_rgb, sigma_s_rgb, lo, hi, scale_=1.0, g=0.0, world_from_medium=None):
        """rgbgrid (media.cpp:411-484): two [nz][ny][nx][3] float32 grids, sigma_scale applied at lookup."""
        sa = np.ascontiguousarray(sigma_a_rgb, dtype=np.float32)
        ss = np.ascontiguousarray(sigma_s_rgb, dtype=np.float32)
        self._keep += [sa, ss]
        self.log.append(("medium", dict(type="rgbgrid", sigma_a_rgb=sa, sigma_s_rgb=ss, lo=[float(x) for x in lo], hi=[float(x) for x in hi],
                                        scale=float(scale_), g=float(g), world_from_medium=world_from_medium, index=self.n_media)))
        self.n_media += 1
        return self.n_media - 1

    def add_nanovdb_medium(self, filename, sigma_a, sigma_s, scale_=1.0, g=0.0, gridname="density", density_offset=0.0, majorant_scale=1.0,
                           world_from_medium=None, written_name=None):
        """nanovdb (media.cpp:683-734): float grid `gridname` of a .nvdb file; the front end reads and flattens it.
        `written_name` is the file name a written .pbrt scene refers to (default: `filename`)."""
        import os
        self.log.append(("medium", dict(type="nanovdb", filename=os.path.abspath(filename), written_name=written_name, gridname=gridname,
                                        sigma_a=[float(x) for x in sigma_a], sigma_s=[float(x) for x in sigma_s], scale=float(scale_), g=float(g),
                                        density_offset=float(density_offset), majorant_scale=float(majorant_scale),
                                        world_from_medium=world_from_medium, index=self.n_media)))
        self.n_media += 1
        return self.n_media - 1

    def finalize(self):
        """Closes the scene description.  The POD desc is built on first use (`.desc`): writing the scene as files for
        the reference renderer needs only the world-space log, not the CUDA library."""
        return self

    @property
    def desc(self):
        if self._desc is None:
            self._build_desc()
        return self._desc

    def _build_desc(self):
        """Parses the scene through the C++ front end (vspg_host_scene_parse_string).  Grids travel by pointer, not
        as text: the scene text carries 1x1x1 placeholders and the parsed vspg_medium records are then pointed at the
        numpy arrays (the front end only stores grids; majorants / cell records are built by vspg_set_scene)."""
        from . import api, pbrt_writer
        self._hs = api.HostScene(text=pbrt_writer.scene_text(_ParseShim(self), placeholder_grids=True))
        d = self._hs.desc
        assert (d.n_quads, d.n_triangles, d.n_spheres, d.n_materials, d.n_lights, d.n_media) == \
               (self.n_quads, self.n_triangles, self.n_spheres, self.n_materials, self.n_lights, self.n_media)
        for kind, e in self.log:
            if kind != "medium" or e["type"] in ("homogeneous", "nanovdb"):   # nanovdb grids are read by the front end itself
                continue
            m = d.media[e["index"]]
            if e["type"] == "uniformgrid":
                m.nz, m.ny, m.nx = e["density"].shape
                m.density = e["density"].ctypes.data_as(A.PF)
                if e.get("Le_scale") is not None:
                    m.le_nz, m.le_ny, m.le_nx = e["Le_scale"].shape
                    m.le_scale = e["Le_scale"].ctypes.data_as(A.PF)
            else:
                m.nz, m.ny, m.nx = e["sigma_a_rgb"].shape[:3]
                m.sigma_a_rgb = e["sigma_a_rgb"].ctypes.data_as(A.PF)
                m.sigma_s_rgb = e["sigma_s_rgb"].ctypes.data_as(A.PF)
        self._desc = d

    @property
    def n_pixels(self):
        return self.camera.width * self.camera.height


class _ParseShim:
    """What pbrt_writer.scene_text needs of a Config, for the finalize() round trip: an unguided integrator block (the
    real vspg_integrator_params are set by the caller through vspg_set_integrator, not taken from this parse)."""

    def __init__(self, scene):
        from . import api
        self.scene, self.spp = scene, 1
        self.ip = api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0)
