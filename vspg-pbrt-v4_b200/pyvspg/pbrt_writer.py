"""Write a pyvspg Config as a .pbrt scene file in the reference's scene-description language, so the same workload
can be fed to the C++ front end (vspg_host_scene_parse / the vspg-pbrt command) or to the reference's own `pbrt`
binary.  Everything is emitted in WORLD space with a LookAt camera; numbers are printed with round-trip precision so
the front end reproduces the harness-built POD descs bit for bit.
"""
import os

import numpy as np

from . import _abi as A


def _f(x):
    return repr(float(x))


def _v(xs):
    return " ".join(_f(x) for x in xs)


def _b(x):
    return "true" if x else "false"


_LS = {A.LIGHTSAMPLER_UNIFORM: "uniform", A.LIGHTSAMPLER_POWER: "power", A.LIGHTSAMPLER_BVH: "bvh"}


def integrator_block(ip, name="guidedvolpathvspg", cache_file=None, isg_file=None, tr_file=None, store_isg_file=None):
    """guidedvolpathvspgintegrator.cpp:1260-1322 parameter names."""
    p = [f'Integrator "{name}"',
         f'  "integer maxdepth" [{ip.max_depth}] "integer minrrdepth" [{ip.min_rr_depth}] "bool usenee" {_b(ip.use_nee)}',
         f'  "bool surfaceguiding" {_b(ip.surface_guiding)} "bool volumeguiding" {_b(ip.volume_guiding)}',
         f'  "string surfaceguidingtype" "{"ris" if ip.surface_guiding_ris else "mis"}" "string volumeguidingtype" "{"ris" if ip.volume_guiding_ris else "mis"}"',
         f'  "bool vspguiding" {_b(ip.vsp_guiding)} "bool vspprimaryguiding" {_b(ip.vsp_primary_guiding)} "bool vspsecondaryguiding" {_b(ip.vsp_secondary_guiding)}',
         f'  "float vspmisratio" [{_f(ip.vsp_mis_ratio)}] "string vspcriterion" "{"variance" if ip.vsp_criterion_variance else "contribution"}"',
         f'  "string vspsamplingmethod" "{"nds" if ip.vsp_sampling_nds else "resampling"}" "bool collisionProbabilityBias" {_b(ip.collision_probability_bias)}',
         f'  "string lightsampler" "{_LS[ip.light_sampler]}" "bool regularize" {_b(ip.regularize)}']
    if cache_file:
        p.append(f'  "bool loadGuidingCache" true "string guidingCacheFileName" "{cache_file}"')
    if isg_file:
        p.append(f'  "bool loadISGBuffer" true "string isgBufferFileName" "{isg_file}"')
    if tr_file:
        p.append(f'  "bool loadTrBuffer" true "string trBufferFileName" "{tr_file}"')
    if store_isg_file:
        p.append(f'  "bool storeISGBuffer" true "string isgBufferFileName" "{store_isg_file}"')
    return "\n".join(p)


def scene_text(cfg, out_dir=None, outfile="out.exr", seed=0, integrator_name="guidedvolpathvspg", cache_file=None, isg_file=None,
               tr_file=None, raw_density=False, placeholder_grids=False, store_isg_file=None):
    """Returns the .pbrt text of `cfg` (a pyvspg.configs.Config).  With raw_density=True grids are written next to
    the scene as raw little-endian float32 files ("string densityfile", the front end's extension for grids too large
    for text) -- needs out_dir."""
    sc = cfg.scene
    cp = sc.camera._params
    fc = sc._desc.camera if getattr(sc, "_desc", None) is not None else sc.camera   # the filter may have been changed on the finalised desc
    L = [f'LookAt {_v(cp["eye"])}  {_v(cp["look"])}  {_v(cp["up"])}',
         f'Camera "perspective" "float fov" [{_f(cp["fov"])}] "float lensradius" [{_f(cp["lens_radius"])}] "float focaldistance" [{_f(cp["focal_distance"])}]',
         f'Film "rgb" "integer xresolution" [{sc.camera.width}] "integer yresolution" [{sc.camera.height}] "string filename" "{outfile}"',
         f'Sampler "independent" "integer pixelsamples" [{cfg.spp}] "integer seed" [{seed}]',
         ('PixelFilter "box"' if fc.filter_type == 0 else
          f'PixelFilter "gaussian" "float xradius" [{_f(fc.filter_radius[0])}] "float yradius" [{_f(fc.filter_radius[1])}] '
          f'"float sigma" [{_f(fc.filter_sigma)}]'),
         integrator_block(cfg.ip, integrator_name, cache_file, isg_file, tr_file, store_isg_file),
         'WorldBegin']
    for kind, e in sc.log:
        if kind == "medium":
            head = f'MakeNamedMedium "medium{e["index"]}" "string type" "{e["type"]}" "float scale" [{_f(e["scale"])}] "float g" [{_f(e["g"])}]'
            pre, post = [], []
            if e.get("world_from_medium") is not None:
                m = np.asarray(e["world_from_medium"], dtype=np.float64)
                pre.append("AttributeBegin\nTransform [" + _v(m.T.reshape(-1)) + "]")   # column-major, as pbrt files give matrices
                post.append("AttributeEnd")
            L += pre
            if e.get("Le") is not None:
                head += f' "rgb Le" [{_v(e["Le"])}]'
                if e["type"] == "homogeneous":
                    head += f' "float Lescale" [{_f(e["Le_scale"])}]'
                elif e.get("Le_scale") is not None:
                    ls = e["Le_scale"][:1, :1, :1] if placeholder_grids else e["Le_scale"]
                    head += ' "float Lescale" [' + " ".join(repr(float(x)) for x in ls.reshape(-1)) + "]"
            if e["type"] == "homogeneous":
                L.append(head + f' "rgb sigma_a" [{_v(e["sigma_a"])}] "rgb sigma_s" [{_v(e["sigma_s"])}]')
            elif e["type"] == "uniformgrid":
                d = e["density"][:1, :1, :1] if placeholder_grids else e["density"]
                nz, ny, nx = d.shape
                head += (f' "rgb sigma_a" [{_v(e["sigma_a"])}] "rgb sigma_s" [{_v(e["sigma_s"])}] "integer nx" [{nx}] "integer ny" [{ny}] "integer nz" [{nz}]'
                         f' "point3 p0" [{_v(e["lo"])}] "point3 p1" [{_v(e["hi"])}]')
                if raw_density:
                    fn = f'medium{e["index"]}_density.f32'
                    d.astype("<f4").tofile(os.path.join(out_dir, fn))
                    L.append(head + f' "string densityfile" "{fn}"')
                else:
                    L.append(head + ' "float density" [' + " ".join(repr(float(x)) for x in d.reshape(-1)) + "]")
            elif e["type"] == "nanovdb":
                fn = e["filename"] if placeholder_grids else (e.get("written_name") or e["filename"])
                L.append(head + f' "rgb sigma_a" [{_v(e["sigma_a"])}] "rgb sigma_s" [{_v(e["sigma_s"])}] "string filename" "{fn}" "string gridname" "{e["gridname"]}"'
                         f' "float densityoffset" [{_f(e["density_offset"])}] "float majorantscale" [{_f(e["majorant_scale"])}]')
            else:
                sa, ss = e["sigma_a_rgb"], e["sigma_s_rgb"]
                if placeholder_grids:
                    sa, ss = sa[:1, :1, :1], ss[:1, :1, :1]
                nz, ny, nx = sa.shape[:3]
                L.append(head + f' "integer nx" [{nx}] "integer ny" [{ny}] "integer nz" [{nz}] "point3 p0" [{_v(e["lo"])}] "point3 p1" [{_v(e["hi"])}]'
                         + ' "rgb sigma_a" [' + " ".join(repr(float(x)) for x in sa.reshape(-1)) + ']'
                         + ' "rgb sigma_s" [' + " ".join(repr(float(x)) for x in ss.reshape(-1)) + ']')
            L += post
        elif kind == "material":
            L.append(f'MakeNamedMaterial "mat{e["index"]}" "string type" "diffuse" "rgb reflectance" [{_v(e["reflectance"])}]')
        elif kind == "quad":
            L.append("AttributeBegin")
            if e["has_interface"]:
                name = lambda i: f"medium{i}" if i >= 0 else ""
                L.append(f'  MediumInterface "{name(e["medium_inside"])}" "{name(e["medium_outside"])}"')
            L.append(f'  NamedMaterial "mat{e["material"]}"' if e["material"] >= 0 else '  Material "interface"')
            if e["emission"] is not None:
                L.append(f'  AreaLightSource "diffuse" "rgb L" [{_v(e["emission"])}] "bool twosided" {_b(e["two_sided"])}')
            L.append('  Shape "bilinearmesh" "point3 P" [' + "  ".join(_v(p) for p in e["corners"]) + "]")
            L.append("AttributeEnd")
        elif kind == "trimesh":
            L.append("AttributeBegin")
            if e["has_interface"]:
                name = lambda i: f"medium{i}" if i >= 0 else ""
                L.append(f'  MediumInterface "{name(e["medium_inside"])}" "{name(e["medium_outside"])}"')
            L.append(f'  NamedMaterial "mat{e["material"]}"' if e["material"] >= 0 else '  Material "interface"')
            if e.get("emission") is not None:
                L.append(f'  AreaLightSource "diffuse" "rgb L" [{_v(e["emission"])}] "bool twosided" {_b(e["two_sided"])}')
            L.append('  Shape "trianglemesh" "point3 P" [' + "  ".join(_v(p) for p in e["points"]) + '] "integer indices" ['
                     + "  ".join(" ".join(str(i) for i in t) for t in e["indices"]) + "]")
            L.append("AttributeEnd")
        elif kind == "plymesh":
            L.append("AttributeBegin")
            if e.get("translate") is not None:
                L.append(f'  Translate {_v(e["translate"])}')
            L.append(f'  NamedMaterial "mat{e["material"]}"' if e["material"] >= 0 else '  Material "interface"')
            if e.get("emission") is not None:
                L.append(f'  AreaLightSource "diffuse" "rgb L" [{_v(e["emission"])}] "bool twosided" {_b(e["two_sided"])}')
            fn = e["filename"] if placeholder_grids else (e.get("written_name") or e["filename"])
            L.append(f'  Shape "plymesh" "string filename" "{fn}"')
            L.append("AttributeEnd")
        elif kind == "sphere":
            L.append("AttributeBegin")
            if e["has_interface"]:
                name = lambda i: f"medium{i}" if i >= 0 else ""
                L.append(f'  MediumInterface "{name(e["medium_inside"])}" "{name(e["medium_outside"])}"')
            L.append(f'  NamedMaterial "mat{e["material"]}"' if e["material"] >= 0 else '  Material "interface"')
            if e.get("emission") is not None:
                L.append(f'  AreaLightSource "diffuse" "rgb L" [{_v(e["emission"])}] "bool twosided" {_b(e["two_sided"])}')
            L.append(f'  Translate {_v(e["center"])}')
            L.append(f'  Shape "sphere" "float radius" [{_f(e["radius"])}]')
            L.append("AttributeEnd")
        elif kind == "point":
            L.append(f'LightSource "point" "rgb I" [{_v(e["I"])}] "point3 from" [{_v(e["pos"])}]')
        elif kind == "spot":
            L.append(f'LightSource "spot" "rgb I" [{_v(e["I"])}] "point3 from" [{_v(e["pos"])}] "point3 to" [{_v(e["to"])}] '
                     f'"float coneangle" [{_f(e["coneangle"])}] "float conedeltaangle" [{_f(e["conedeltaangle"])}]')
        elif kind == "distant":
            L.append(f'LightSource "distant" "rgb L" [{_v(e["L"])}] "point3 from" [{_v(e["to_light"])}] "point3 to" [0 0 0]')
        elif kind == "infinite":
            L.append(f'LightSource "infinite" "rgb L" [{_v(e["L"])}]')
    return "\n".join(L) + "\n"


def write_scene(cfg, path, **kw):
    out_dir = os.path.dirname(os.path.abspath(path))
    os.makedirs(out_dir, exist_ok=True)
    with open(path, "w") as f:
        f.write(scene_text(cfg, out_dir=out_dir, **kw))
    return path
