"""Writes a pyvspg Config as the set of FILES the reference's scene-description surface names: scene.pbrt, the guiding
cache (OpenPGL field file), the ISG buffer and the Tr buffer (EXR).  The reference renderer and the C++ front end then
consume the same bytes -- this is how bench.py feeds both arms and how the full-size parity tests reach the reference."""
import os

import numpy as np

from . import pbrt_writer, pgl_writer


def _exr_float(path, channels):
    """Uncompressed scanline EXR with FLOAT channels (pure Python: usable by a process that must not load libvspg)."""
    import struct
    names = sorted(channels)
    h, w = channels[names[0]].shape
    def attr(n, t, data):
        return n.encode() + b"\0" + t.encode() + b"\0" + struct.pack("<i", len(data)) + data
    ch = b"".join(n.encode() + b"\0" + struct.pack("<iBxxxii", 2, 0, 1, 1) for n in names) + b"\0"
    hdr = struct.pack("<II", 20000630, 2)
    hdr += attr("channels", "chlist", ch) + attr("compression", "compression", b"\0")
    hdr += attr("dataWindow", "box2i", struct.pack("<4i", 0, 0, w - 1, h - 1)) + attr("displayWindow", "box2i", struct.pack("<4i", 0, 0, w - 1, h - 1))
    hdr += attr("lineOrder", "lineOrder", b"\0") + attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
    hdr += attr("screenWindowCenter", "v2f", struct.pack("<2f", 0, 0)) + attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0"
    row = 4 * w * len(names)
    pos = len(hdr) + 8 * h
    table = b"".join(struct.pack("<Q", pos + y * (8 + row)) for y in range(h))
    planes = [np.ascontiguousarray(channels[n], dtype="<f4") for n in names]
    with open(path, "wb") as f:
        f.write(hdr + table)
        for y in range(h):
            f.write(struct.pack("<ii", y, row) + b"".join(p[y].tobytes() for p in planes))


def export_config(cfg, out_dir, template_field, name="scene", outfile="out.pfm", seed=0, spp=None, raw_density=False):
    """Returns the path of <out_dir>/<name>.pbrt.  File names inside the scene are relative to out_dir (the reference
    resolves the cache / buffer names against its working directory: run it with cwd = out_dir)."""
    os.makedirs(out_dir, exist_ok=True)
    g = cfg.guiding
    kw = {}
    if g is not None:
        if g.fields:
            pgl_writer.write_field(os.path.join(out_dir, name + "_cache.pgl"), g.fields["surface"], g.points["surface"], g.fields["volume"],
                                   g.points["volume"], template_field)
            kw["cache_file"] = name + "_cache.pgl"
        if getattr(g, "isg", None) is not None:
            z = np.zeros_like(g.isg)   # the reference's loader keys "vspEstimate" on a surfContrib layer being present
            _exr_float(os.path.join(out_dir, name + "_isg.exr"), {"VSP.Y": g.isg, "surfContrib.R": z, "surfContrib.G": z, "surfContrib.B": z})
            kw["isg_file"] = name + "_isg.exr"
        if getattr(g, "tr", None) is not None and cfg.ip.collision_probability_bias:
            _exr_float(os.path.join(out_dir, name + "_tr.exr"), {"Transmittance.R": g.tr[..., 0], "Transmittance.G": g.tr[..., 1], "Transmittance.B": g.tr[..., 2]})
            kw["tr_file"] = name + "_tr.exr"
    if spp is not None:
        import copy
        cfg = copy.copy(cfg)
        cfg.spp = spp
    return pbrt_writer.write_scene(cfg, os.path.join(out_dir, name + ".pbrt"), outfile=outfile, seed=seed, raw_density=raw_density, **kw)
