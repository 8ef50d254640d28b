"""Writes a vspg_field pair as an OpenPGL 0.7.0 (VSP branch) field file -- the format the reference's loadGuidingCache
path reads (pglFieldLoadFromFile -> SurfaceVolumeField::deserialize -> Field::deserialize, field/Field.h:308-349) and
csrc/host_field.cpp documents member by member.  Used to hand the synthetic caches of the BASELINE configs to the
reference renderer and to the C++ front end as FILES, so both consume the same bytes.

The members a render pass never reads (factory / builder configuration, timers) are taken from `template`, a field file
the reference itself wrote (tests/golden/field_small.pgl); per-region training statistics are zero."""
import ctypes as C
import struct

import numpy as np

from . import _abi as A

_HEADER = b"OPENPGL_0.7.0_FIELD\0"
_FIELD_FIXED = 1 + 4 + 8 + 8 + 1 + 1 + 1 + 1 + 24 + 1 + 16      # isSurface ... timers
_CFG = 77 + 24                                                   # factory configuration + builder settings
_TRAINING_STATS = 2349


def _field_bytes(f, points, is_surface, cfg_bytes, iteration):
    out = bytearray()
    bmin, bmax = list(f.bounds_min), list(f.bounds_max)
    out += struct.pack("<BfQQBBBB", 1 if is_surface else 0, 0.25, iteration, iteration, 0, 1, 1, 1)
    out += struct.pack("<6f", *bmin, *bmax)
    out += struct.pack("<B4f", 1, 0, 0, 0, 0)
    out += cfg_bytes
    out += struct.pack("<B6fQ", 1, *bmin, *bmax, f.n_nodes)
    out += bytes(C.string_at(f.nodes, f.n_nodes * C.sizeof(A.KdNode)))
    out += struct.pack("<Q", f.n_regions)
    regs = np.frombuffer(C.string_at(f.regions, f.n_regions * C.sizeof(A.Mixture)), dtype=np.uint8).reshape(f.n_regions, C.sizeof(A.Mixture))
    pts = np.asarray(points, dtype=np.float32).reshape(f.n_regions, 3)
    zeros_stats = bytes(_TRAINING_STATS)
    for i in range(f.n_regions):
        m = f.regions[i]
        out += struct.pack("<B", 1 if m.valid else 0)
        out += regs[i, :10 * 128].tobytes()                      # w, kappa, mu x/y/z, N, e^-2k, meanCos, dist, vspW: 10 x float[32]
        out += struct.pack("<Q3f", m.num_components, *list(m.pivot))
        out += struct.pack("<6f", *bmin, *bmax)                  # regionBounds (training only)
        out += zeros_stats
        out += struct.pack("<3f3f2f6f", *pts[i], 0, 0, 0, 1.0, 0.0, *bmin, *bmax)   # SampleStatistics: mean = the KNN point
        out += struct.pack("<QB", 0, 0)                          # numZeroValueSamples, splitFlag
        out += struct.pack("<4Q", 0, 0, 0, 0)                    # Range
    out += struct.pack("<BBB", 1 if f.stochastic_nn else 0, 0, 1)   # useStochasticNNLookUp, useISNNLookUp, KNN tree built
    out += struct.pack("<I", f.n_regions)
    out += np.concatenate([pts, np.zeros((f.n_regions, 1), np.float32)], axis=1).astype("<f4").tobytes()
    return bytes(out)


def write_field(path, surface, surface_points, volume, volume_points, template, iteration=1):
    """surface / volume: A.Field structs (pyvspg.synthetic.synthetic_field); *_points: the regions' sample means (the
    points the reference's deserialize rebuilds the neighbour tables from)."""
    t = open(template, "rb").read()
    assert t[:len(_HEADER)] == _HEADER, "template is not an OpenPGL 0.7.0 field file"
    cfg = t[len(_HEADER) + 8 + 16 + _FIELD_FIXED:][:_CFG]
    with open(path, "wb") as fp:
        fp.write(_HEADER)
        fp.write(struct.pack("<IIQQ", 0, 0, iteration, iteration))
        fp.write(_field_bytes(surface, surface_points, True, cfg, iteration))
        fp.write(_field_bytes(volume, volume_points, False, cfg, iteration))
    return path
