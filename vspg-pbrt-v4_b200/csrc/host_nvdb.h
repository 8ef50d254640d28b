// Minimal reader for NanoVDB float grids (file format and in-memory layout of NanoVDB 32.x, the version the reference
// vendors under src/ext/openvdb/nanovdb: util/IO.h:109-165 segment header + per-grid meta data, PNanoVDB.h:702-1111
// byte offsets of GridData / TreeData / root / upper / lower / leaf nodes).  The sparse tree is flattened into a dense
// float array over the grid's index bounding box -- what the CUDA path turns into 32-byte cell records -- and the
// index<->world map is kept in both precisions, as media.cpp:586-670 uses both.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace vspg_host {

struct NvdbGrid {
    std::string name;
    int bbox_min[3], bbox_max[3];          // Tree::bbox(), inclusive (media.cpp:638 indexBBox())
    double world_min[3], world_max[3];     // GridData::mWorldBBox
    float mat_f[9], inv_mat_f[9], vec_f[3];   // Map (NanoVDB.h:1946-1960), single precision
    double mat_d[9], inv_mat_d[9], vec_d[3];  // double precision
    float background;
    // dense copy over [bbox_min, bbox_max] (inclusive), x fastest: value(i,j,k) = dense[((k-k0)*ny + (j-j0))*nx + (i-i0)];
    // voxels the tree does not cover hold the value ReadAccessor::getValue returns for them (tile value or background)
    int nx, ny, nz;
    std::vector<float> dense;
};

// Reads the first float grid called `gridName` (util/IO.h readGrid(fileName, gridName)).  Returns false with *err set
// when the file cannot be read, holds no such grid, or uses something this reader does not cover (BLOSC codec,
// non-float value types, another major version).
bool read_nvdb_float_grid(const std::string &path, const std::string &gridName, NvdbGrid *out, std::string *err);

// True when the file holds a grid of that name (any type): Create() looks for an optional "temperature" grid this way.
bool nvdb_has_grid(const std::string &path, const std::string &gridName);

}  // namespace vspg_host
