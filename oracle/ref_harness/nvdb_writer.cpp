// Test infrastructure (oracle/): writes .nvdb files with the reference's own vendored NanoVDB (header-only,
// /root/reference/src/ext/openvdb/nanovdb), so that the product's reader (csrc/host_nvdb.cpp) is exercised on files
// the reference's library produced.  Built by oracle/build_ref_pbrt.sh into oracle/_ref/nvdb_writer.
//   nvdb_writer <out.nvdb> <raw float32 grid file> nx ny nz  i0 j0 k0  voxelSize  ox oy oz  [zip]
// Voxel (i0+x, j0+y, k0+z) of a FogVolume grid named "density" gets raw[(z*ny + y)*nx + x] when that value is > 0
// (zeros stay background, so the tree is sparse); index -> world is  p = voxelSize * ijk + (ox,oy,oz).
#include <nanovdb/util/GridBuilder.h>
#include <nanovdb/util/IO.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

int main(int argc, char **argv) {
    if (argc < 13) { fprintf(stderr, "usage: nvdb_writer out.nvdb raw.f32 nx ny nz i0 j0 k0 voxelSize ox oy oz [zip]\n"); return 2; }
    const int nx = atoi(argv[3]), ny = atoi(argv[4]), nz = atoi(argv[5]);
    const int i0 = atoi(argv[6]), j0 = atoi(argv[7]), k0 = atoi(argv[8]);
    const double dx = atof(argv[9]);
    const nanovdb::Vec3d origin(atof(argv[10]), atof(argv[11]), atof(argv[12]));
    std::vector<float> raw((size_t)nx * ny * nz);
    FILE *f = fopen(argv[2], "rb");
    if (!f || fread(raw.data(), sizeof(float), raw.size(), f) != raw.size()) { fprintf(stderr, "nvdb_writer: cannot read %s\n", argv[2]); return 1; }
    fclose(f);
    try {
        nanovdb::GridBuilder<float> builder(0.0f, nanovdb::GridClass::FogVolume);
        auto acc = builder.getAccessor();
        for (int z = 0; z < nz; z++)
            for (int y = 0; y < ny; y++)
                for (int x = 0; x < nx; x++) {
                    float v = raw[((size_t)z * ny + y) * nx + x];
                    if (v > 0) acc.setValue(nanovdb::Coord(i0 + x, j0 + y, k0 + z), v);
                }
        auto handle = builder.getHandle<>(dx, origin, "density");
        const bool zip = argc > 13 && !strcmp(argv[13], "zip");
        nanovdb::io::writeGrid(argv[1], handle, zip ? nanovdb::io::Codec::ZIP : nanovdb::io::Codec::NONE);
        auto *g = handle.grid<float>();
        auto bb = g->indexBBox();
        printf("nvdb_writer: %s  index bbox [%d %d %d]..[%d %d %d]  %llu active voxels\n", argv[1], bb.min()[0], bb.min()[1], bb.min()[2],
               bb.max()[0], bb.max()[1], bb.max()[2], (unsigned long long)g->activeVoxelCount());
    } catch (const std::exception &e) { fprintf(stderr, "nvdb_writer: %s\n", e.what()); return 1; }
    return 0;
}
