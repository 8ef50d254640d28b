// NanoVDB float-grid reader (see host_nvdb.h).  Written against the documented byte layout of NanoVDB 32.x; it does not
// include or link the NanoVDB headers.
#include "host_nvdb.h"

#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstring>

namespace vspg_host {
namespace {

constexpr uint64_t kMagic = 0x304244566f6e614eULL;   // "NanoVDB0" (NanoVDB.h:102)
constexpr uint32_t kMajor = 32;                      // NanoVDB.h:104: the file format follows the ABI's major version
enum { CODEC_NONE = 0, CODEC_ZIP = 1, CODEC_BLOSC = 2 };
constexpr uint32_t kGridTypeFloat = 1;               // NanoVDB.h:216-217

// util/IO.h:109-165
#pragma pack(push, 1)
struct SegHeader { uint64_t magic; uint32_t version; uint16_t gridCount; uint16_t codec; };
struct MetaData {
    uint64_t gridSize, fileSize, nameKey, voxelCount;
    uint32_t gridType, gridClass;
    double worldBBox[6];
    int32_t indexBBox[6];
    double voxelSize[3];
    uint32_t nameSize;
    uint32_t nodeCount[4], tileCount[3];
    uint16_t codec, padding;
    uint32_t version;
};
#pragma pack(pop)
static_assert(sizeof(SegHeader) == 16, "segment header is 16 bytes");
static_assert(sizeof(MetaData) == 176, "grid meta data is 176 bytes");

// PNanoVDB.h:702-777, 904-916, 964-996, 1022-1111 and the Float row of pnanovdb_grid_type_constants (:1162)
constexpr size_t GRID_SIZE = 672, GRID_OFF_VERSION = 16, GRID_OFF_NAME = 40, GRID_OFF_MAP = 296, GRID_OFF_WORLD_BBOX = 560, GRID_OFF_TYPE = 636;
constexpr size_t MAP_OFF_MATF = 0, MAP_OFF_INVMATF = 36, MAP_OFF_VECF = 72, MAP_OFF_MATD = 88, MAP_OFF_INVMATD = 160, MAP_OFF_VECD = 232;
constexpr size_t TREE_OFF_ROOT = 24;
constexpr size_t ROOT_OFF_BBOX_MIN = 0, ROOT_OFF_BBOX_MAX = 12, ROOT_OFF_TABLE_SIZE = 24, ROOT_OFF_BACKGROUND = 28, ROOT_SIZE = 64;
constexpr size_t ROOT_TILE_OFF_KEY = 0, ROOT_TILE_OFF_CHILD = 8, ROOT_TILE_OFF_VALUE = 20, ROOT_TILE_SIZE = 32;
constexpr size_t UPPER_OFF_CHILD_MASK = 4128, UPPER_OFF_TABLE = 8256, UPPER_SIZE = 270400;
constexpr size_t LOWER_OFF_CHILD_MASK = 544, LOWER_OFF_TABLE = 1088, LOWER_SIZE = 33856;
constexpr size_t LEAF_OFF_TABLE = 96, LEAF_SIZE = 2144;

template <typename T> T rd(const std::vector<uint8_t> &b, size_t off) { T v; memcpy(&v, b.data() + off, sizeof(T)); return v; }
bool mask_bit(const std::vector<uint8_t> &b, size_t maskOff, uint32_t n) { return (rd<uint32_t>(b, maskOff + 4 * (n >> 5)) >> (n & 31)) & 1u; }

struct Dense {
    NvdbGrid *g;
    bool inside(int i, int j, int k) const {
        return i >= g->bbox_min[0] && j >= g->bbox_min[1] && k >= g->bbox_min[2] && i <= g->bbox_max[0] && j <= g->bbox_max[1] && k <= g->bbox_max[2];
    }
    void set(int i, int j, int k, float v) {
        if (inside(i, j, k)) g->dense[((size_t)(k - g->bbox_min[2]) * g->ny + (j - g->bbox_min[1])) * g->nx + (i - g->bbox_min[0])] = v;
    }
    // constant value over the cube [o, o + dim) (a tile), clipped to the dense box
    void fill(const int o[3], int dim, float v) {
        int lo[3], hi[3];
        for (int a = 0; a < 3; a++) { lo[a] = std::max(o[a], g->bbox_min[a]); hi[a] = std::min(o[a] + dim - 1, g->bbox_max[a]); if (lo[a] > hi[a]) return; }
        for (int k = lo[2]; k <= hi[2]; k++)
            for (int j = lo[1]; j <= hi[1]; j++)
                for (int i = lo[0]; i <= hi[0]; i++) set(i, j, k, v);
    }
};

bool flatten(const std::vector<uint8_t> &buf, NvdbGrid *g, std::string *err) {
    if (buf.size() < GRID_SIZE + 64 + ROOT_SIZE) { *err = "grid buffer too small"; return false; }
    if (rd<uint64_t>(buf, 0) != kMagic) { *err = "grid buffer does not start with the NanoVDB magic number"; return false; }
    if (rd<uint32_t>(buf, GRID_OFF_TYPE) != kGridTypeFloat) { *err = "grid is not a float grid"; return false; }
    const size_t map = GRID_OFF_MAP;
    for (int i = 0; i < 9; i++) {
        g->mat_f[i] = rd<float>(buf, map + MAP_OFF_MATF + 4 * i); g->inv_mat_f[i] = rd<float>(buf, map + MAP_OFF_INVMATF + 4 * i);
        g->mat_d[i] = rd<double>(buf, map + MAP_OFF_MATD + 8 * i); g->inv_mat_d[i] = rd<double>(buf, map + MAP_OFF_INVMATD + 8 * i);
    }
    for (int i = 0; i < 3; i++) {
        g->vec_f[i] = rd<float>(buf, map + MAP_OFF_VECF + 4 * i); g->vec_d[i] = rd<double>(buf, map + MAP_OFF_VECD + 8 * i);
        g->world_min[i] = rd<double>(buf, GRID_OFF_WORLD_BBOX + 8 * i); g->world_max[i] = rd<double>(buf, GRID_OFF_WORLD_BBOX + 24 + 8 * i);
    }
    const size_t tree = GRID_SIZE;
    const size_t root = tree + (size_t)rd<uint64_t>(buf, tree + TREE_OFF_ROOT);
    if (root + ROOT_SIZE > buf.size()) { *err = "root node outside the grid buffer"; return false; }
    for (int i = 0; i < 3; i++) { g->bbox_min[i] = rd<int32_t>(buf, root + ROOT_OFF_BBOX_MIN + 4 * i); g->bbox_max[i] = rd<int32_t>(buf, root + ROOT_OFF_BBOX_MAX + 4 * i); }
    g->background = rd<float>(buf, root + ROOT_OFF_BACKGROUND);
    const uint32_t tiles = rd<uint32_t>(buf, root + ROOT_OFF_TABLE_SIZE);
    for (int i = 0; i < 3; i++)
        if (g->bbox_max[i] < g->bbox_min[i]) { g->nx = g->ny = g->nz = 0; g->dense.clear(); return true; }   // empty grid
    const long long ex = (long long)g->bbox_max[0] - g->bbox_min[0] + 1, ey = (long long)g->bbox_max[1] - g->bbox_min[1] + 1, ez = (long long)g->bbox_max[2] - g->bbox_min[2] + 1;
    if (ex * ey * ez > (1ll << 33)) { *err = "index bounding box too large to flatten"; return false; }
    g->nx = (int)ex; g->ny = (int)ey; g->nz = (int)ez;
    g->dense.assign((size_t)(ex * ey * ez), g->background);
    Dense D{g};
    for (uint32_t t = 0; t < tiles; t++) {
        const size_t tile = root + ROOT_SIZE + (size_t)t * ROOT_TILE_SIZE;
        if (tile + ROOT_TILE_SIZE > buf.size()) { *err = "root tile outside the grid buffer"; return false; }
        const uint64_t key = rd<uint64_t>(buf, tile + ROOT_TILE_OFF_KEY);
        // PNanoVDB.h:1236-1242 coord_to_key: 21 bits per axis of (coord >> 12), x in the top bits; sign-extend back
        auto unkey = [](uint64_t v) { int32_t c = (int32_t)(uint32_t)(v & 0x1FFFFFu); if (c & 0x100000) c |= ~0x1FFFFF; return (int)((uint32_t)c << 12); };
        const int uo[3] = {unkey(key >> 42), unkey(key >> 21), unkey(key)};
        const int64_t child = rd<int64_t>(buf, tile + ROOT_TILE_OFF_CHILD);
        if (child == 0) { D.fill(uo, 4096, rd<float>(buf, tile + ROOT_TILE_OFF_VALUE)); continue; }
        const size_t upper = root + (size_t)child;
        if (upper + UPPER_SIZE > buf.size()) { *err = "upper node outside the grid buffer"; return false; }
        for (uint32_t n = 0; n < 32768; n++) {
            const int lo_[3] = {uo[0] + (int)((n >> 10) << 7), uo[1] + (int)(((n >> 5) & 31) << 7), uo[2] + (int)((n & 31) << 7)};
            const size_t ue = upper + UPPER_OFF_TABLE + 8 * (size_t)n;
            if (!mask_bit(buf, upper + UPPER_OFF_CHILD_MASK, n)) { D.fill(lo_, 128, rd<float>(buf, ue)); continue; }
            const size_t lower = upper + (size_t)rd<int64_t>(buf, ue);
            if (lower + LOWER_SIZE > buf.size()) { *err = "lower node outside the grid buffer"; return false; }
            for (uint32_t m = 0; m < 4096; m++) {
                const int fo[3] = {lo_[0] + (int)((m >> 8) << 3), lo_[1] + (int)(((m >> 4) & 15) << 3), lo_[2] + (int)((m & 15) << 3)};
                const size_t le = lower + LOWER_OFF_TABLE + 8 * (size_t)m;
                if (!mask_bit(buf, lower + LOWER_OFF_CHILD_MASK, m)) { D.fill(fo, 8, rd<float>(buf, le)); continue; }
                const size_t leaf = lower + (size_t)rd<int64_t>(buf, le);
                if (leaf + LEAF_SIZE > buf.size()) { *err = "leaf node outside the grid buffer"; return false; }
                for (uint32_t v = 0; v < 512; v++)   // PNanoVDB.h:1271-1276: (x & 7) << 6 | (y & 7) << 3 | (z & 7)
                    D.set(fo[0] + (int)(v >> 6), fo[1] + (int)((v >> 3) & 7), fo[2] + (int)(v & 7), rd<float>(buf, leaf + LEAF_OFF_TABLE + 4 * (size_t)v));
            }
        }
    }
    return true;
}

// Walks the segments of a file; calls visit(meta, name, offset of the grid's bytes) for every grid until it returns true.
template <typename F> bool walk(FILE *f, std::string *err, F visit) {
    while (true) {
        SegHeader h;
        size_t got = fread(&h, 1, sizeof h, f);
        if (got == 0) return false;                       // clean end of file
        if (got != sizeof h || h.magic != kMagic) { *err = "not a NanoVDB file (magic number)"; return false; }
        if (((h.version >> 21) & 0x7ffu) != kMajor) { *err = "NanoVDB file of major version " + std::to_string((h.version >> 21) & 0x7ffu) + " (this reader: 32)"; return false; }
        std::vector<MetaData> metas(h.gridCount);
        std::vector<std::string> names(h.gridCount);
        for (int i = 0; i < h.gridCount; i++) {
            if (fread(&metas[i], 1, sizeof(MetaData), f) != sizeof(MetaData)) { *err = "truncated grid meta data"; return false; }
            std::vector<char> nm(metas[i].nameSize + 1, 0);
            if (metas[i].nameSize && fread(nm.data(), 1, metas[i].nameSize, f) != metas[i].nameSize) { *err = "truncated grid name"; return false; }
            names[i] = nm.data();
        }
        for (int i = 0; i < h.gridCount; i++) {
            long at = ftell(f);
            if (visit(metas[i], names[i], h.codec)) return true;
            if (!err->empty()) return false;
            if (fseek(f, at + (long)metas[i].fileSize, SEEK_SET) != 0) { *err = "truncated grid data"; return false; }
        }
    }
}

}  // namespace

bool read_nvdb_float_grid(const std::string &path, const std::string &gridName, NvdbGrid *out, std::string *err) {
    err->clear();
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { *err = "cannot open " + path; return false; }
    bool found = walk(f, err, [&](const MetaData &m, const std::string &name, int codec) {
        if (name != gridName) return false;
        if (m.gridType != kGridTypeFloat) { *err = "grid \"" + gridName + "\" is not a float grid"; return false; }
        std::vector<uint8_t> buf((size_t)m.gridSize);
        if (codec == CODEC_NONE) {
            if (fread(buf.data(), 1, buf.size(), f) != buf.size()) { *err = "truncated grid data"; return false; }
        } else if (codec == CODEC_ZIP) {   // util/IO.h:314-330: u64 compressed size, then one zlib stream
            uint64_t csize = 0;
            if (fread(&csize, 1, 8, f) != 8) { *err = "truncated grid data"; return false; }
            std::vector<uint8_t> tmp((size_t)csize);
            if (fread(tmp.data(), 1, tmp.size(), f) != tmp.size()) { *err = "truncated grid data"; return false; }
            uLongf n = (uLongf)buf.size();
            if (uncompress(buf.data(), &n, tmp.data(), (uLong)tmp.size()) != Z_OK || n != buf.size()) { *err = "ZIP codec: inflate failed"; return false; }
        } else { *err = "BLOSC-compressed NanoVDB files are not supported (re-save with codec NONE or ZIP)"; return false; }
        out->name = name;
        return flatten(buf, out, err);
    });
    fclose(f);
    if (!found && err->empty()) *err = "no grid named \"" + gridName + "\" in " + path;
    return found;
}

bool nvdb_has_grid(const std::string &path, const std::string &gridName) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    std::string err;
    bool found = walk(f, &err, [&](const MetaData &, const std::string &name, int) { return name == gridName; });
    fclose(f);
    return found;
}

}  // namespace vspg_host
