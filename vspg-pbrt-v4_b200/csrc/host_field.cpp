// Guiding-cache loader: reads an OpenPGL 0.7.0 (VSP branch) field file -- what the reference's loadGuidingCache
// path consumes through pglFieldLoadFromFile (cpu/guidedvolpathvspgintegrator.cpp:117-122 ->
// openpgl/api/api.cpp -> device/Device.h:236-290 -> field/SurfaceVolumeField.h:191-197) -- and flattens it into the
// POD arrays of include/vspg.h (KD nodes, one vspg_mixture per region, 8-NN neighbour tables).
//
// The file is a member-by-member dump of the reference's C++ objects (x86-64, little endian), built with
// OPENPGL_EF_VSP_GUIDING=ON, radiance caches OFF (src/ext/CMakeLists.txt:262-273), device CPU_4
// (ParallaxAwareVonMisesFisherMixture<4,32,true>).  Layout, with the serialize() that writes each piece:
//   "OPENPGL_0.7.0_FIELD\0"  u32 spatialStructureType  u32 directionalDistributionType      SurfaceVolumeField.h:209-224
//   u64 iteration  u64 totalSPP                                                              SurfaceVolumeField.h:183-189
//   Field surface, Field volume                                                              field/Field.h:273-306
// Field:
//   u8 isSurface  f32 decayOnSpatialSplit  u64 iteration  u64 totalSPP  u8 deterministic  u8 vspVarianceBased
//   u8 fitRegions  u8 isSceneBoundsSet  f32[6] sceneBounds  u8 initialized  f32[4] timers
//   factory configuration (77 B)    directional/vmm/AdaptiveSplitandMergeFactory.h:249-262,
//                                   ParallaxAwareVonMisesFisherWeightedEMFactory.h:1593-1608
//   builder settings (3 x u64)      spatial/kdtree/KDTreeBuilder.h:490-495
//   KD tree: u8 isInit  f32[6] bounds  u64 n  n x {f32 split, u32 dimAndIdx}                 spatial/kdtree/KDTree.h:143-147, 616-626
//   u64 nRegions, nRegions x { Region (3739 B)  spatial/Region.h:76-89 ; Range (4 x u64)  data/Range.h:45-52 }
//   u8 useStochasticNNLookUp  u8 useISNNLookUp
//   KNN tree: u8 isBuild [u32 n, n x 16 B points]                                            spatial/KNN.h:455-466
// Region:
//   u8 valid; mixture = f32[32] x {w, kappa, mu.x, mu.y, mu.z, N, e^-2k, meanCos, dist, vspW}, u64 numComponents,
//   f32[3] pivot (ParallaxAwareVonMisesFisherMixture.h:521-546); f32[6] regionBounds; training statistics (2349 B,
//   skipped); SampleStatistics {f32[3] mean, f32[3] variance, f32 n, f32 nZero, f32[6] bounds} (data/SampleStatistics.h:177-184);
//   u64 numZeroValueSamples; u8 splitFlag.
// Neighbour tables are not stored: Field::deserialize rebuilds them (Field.h:345-348 -> KNN.h:343-393, the 8 nearest
// region points, self included, ascending squared distance); done here the same way.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vspg.h"

namespace {

struct Reader {
    const unsigned char *p, *end;
    bool ok = true;
    template <typename T> T get() {
        T v{};
        if ((size_t)(end - p) < sizeof(T)) { ok = false; p = end; return v; }
        memcpy(&v, p, sizeof(T)); p += sizeof(T);
        return v;
    }
    void skip(size_t n) { if ((size_t)(end - p) < n) { ok = false; p = end; } else p += n; }
    void floats(float *dst, size_t n) {
        if ((size_t)(end - p) < 4 * n) { ok = false; p = end; return; }
        memcpy(dst, p, 4 * n); p += 4 * n;
    }
};

constexpr size_t kFactoryCfgBytes = (8 + 4 + 8 + 8 + 6 * 4) + 4 + 4 + 1 + 4 * 4;          // 77
constexpr size_t kSufficientStatsBytes = (3 + 1 + 1) * 128 + 3 * 4 + 8 + 1 + 4 * 128;     // 1173
constexpr size_t kSplitStatsBytes = (1 + 2 + 3 + 1 + 1 + 1) * 128 + 8;                    // 1160
constexpr size_t kTrainingStatsBytes = kSufficientStatsBytes + kSplitStatsBytes + 2 * 8;  // 2349

struct HostFieldPart {
    vspg_field desc{};
    std::vector<vspg_kdnode> nodes;
    std::vector<vspg_mixture> regions;
    std::vector<vspg_neighbours> neigh;
    std::vector<float> points;  // 3 per region (SampleStatistics::mean = the KNN point)
    bool initialized = false;
    unsigned long long iteration = 0;
};

bool read_field(Reader &r, HostFieldPart &f, std::string &err, const char *which) {
    (void)r.get<uint8_t>();                 // isSurface
    (void)r.get<float>();                   // decayOnSpatialSplit
    f.iteration = r.get<uint64_t>();
    (void)r.get<uint64_t>();                // totalSPP
    (void)r.get<uint8_t>();                 // deterministic
    (void)r.get<uint8_t>();                 // vspVarianceBased
    (void)r.get<uint8_t>();                 // fitRegions
    (void)r.get<uint8_t>();                 // isSceneBoundsSet
    r.skip(24);                             // sceneBounds
    f.initialized = r.get<uint8_t>() != 0;
    r.skip(16);                             // timers
    r.skip(kFactoryCfgBytes);
    r.skip(24);                             // builder settings
    (void)r.get<uint8_t>();                 // KDTree::m_isInit
    float b[6]; r.floats(b, 6);
    uint64_t nNodes = r.get<uint64_t>();
    if (!r.ok || nNodes > (1ull << 30)) { err = std::string(which) + " field: truncated or corrupt KD tree header"; return false; }
    f.nodes.resize(nNodes);
    for (uint64_t i = 0; i < nNodes; i++) { f.nodes[i].split = r.get<float>(); f.nodes[i].dim_idx = r.get<uint32_t>(); }
    uint64_t nRegions = r.get<uint64_t>();
    if (!r.ok || nRegions > (1ull << 28)) { err = std::string(which) + " field: truncated or corrupt region table"; return false; }
    f.regions.resize(nRegions);
    f.points.resize(3 * nRegions);
    for (uint64_t i = 0; i < nRegions; i++) {
        vspg_mixture &m = f.regions[i];
        memset(&m, 0, sizeof m);
        m.valid = r.get<uint8_t>() != 0;
        r.floats(m.weight, 32); r.floats(m.kappa, 32);
        r.floats(m.mu_x, 32); r.floats(m.mu_y, 32); r.floats(m.mu_z, 32);
        r.floats(m.norm, 32); r.floats(m.e_minus_2k, 32); r.floats(m.mean_cos, 32); r.floats(m.distance, 32);
        r.floats(m.vsp_weight, 32);
        uint64_t nc = r.get<uint64_t>();
        r.floats(m.pivot, 3);
        if (nc > VSPG_MAX_LOBES) { err = std::string(which) + " field: region with more than 32 lobes"; return false; }
        m.num_components = (int)nc;
        r.skip(24);                          // regionBounds
        r.skip(kTrainingStatsBytes);
        r.floats(&f.points[3 * i], 3);       // SampleStatistics::mean
        r.skip(12 + 4 + 4 + 24);             // variance, numSamples, numZeroValueSamples, sampleBounds
        r.skip(8 + 1);                       // numZeroValueSamples, splitFlag
        r.skip(32);                          // Range
        if (!r.ok) { err = std::string(which) + " field: truncated inside region " + std::to_string(i); return false; }
    }
    bool stochasticNN = r.get<uint8_t>() != 0, isNN = r.get<uint8_t>() != 0;
    bool knnBuilt = r.get<uint8_t>() != 0;
    if (knnBuilt) {
        uint32_t np = r.get<uint32_t>();
        if (!r.ok || np != nRegions) { err = std::string(which) + " field: KNN point count does not match the region count"; return false; }
        for (uint32_t i = 0; i < np; i++) { float q[4]; r.floats(q, 4); f.points[3 * i] = q[0]; f.points[3 * i + 1] = q[1]; f.points[3 * i + 2] = q[2]; }
    }
    if (!r.ok) { err = std::string(which) + " field: truncated"; return false; }
    // validate the tree against the region table (a corrupt index must not reach the GPU)
    for (uint64_t i = 0; i < nNodes; i++) {
        uint32_t v = f.nodes[i].dim_idx, dim = v >> 30, idx = (v << 2) >> 2;
        if (dim == 3 ? idx >= nRegions : (uint64_t)idx + 1 >= nNodes) { err = std::string(which) + " field: KD node points outside the tree"; return false; }
    }
    // Field.h:345-348: neighbours only when stochastic lookup is on and the KNN tree was built
    if (stochasticNN && knnBuilt && nRegions > 0) {
        f.neigh.resize(nRegions);
        const int n = (int)nRegions, k = std::min(8, n);
        std::vector<std::pair<float, int>> cand(n);
        for (int i = 0; i < n; i++) {
            const float *q = &f.points[3 * i];
            for (int j = 0; j < n; j++) {
                const float *p = &f.points[3 * j];
                float d0 = q[0] - p[0], d1 = q[1] - p[1], d2 = q[2] - p[2];
                float d = 0.f; d += d0 * d0; d += d1 * d1; d += d2 * d2;   // nanoflann L2_Simple_Adaptor::evalMetric
                cand[j] = {d, j};
            }
            std::partial_sort(cand.begin(), cand.begin() + k, cand.end());
            vspg_neighbours &nb = f.neigh[i];
            memset(&nb, 0, sizeof nb);
            for (int s = 0; s < 8; s++) {
                if (s < k) { int j = cand[s].second; nb.ids[s] = (uint32_t)j; nb.px[s] = f.points[3 * j]; nb.py[s] = f.points[3 * j + 1]; nb.pz[s] = f.points[3 * j + 2]; }
                else nb.ids[s] = ~0u;
            }
            nb.size = (uint32_t)k;
        }
    }
    vspg_field &d = f.desc;
    d.usable = f.iteration > 0 ? 1 : 0;      // Field::getRegion: `if (m_iteration > 0 && ...)`  (Field.h:116)
    for (int a = 0; a < 3; a++) { d.bounds_min[a] = b[a]; d.bounds_max[a] = b[3 + a]; }
    d.n_nodes = (int)nNodes; d.nodes = f.nodes.data();
    d.n_regions = (int)nRegions; d.regions = f.regions.data();
    d.neighbours = f.neigh.empty() ? nullptr : f.neigh.data();
    d.stochastic_nn = (stochasticNN && !f.neigh.empty()) ? 1 : 0;
    d.is_nn = isNN ? 1 : 0;
    return true;
}

}  // namespace

struct vspg_host_field {
    HostFieldPart surf, vol;
    unsigned long long iteration = 0;
};

static void set_err(char *err, size_t n, const std::string &s) { if (err && n) snprintf(err, n, "%s", s.c_str()); }

extern "C" {

int vspg_host_field_load(const char *pgl_path, vspg_host_field **out, char *err, size_t err_len) {
    if (!pgl_path || !out) return VSPG_ERR_INVALID;
    *out = nullptr;
    FILE *fp = fopen(pgl_path, "rb");
    if (!fp) { set_err(err, err_len, std::string("cannot open guiding cache file '") + pgl_path + "'"); return VSPG_ERR_IO; }
    std::vector<unsigned char> buf;
    fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
    if (sz < 0) { fclose(fp); return VSPG_ERR_IO; }
    buf.resize((size_t)sz);
    size_t got = fread(buf.data(), 1, buf.size(), fp);
    fclose(fp);
    if (got != buf.size()) { set_err(err, err_len, "short read"); return VSPG_ERR_IO; }
    static const char kHeader[] = "OPENPGL_0.7.0_FIELD";
    if (buf.size() < sizeof kHeader + 8 || memcmp(buf.data(), kHeader, sizeof kHeader) != 0) {
        // Device.h:243-262: the reference refuses files whose header string (library version) differs
        set_err(err, err_len, std::string("'") + pgl_path + "' is not an OpenPGL 0.7.0 field file (header mismatch)");
        return VSPG_ERR_IO;
    }
    Reader r{buf.data() + sizeof kHeader, buf.data() + buf.size()};
    uint32_t spatialType = r.get<uint32_t>(), dirType = r.get<uint32_t>();
    // PGL_SPATIAL_STRUCTURE_KDTREE = 0; PGL_DIRECTIONAL_DISTRIBUTION_PARALLAX_AWARE_VMM = 0 (include/openpgl/types.h)
    if (spatialType != 0 || dirType != 0) {
        set_err(err, err_len, "guiding cache uses a spatial structure / directional distribution other than KD-tree + parallax-aware VMM");
        return VSPG_ERR_UNSUPPORTED;
    }
    vspg_host_field *f = new vspg_host_field();
    f->iteration = r.get<uint64_t>();
    (void)r.get<uint64_t>();
    std::string e;
    if (!read_field(r, f->surf, e, "surface") || !read_field(r, f->vol, e, "volume")) { set_err(err, err_len, e); delete f; return VSPG_ERR_IO; }
    if (r.p != r.end) { set_err(err, err_len, "guiding cache file has trailing bytes (built with different OpenPGL feature flags?)"); delete f; return VSPG_ERR_IO; }
    *out = f;
    return VSPG_OK;
}

void vspg_host_field_free(vspg_host_field *f) { delete f; }
const vspg_field *vspg_host_field_surface(const vspg_host_field *f) { return f && f->surf.desc.n_regions > 0 ? &f->surf.desc : nullptr; }
const vspg_field *vspg_host_field_volume(const vspg_host_field *f) { return f && f->vol.desc.n_regions > 0 ? &f->vol.desc : nullptr; }

}  // extern "C"
