// C-ABI implementation (include/vspg.h): context, HBM staging of the scene / guiding inputs, the wavefront
// render loop, film and counters, and the per-kernel parity hooks.  No CPU fallback: every compute entry point
// launches CUDA kernels and returns VSPG_ERR_CUDA if the device is unavailable.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <unordered_map>
#include <mutex>
#include <vector>

#include "host_util.h"
#include "host_lightbvh.h"
#include "host_lightpower.h"
#include "kernels2.cuh"
#include "kbin.cuh"
#include "kshadow_p.cuh"
#include "debug_kernels.cuh"

using namespace vspg;
using namespace vspg_host;

extern "C" const uint32_t vspg_hg_vmm_table_bits[128 * 18];  // hg_table_data.cpp (generated data)

struct vspg_ctx {
    int device = 0;
    int sm_count = 148;
    std::string err;
    std::vector<void *> scene_allocs, guide_allocs, wf_allocs;
    // Device blocks released by the previous set_scene / set_guiding, keyed by size: a caller that re-stages inputs of
    // the same shape every frame (bench.py's e2e leg, an animation) reuses them instead of paying cudaFree + cudaMalloc
    // (each a device-wide synchronisation, 25-30 ms for the 67 MB cell grid).  Blocks not reused by the end of the call
    // that could have reused them are returned to the driver (dev_trim).
    std::multimap<size_t, void *> block_cache;
    std::unordered_map<void *, size_t> block_size;
    DScene sc{};
    bool have_scene = false;
    Wavefront W{};
    Queue phys_q[2]{};            // the two ping-pong path queues (roles swap each wave iteration)
    int capacity = 0, requested_capacity = 0;
    int *qcounts = nullptr;       // device: 8 queue counters
    int *h_pinned = nullptr;      // pinned readback (int[0]) + RenderVars staging (from int[4])
    RenderVars *d_rv = nullptr;   // per-render variables the kernels read through Wavefront::rv
    int *d_iters = nullptr;       // trips of the wave graph's WHILE node in the current render
    bool use_graph = true;        // device-driven wave loop (VSPG_GRAPH=0: host-driven loop)
    cudaGraph_t wave_graph = nullptr; cudaGraphExec_t wave_exec = nullptr; struct WaveGraphKey *wave_key = nullptr;
    cudaStream_t cap_stream = nullptr;
    unsigned graph_builds = 0;
    const void *cells_ptr = nullptr; size_t cells_bytes = 0;      // cell-record grid of the (first) uniformgrid medium
    size_t l2_persist_bytes = 0, l2_window_max = 0;               // persisting-L2 carve-out granted by the device (0 = off)
    double *film = nullptr;
    float *isg_rec = nullptr, *isg_trained = nullptr; size_t isg_pixels = 0; bool isg_active = false;   // storeISGBuffer (vspg_isg_store_*)
    const float *isg_saved_vsp = nullptr; int isg_saved_w = 0, isg_saved_h = 0;
    double *tr_acc = nullptr; size_t tr_acc_pixels = 0;   // storeTrBuffer accumulators (vspg_tr_store_*)
    unsigned long long *counters = nullptr;
    unsigned long long launches = 0, iterations = 0;
    int film_w = 0, film_h = 0;
    int grid_layout = GRID_CELL8;
    int scatter_block = 256;   // VSPG_SCATTER_BLOCK overrides (A/B runs)
    int volscatter_ctas = 2, shade_ctas = 2, scatter_ctas_block = 0;   // resident CTAs/SM of the scatter kernels (occupancy API, per scatter_block)
    int shadow_persist = 0;     // VSPG_SHADOW_PERSIST=1: shadow rays through the persistent-lane kernel (kshadow_p.cuh); measured slower than one thread per ray (profiles/r02_summary.md)
    int shadow_refill = 16;
    int band_mode = -1;         // sample order of a render: 0 = pass-major, -1 = band-major with automatic band height, n > 0 = n rows (VSPG_BAND_ROWS)
    int sort_mode = 0;          // coherence pass (kbin.cuh): bit 0 = medium queue, bit 1 = shadow queue (VSPG_SORT)
    bool bin_ok = false;        // the scene has a grid medium to bin over
    int *bin_keys = nullptr, *bin_hist = nullptr, *bin_offsets = nullptr, *sorted_medium = nullptr, *sorted_shadow = nullptr;
    int prim_kinds = 7;         // primitive kinds of the scene (dshapes.cuh KINDS): 1 = quads only
    bool any_grid_medium = false;   // some medium is not homogeneous: k_medium at full occupancy (kernels.cuh)
    bool any_emissive = false;  // some medium emits: the delta-tracking kernel with the emission term compiled in
    bool any_nvdb = false;      // some medium is a NanoVDB grid: the resampling kernel with the NanoVDB sampler compiled in
    bool grids_ready = false;   // one-wave grids of the grid-stride kernels (occupancy API, vspg_ctx_render.inl)
    int g_generate = 0, g_intersect = 0, g_medium = 0, g_medium_hom = 0, g_lookup0 = 0, g_lookup1 = 0, g_surface = 0, g_volscatter0 = 0, g_shade0 = 0, g_finish = 0,
        g_bin = 148, g_shadow_gray = 0, g_shadow_rgb = 0;
    int rs_occ = 5;               // CTAs/SM the gray resampling kernel is compiled for (VSPG_RS_OCC=4|5, A/B runs)
    int rs_refill = kRsRefillIdleDefault;   // VSPG_RS_REFILL overrides (A/B runs)
    bool all_gray = false;        // every medium has sigma_a / sigma_s equal across RGB => scalar tracking kernels (kmedium_rs.cuh)
    // timing (vspg_kernel_stats_get)
    bool timing = false;
    cudaEvent_t ev[VSPG_K_COUNT + 1] = {};
    double ms_kernel[VSPG_K_COUNT] = {};
    unsigned long long launches_kernel[VSPG_K_COUNT] = {};
    double ms_total = 0;
};

static std::string g_create_error;

static int fail(vspg_ctx *c, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}
#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) return fail(ctx, VSPG_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

static cudaError_t dev_alloc(vspg_ctx *ctx, std::vector<void *> &owner, void **p, size_t bytes) {
    auto it = ctx->block_cache.find(bytes);
    if (it != ctx->block_cache.end()) { *p = it->second; ctx->block_cache.erase(it); owner.push_back(*p); return cudaSuccess; }
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) return e;
    ctx->block_size[*p] = bytes;
    owner.push_back(*p);
    return e;
}
#define CK_STATE(cond, msg) do { if (!(cond)) return fail(ctx, VSPG_ERR_STATE, msg); } while (0)

template <typename T>
static cudaError_t upload(vspg_ctx *ctx, std::vector<void *> &owner, const T *host, size_t n, const T **dev) {
    *dev = nullptr;
    if (n == 0) return cudaSuccess;
    void *p = nullptr;
    cudaError_t e = dev_alloc(ctx, owner, &p, n * sizeof(T));
    if (e != cudaSuccess) return e;
    e = cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice);
    *dev = (const T *)p;
    return e;
}
// release: park the blocks in the cache;  trim: give whatever is still parked back to the driver
static void dev_release(vspg_ctx *ctx, std::vector<void *> &v) {
    for (void *p : v) {
        auto it = ctx->block_size.find(p);
        if (it == ctx->block_size.end()) cudaFree(p); else ctx->block_cache.emplace(it->second, p);
    }
    v.clear();
}
static void dev_trim(vspg_ctx *ctx) {
    for (auto &kv : ctx->block_cache) { ctx->block_size.erase(kv.second); cudaFree(kv.second); }
    ctx->block_cache.clear();
}
static void free_all(std::vector<void *> &v) { for (void *p : v) cudaFree(p); v.clear(); }

extern "C" {

const char *vspg_version(void) { return "vspg-b200 0.1 (sm_100a)"; }

const char *vspg_last_error(const vspg_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void vspg_integrator_defaults(vspg_integrator_params *p) {  // vspg.cpp:1263-1319
    p->max_depth = 5; p->min_rr_depth = 1; p->use_nee = 1; p->surface_guiding = 1; p->volume_guiding = 1;
    p->surface_guiding_ris = 1; p->volume_guiding_ris = 0; p->vsp_guiding = 1; p->vsp_primary_guiding = 1;
    p->vsp_secondary_guiding = 1; p->vsp_mis_ratio = 0.5f; p->vsp_criterion_variance = 1; p->vsp_sampling_nds = 0;
    p->collision_probability_bias = 0; p->light_sampler = VSPG_LIGHTSAMPLER_BVH; p->regularize = 0;
}

int vspg_create(vspg_ctx **out, int device_ordinal) {
    if (!out) return fail(nullptr, VSPG_ERR_INVALID, "vspg_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, VSPG_ERR_CUDA, "vspg_create: no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device_ordinal < 0 || device_ordinal >= n) return fail(nullptr, VSPG_ERR_INVALID, "vspg_create: device %d out of range [0,%d)", device_ordinal, n);
    vspg_ctx *ctx = new vspg_ctx();
    ctx->device = device_ordinal;
    if ((e = cudaSetDevice(device_ordinal)) != cudaSuccess) { delete ctx; return fail(nullptr, VSPG_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e)); }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_ordinal) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    vspg_integrator_defaults(&ctx->sc.ip);
    ctx->sc.ip.light_sampler = VSPG_LIGHTSAMPLER_UNIFORM;
    if (const char *sb = getenv("VSPG_SCATTER_BLOCK")) { int v = atoi(sb); if (v >= 32 && v <= 256 && v % 32 == 0) ctx->scatter_block = v; }
    {   // persisting-L2 carve-out for the density cell grid (vspg_ctx_render.inl)
        const char *lp = getenv("VSPG_L2_PERSIST");
        cudaDeviceProp prop;
        // Off unless asked for: measured on C2 (profiles/r01z_summary.md) the set-aside takes L2 away from the state-bound
        // kernels (k_surface 22.6 -> 63.6 ms per 3 renders) for a 3 ms gain in k_medium_rs.
        if (lp && atoi(lp) != 0 && cudaGetDeviceProperties(&prop, ctx->device) == cudaSuccess && prop.persistingL2CacheMaxSize > 0) {
            size_t want = (size_t)prop.persistingL2CacheMaxSize;
            if (lp && atoi(lp) > 1) want = std::min(want, (size_t)atoi(lp) << 20);     // VSPG_L2_PERSIST=<MiB>
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
                ctx->l2_persist_bytes = want; ctx->l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
            } else cudaGetLastError();
        }
    }
    if (const char *rr = getenv("VSPG_RS_REFILL")) { int v = atoi(rr); if (v >= 1 && v <= 32) ctx->rs_refill = v; }
    if (const char *g = getenv("VSPG_GRAPH")) ctx->use_graph = atoi(g) != 0;
    if (const char *g = getenv("VSPG_SORT")) ctx->sort_mode = atoi(g) & 3;
    if (const char *g = getenv("VSPG_BAND_ROWS")) ctx->band_mode = atoi(g);
    if (const char *g = getenv("VSPG_SHADOW_PERSIST")) ctx->shadow_persist = atoi(g) != 0;
    if (const char *g = getenv("VSPG_SHADOW_REFILL")) { int v = atoi(g); if (v >= 1 && v <= 32) ctx->shadow_refill = v; }
    if (const char *ro = getenv("VSPG_RS_OCC")) { int v = atoi(ro); if (v == 4 || v == 5) ctx->rs_occ = v; }
    const char *lay = getenv("VSPG_GRID_LAYOUT");
    if (lay && std::string(lay) == "linear") ctx->grid_layout = GRID_LINEAR;
    *out = ctx;
    return VSPG_OK;
}

void vspg_destroy(vspg_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    dev_release(ctx, ctx->scene_allocs); dev_release(ctx, ctx->guide_allocs); dev_trim(ctx); free_all(ctx->wf_allocs);
    if (ctx->wave_exec) cudaGraphExecDestroy(ctx->wave_exec);
    if (ctx->wave_graph) cudaGraphDestroy(ctx->wave_graph);
    if (ctx->cap_stream) cudaStreamDestroy(ctx->cap_stream);
    free(ctx->wave_key);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    if (ctx->film) cudaFree(ctx->film);
    if (ctx->tr_acc) cudaFree(ctx->tr_acc);
    if (ctx->isg_rec) cudaFree(ctx->isg_rec);
    if (ctx->isg_trained) cudaFree(ctx->isg_trained);
    if (ctx->counters) cudaFree(ctx->counters);
    for (auto e : ctx->ev) if (e) cudaEventDestroy(e);
    delete ctx;
}

int vspg_build_majorant_grid(const vspg_medium *m, float *out) {
    if (!m || !out || m->type == VSPG_MEDIUM_HOMOGENEOUS) return VSPG_ERR_INVALID;
    build_majorant_grid(m, out);
    return VSPG_OK;
}

int vspg_set_scene(vspg_ctx *ctx, const vspg_scene_desc *s) {
    if (!ctx || !s) return VSPG_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    if (s->camera.filter_type != 0 && s->camera.filter_type != 1) return fail(ctx, VSPG_ERR_UNSUPPORTED, "set_scene: pixel filter must be box (0) or gaussian (1)");
    if (s->camera.filter_type == 1 && (int(32 * s->camera.filter_radius[0]) < 1 || int(32 * s->camera.filter_radius[1]) < 1 || !(s->camera.filter_sigma > 0)))
        return fail(ctx, VSPG_ERR_INVALID, "set_scene: gaussian filter needs radius >= 1/32 and sigma > 0");
    if (s->camera.width <= 0 || s->camera.height <= 0) return fail(ctx, VSPG_ERR_INVALID, "set_scene: bad film resolution");
    if (s->camera.medium >= s->n_media) return fail(ctx, VSPG_ERR_INVALID, "set_scene: camera medium %d does not exist", s->camera.medium);
    for (int i = 0; i < s->n_lights; i++)
        if (s->lights[i].medium >= s->n_media) return fail(ctx, VSPG_ERR_INVALID, "set_scene: light %d: medium %d does not exist", i, s->lights[i].medium);
    dev_release(ctx, ctx->scene_allocs);
    ctx->have_scene = false;
    ctx->cells_ptr = nullptr; ctx->cells_bytes = 0;
    vspg_isg_store_end(ctx);    // so does an ISG store
    ctx->sc.tr_acc = nullptr;   // a Tr store belongs to one scene (resolution): vspg_tr_store_begin again after set_scene
    struct Trim { vspg_ctx *c; ~Trim() { dev_trim(c); } } trimOnExit{ctx};   // every return path drops the blocks this call did not reuse
    DScene &sc = ctx->sc;
    sc.cam = s->camera;
    // ---- media
    std::vector<DMedium> media(s->n_media);
    ctx->any_emissive = false; ctx->any_grid_medium = false;
    for (int i = 0; i < s->n_media; i++) {
        const vspg_medium &m = s->media[i];
        DMedium &d = media[i];
        memset(&d, 0, sizeof d);
        d.type = m.type;
        for (int a = 0; a < 3; a++) { d.bmin[a] = m.bounds_min[a]; d.bmax[a] = m.bounds_max[a]; d.sigma_a[a] = m.sigma_a[a]; d.sigma_s[a] = m.sigma_s[a]; d.mres[a] = m.maj_res[a]; }
        memcpy(d.mInv, m.render_from_medium.mInv, sizeof d.mInv);
        d.sigma_scale = m.sigma_scale; d.g = m.g; d.nx = m.nx; d.ny = m.ny; d.nz = m.nz;
        d.layout = GRID_LINEAR;
        {   // exact shortcuts (dscene.cuh DMedium::xlateOnly / extPow2)
            const float(*I)[4] = m.render_from_medium.mInv;
            bool x = true;
            for (int r = 0; r < 4; r++)
                for (int c = 0; c < 4; c++) {
                    if (r < 3 && c == 3) { if (!std::isfinite(I[r][c])) x = false; continue; }
                    if (I[r][c] != (r == c ? 1.f : 0.f)) x = false;
                }
            d.xlateOnly = x && !getenv("VSPG_NO_SHORTCUTS");
            d.extPow2 = 0;
            for (int a = 0; a < 3; a++) {
                float e = m.bounds_max[a] - m.bounds_min[a];
                int ex; float mant = std::frexp(e, &ex);
                d.invExt[a] = 1.f / e;
                if (e > 0 && mant == 0.5f && ex > -100 && ex < 100 && !getenv("VSPG_NO_SHORTCUTS")) d.extPow2 |= 1 << a;
            }
        }
        if (std::max({m.Le[0], m.Le[1], m.Le[2]}) > 0) {   // IsEmissive(): Le_spec.MaxValue() > 0 (media.h:250, media.cpp:261)
            if (m.type != VSPG_MEDIUM_HOMOGENEOUS && m.type != VSPG_MEDIUM_GRID)
                return fail(ctx, VSPG_ERR_UNSUPPORTED, "set_scene: medium %d: emission is built for homogeneous and uniformgrid media only", i);
            d.emissive = 1; memcpy(d.Le, m.Le, 12);
            ctx->any_emissive = true;
            if (m.type == VSPG_MEDIUM_GRID) {
                static const float one = 1.f;
                const bool given = m.le_scale != nullptr;
                if (given && (m.le_nx <= 0 || m.le_ny <= 0 || m.le_nz <= 0)) return fail(ctx, VSPG_ERR_INVALID, "set_scene: medium %d: le_scale grid without dimensions", i);
                d.leN[0] = given ? m.le_nx : 1; d.leN[1] = given ? m.le_ny : 1; d.leN[2] = given ? m.le_nz : 1;
                CK(upload(ctx, ctx->scene_allocs, given ? m.le_scale : &one, (size_t)d.leN[0] * d.leN[1] * d.leN[2], &d.leScale));
            }
        }
        if (m.type == VSPG_MEDIUM_HOMOGENEOUS) continue;
        ctx->any_grid_medium = true;
        const bool scalarGrid = m.type == VSPG_MEDIUM_GRID || m.type == VSPG_MEDIUM_NANOVDB;
        if (m.type == VSPG_MEDIUM_NANOVDB) {
            if (!m.majorant) return fail(ctx, VSPG_ERR_INVALID, "set_scene: medium %d (nanovdb) comes without its majorant grid", i);
            for (int a = 0; a < 3; a++) { d.imin[a] = m.index_min[a]; d.nvVec[a] = m.index_origin[a]; }
            for (int a = 0; a < 9; a++) d.nvInv[a] = m.index_from_medium[a];
            d.densityOffset = m.density_offset; d.bg = m.background;
        }
        if (!scalarGrid && m.type != VSPG_MEDIUM_RGBGRID) return fail(ctx, VSPG_ERR_UNSUPPORTED, "set_scene: medium %d has unsupported type %d", i, m.type);
        size_t nvox = (size_t)m.nx * m.ny * m.nz;
        if (nvox == 0 || m.maj_res[0] <= 0 || m.maj_res[1] <= 0 || m.maj_res[2] <= 0) return fail(ctx, VSPG_ERR_INVALID, "set_scene: medium %d has empty grid", i);
        if (scalarGrid) {
            if (!m.density) return fail(ctx, VSPG_ERR_INVALID, "set_scene: medium %d has no density grid", i);
            CK(upload(ctx, ctx->scene_allocs, m.density, nvox, &d.density));
        } else {
            if (m.sigma_a_rgb) CK(upload(ctx, ctx->scene_allocs, m.sigma_a_rgb, nvox * 3, &d.sa_rgb));
            if (m.sigma_s_rgb) CK(upload(ctx, ctx->scene_allocs, m.sigma_s_rgb, nvox * 3, &d.ss_rgb));
        }
        size_t nmaj = (size_t)m.maj_res[0] * m.maj_res[1] * m.maj_res[2];
        if (m.majorant) CK(upload(ctx, ctx->scene_allocs, m.majorant, nmaj, &d.majorant));
        else if (getenv("VSPG_HOST_MAJORANT")) {
            std::vector<float> maj(nmaj);
            build_majorant_grid(&m, maj.data());
            CK(upload(ctx, ctx->scene_allocs, maj.data(), nmaj, &d.majorant));
        } else {   // on the device, from the grids just uploaded (exact: a maximum has no rounding)
            void *dm = nullptr;
            CK(dev_alloc(ctx, ctx->scene_allocs, &dm, nmaj * sizeof(float)));
            if (m.type == VSPG_MEDIUM_GRID) k_build_majorant<<<(int)nmaj, 256>>>(d.density, nullptr, 1, 1.f, m.nx, m.ny, m.nz, m.maj_res[0], m.maj_res[1], m.maj_res[2], (float *)dm);
            else k_build_majorant<<<(int)nmaj, 256>>>(d.sa_rgb, d.ss_rgb, 3, m.sigma_scale, m.nx, m.ny, m.nz, m.maj_res[0], m.maj_res[1], m.maj_res[2], (float *)dm);
            CK(cudaGetLastError());
            d.majorant = (const float *)dm;
        }
        if (m.type == VSPG_MEDIUM_RGBGRID && d.sa_rgb && d.ss_rgb && ctx->grid_layout == GRID_CELL8) {
            size_t ncell = (size_t)(m.nx + 1) * (m.ny + 1) * (m.nz + 1);
            void *cells = nullptr;
            CK(dev_alloc(ctx, ctx->scene_allocs, &cells, ncell * 192));
            int blocks = (int)std::min<size_t>((2 * ncell + 255) / 256, 148 * 32);
            k_build_cells_rgb<<<blocks, 256>>>(d.sa_rgb, d.ss_rgb, m.nx, m.ny, m.nz, (float4 *)cells);
            CK(cudaGetLastError());
            d.cells_rgb = (const float4 *)cells;
            if (!ctx->cells_ptr) { ctx->cells_ptr = cells; ctx->cells_bytes = ncell * 192; }
        }
        if (scalarGrid && ctx->grid_layout == GRID_CELL8) {
            size_t ncell = (size_t)(m.nx + 1) * (m.ny + 1) * (m.nz + 1);
            void *cells = nullptr;
            CK(dev_alloc(ctx, ctx->scene_allocs, &cells, ncell * 32));
            int blocks = (int)std::min<size_t>((ncell + 255) / 256, 148 * 32);
            k_build_cells<<<blocks, 256>>>(d.density, m.nx, m.ny, m.nz, d.bg, (float4 *)cells);
            CK(cudaGetLastError());
            d.cells = (const float4 *)cells; d.layout = GRID_CELL8;
            if (!ctx->cells_ptr) { ctx->cells_ptr = cells; ctx->cells_bytes = ncell * 32; }
        }
    }
    CK(upload(ctx, ctx->scene_allocs, media.data(), media.size(), &sc.media));
    sc.n_media = s->n_media;
    ctx->all_gray = !getenv("VSPG_NO_GRAY");
    ctx->any_nvdb = false;
    ctx->bin_ok = false;
    for (int i = 0; i < s->n_media && !ctx->bin_ok; i++) {   // binning box = render-space bounds of the first grid medium
        const vspg_medium &m = s->media[i];
        if (m.type == VSPG_MEDIUM_HOMOGENEOUS) continue;
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int c = 0; c < 8; c++) {
            const float p[3] = {c & 1 ? m.bounds_max[0] : m.bounds_min[0], c & 2 ? m.bounds_max[1] : m.bounds_min[1], c & 4 ? m.bounds_max[2] : m.bounds_min[2]};
            const float(*M)[4] = m.render_from_medium.m;
            for (int a = 0; a < 3; a++) {
                float v = M[a][0] * p[0] + M[a][1] * p[1] + M[a][2] * p[2] + M[a][3];
                lo[a] = std::min(lo[a], v); hi[a] = std::max(hi[a], v);
            }
        }
        for (int a = 0; a < 3; a++) { sc.bin_lo[a] = lo[a]; sc.bin_scale[a] = hi[a] > lo[a] ? kBinCells / (hi[a] - lo[a]) : 0.f; }
        ctx->bin_ok = true;
    }
    for (const DMedium &d : media) ctx->any_nvdb |= d.type == VSPG_MEDIUM_NANOVDB;
    for (const DMedium &d : media)
        if (d.type == VSPG_MEDIUM_RGBGRID || d.sigma_a[0] != d.sigma_a[1] || d.sigma_a[0] != d.sigma_a[2] || d.sigma_s[0] != d.sigma_s[1] || d.sigma_s[0] != d.sigma_s[2])
            ctx->all_gray = false;
    // ---- quads + BVH
    // prims = quads, then triangles stored in the same record (p0,p1,p2 in p00,p10,p01; kind = PRIM_TRIANGLE), then full
    // spheres (kind = PRIM_SPHERE; the host-side record holds the corners of Sphere::Bounds for the BVH build)
    const int nPrims = s->n_quads + s->n_triangles + s->n_spheres;
    ctx->prim_kinds = (s->n_triangles == 0 && s->n_spheres == 0 && !getenv("VSPG_ALL_KINDS")) ? kKindsQuads : kKindsAll;
    std::vector<vspg_quad> hostPrims(nPrims);
    std::vector<DQuad> quads(nPrims);
    for (int i = 0; i < s->n_triangles; i++) {
        const vspg_triangle &t = s->triangles[i];
        vspg_quad &q = hostPrims[s->n_quads + i];
        memcpy(q.p00, t.p0, 12); memcpy(q.p10, t.p1, 12); memcpy(q.p01, t.p2, 12); memcpy(q.p11, t.p2, 12);
        q.material = t.material; q.area_light = t.area_light; q.medium_inside = t.medium_inside; q.medium_outside = t.medium_outside;
        q.has_interface = t.has_interface; q.flip_normal = t.flip_normal;
    }
    for (int i = 0; i < s->n_spheres; i++) {
        const vspg_sphere &t = s->spheres[i];
        vspg_quad &q = hostPrims[s->n_quads + s->n_triangles + i];
        if (!(t.radius > 0)) return fail(ctx, VSPG_ERR_INVALID, "set_scene: sphere %d has radius <= 0", i);
        for (int a = 0; a < 3; a++) {   // shapes.cpp:33-36 Sphere::Bounds through translate(center)
            q.p00[a] = q.p10[a] = q.p01[a] = -t.radius + t.center[a];
            q.p11[a] = t.radius + t.center[a];
        }
        q.material = t.material; q.area_light = t.area_light; q.medium_inside = t.medium_inside; q.medium_outside = t.medium_outside;
        q.has_interface = t.has_interface; q.flip_normal = t.flip_normal;
    }
    float bmin[3] = {INFINITY, INFINITY, INFINITY}, bmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < s->n_quads; i++) hostPrims[i] = s->quads[i];
    for (int i = 0; i < nPrims; i++) {
        const vspg_quad &q = hostPrims[i];
        DQuad &d = quads[i];
        d.kind = i < s->n_quads ? PRIM_QUAD : i < s->n_quads + s->n_triangles ? PRIM_TRIANGLE : PRIM_SPHERE;
        memcpy(d.p00, q.p00, 12); memcpy(d.p10, q.p10, 12); memcpy(d.p01, q.p01, 12); memcpy(d.p11, q.p11, 12);
        if (d.kind == PRIM_SPHERE) {
            const vspg_sphere &t = s->spheres[i - s->n_quads - s->n_triangles];
            memcpy(d.p00, t.center, 12); d.p10[0] = t.radius; d.p10[1] = d.p10[2] = 0;
            memset(d.p01, 0, 12); memset(d.p11, 0, 12);
        }
        d.material = q.material; d.area_light = q.area_light; d.medium_inside = q.medium_inside; d.medium_outside = q.medium_outside;
        d.has_interface = q.has_interface; d.flip_normal = q.flip_normal;
        if (q.material >= s->n_materials || q.area_light >= s->n_lights || q.medium_inside >= s->n_media || q.medium_outside >= s->n_media)
            return fail(ctx, VSPG_ERR_INVALID, "set_scene: primitive %d references an out-of-range material/light/medium", i);
        H3 p00(q.p00), p10(q.p10), p01(q.p01), p11(q.p11);
        if (d.kind == PRIM_QUAD) { d.is_rect = quad_is_rectangle(p00, p10, p01, p11) ? 1 : 0; d.area = quad_area(p00, p10, p01, p11); }
        else if (d.kind == PRIM_TRIANGLE) { d.is_rect = 0; d.area = 0.5f * hlen(hcross(p10 - p00, p01 - p00)); }   // shapes.h Triangle::Area
        else { d.is_rect = 0; d.area = 6.28318548f * d.p10[0] * (d.p10[0] - -d.p10[0]); }   // Sphere::Area = phiMax * radius * (zMax - zMin)
        const float *P[4] = {q.p00, q.p10, q.p01, q.p11};
        for (int k = 0; k < 4; k++)
            for (int a = 0; a < 3; a++) { bmin[a] = std::min(bmin[a], P[k][a]); bmax[a] = std::max(bmax[a], P[k][a]); }
    }
    CK(upload(ctx, ctx->scene_allocs, quads.data(), quads.size(), &sc.quads));
    sc.n_quads = nPrims;
    HostBvh bvh = build_bvh(hostPrims.data(), nPrims);
    static_assert(sizeof(HostBvhNode) == sizeof(DBvhNode) && sizeof(DBvhNode) == 48, "BVH node layout");
    CK(upload(ctx, ctx->scene_allocs, (const DBvhNode *)bvh.nodes.data(), bvh.nodes.size(), &sc.bvh));
    CK(upload(ctx, ctx->scene_allocs, bvh.prim_ids.data(), bvh.prim_ids.size(), &sc.prim_ids));
    sc.n_bvh = (int)bvh.nodes.size();
    // Bounds3::BoundingSphere of the aggregate bounds (lights' Preprocess)
    if (nPrims) {
        H3 c((bmin[0] + bmax[0]) / 2, (bmin[1] + bmax[1]) / 2, (bmin[2] + bmax[2]) / 2);
        sc.scene_center[0] = c.x; sc.scene_center[1] = c.y; sc.scene_center[2] = c.z;
        sc.scene_radius = hdist(c, H3(bmax[0], bmax[1], bmax[2]));
    } else { sc.scene_center[0] = sc.scene_center[1] = sc.scene_center[2] = 0; sc.scene_radius = 0; }
    // ---- materials, lights
    for (int i = 0; i < s->n_materials; i++)
        if (s->materials[i].type != VSPG_MATERIAL_DIFFUSE) return fail(ctx, VSPG_ERR_UNSUPPORTED, "set_scene: material %d: only diffuse is supported", i);
    CK(upload(ctx, ctx->scene_allocs, s->materials, (size_t)s->n_materials, &sc.materials));
    sc.n_materials = s->n_materials;
    std::vector<DLight> lights(s->n_lights);
    std::vector<int> inf;
    std::vector<float> lightArea(s->n_lights, 0.f);
    for (int i = 0; i < s->n_lights; i++) {
        const vspg_light &l = s->lights[i];
        DLight &d = lights[i];
        d.type = l.type; memcpy(d.pos, l.pos, 12);
        for (int c = 0; c < 3; c++) d.L[c] = l.scale * l.L[c];   // scale * Lemit->Sample(lambda) (lights.h:169-171, 495-506; lights.cpp:267, 1031)
        d.shape_index = l.shape_index; d.two_sided = l.two_sided; d.medium = l.medium; d.pmf = 0; d.q = 1; d.alias = -1;
        switch (l.type) {
        case VSPG_LIGHT_DIFFUSE_AREA:
        {   // the device light carries the GLOBAL primitive id (quads, then triangles, then spheres)
            const int nOf[3] = {s->n_quads, s->n_triangles, s->n_spheres}, baseOf[3] = {0, s->n_quads, s->n_quads + s->n_triangles};
            if (l.shape_kind < 0 || l.shape_kind > 2 || l.shape_index < 0 || l.shape_index >= nOf[l.shape_kind])
                return fail(ctx, VSPG_ERR_INVALID, "set_scene: light %d: area light references shape kind %d index %d, which does not exist", i, l.shape_kind, l.shape_index);
            d.shape_index = baseOf[l.shape_kind] + l.shape_index;
            if (quads[d.shape_index].area_light != i) return fail(ctx, VSPG_ERR_INVALID, "set_scene: light %d and its shape do not reference each other", i);
            lightArea[i] = quads[d.shape_index].area;
            break;
        }
        case VSPG_LIGHT_POINT: break;
        case VSPG_LIGHT_SPOT:
            if (!(l.cos_falloff_end <= l.cos_falloff_start)) return fail(ctx, VSPG_ERR_INVALID, "set_scene: light %d: spot light with cos_falloff_end %g > cos_falloff_start %g", i, l.cos_falloff_end, l.cos_falloff_start);
            memcpy(d.I, l.L, 12); d.scale = l.scale; memcpy(d.toLight, l.light_from_render, 36); d.cosStart = l.cos_falloff_start; d.cosEnd = l.cos_falloff_end;
            break;
        case VSPG_LIGHT_DISTANT: case VSPG_LIGHT_UNIFORM_INFINITE: inf.push_back(i); break;   // Integrator::infiniteLights (cpu/integrators.h:77-81)
        default: return fail(ctx, VSPG_ERR_UNSUPPORTED, "set_scene: light %d has unsupported type %d", i, l.type);
        }
    }
    {   // PowerLightSampler: alias table over Phi / lambda.PDF() (host_lightpower.h)
        HostPowerSampler ps = build_power_sampler(s->lights, s->n_lights, lightArea.data(), sc.scene_radius);
        for (int i = 0; i < s->n_lights; i++) { lights[i].pmf = ps.p[i]; lights[i].q = ps.q[i]; lights[i].alias = ps.alias[i]; }
    }
    CK(upload(ctx, ctx->scene_allocs, lights.data(), lights.size(), &sc.lights));
    CK(upload(ctx, ctx->scene_allocs, inf.data(), inf.size(), &sc.inf_lights));
    sc.n_lights = s->n_lights; sc.n_inf_lights = (int)inf.size();
    {   // BVHLightSampler (the reference's default "lightsampler"), lightsamplers.cpp:107-236
        std::vector<float> primArea(nPrims);
        for (int i = 0; i < nPrims; i++) primArea[i] = quads[i].area;
        HostLightBvh lb = build_light_bvh(s, primArea.data());
        static_assert(sizeof(HostLightNode) == sizeof(DLightNode) && sizeof(DLightNode) == 64, "light BVH node layout");
        CK(upload(ctx, ctx->scene_allocs, (const DLightNode *)lb.nodes.data(), lb.nodes.size(), &sc.lbvh));
        CK(upload(ctx, ctx->scene_allocs, (const unsigned *)lb.bitTrail.data(), lb.bitTrail.size(), &sc.light_trail));
        CK(upload(ctx, ctx->scene_allocs, lb.inTree.data(), lb.inTree.size(), &sc.light_in_tree));
        CK(upload(ctx, ctx->scene_allocs, lb.infinite.data(), lb.infinite.size(), &sc.unbounded_lights));
        sc.n_lbvh = (int)lb.nodes.size(); sc.n_unbounded = (int)lb.infinite.size();
    }
    // ---- pixel filter
    sc.filt_nx = sc.filt_ny = 0; sc.filt_f = sc.filt_cond_cdf = sc.filt_cond_int = sc.filt_marg_cdf = nullptr; sc.filt_marg_int = 0;
    if (s->camera.filter_type == 1) {
        HostFilterTable T = build_gaussian_filter(s->camera.filter_radius[0], s->camera.filter_radius[1], s->camera.filter_sigma);
        CK(upload(ctx, ctx->scene_allocs, T.f.data(), T.f.size(), &sc.filt_f));
        CK(upload(ctx, ctx->scene_allocs, T.condCdf.data(), T.condCdf.size(), &sc.filt_cond_cdf));
        CK(upload(ctx, ctx->scene_allocs, T.condInt.data(), T.condInt.size(), &sc.filt_cond_int));
        CK(upload(ctx, ctx->scene_allocs, T.margCdf.data(), T.margCdf.size(), &sc.filt_marg_cdf));
        sc.filt_nx = T.nx; sc.filt_ny = T.ny; sc.filt_marg_int = T.margInt;
    }
    // ---- film
    if (ctx->film_w != s->camera.width || ctx->film_h != s->camera.height || !ctx->film) {
        if (ctx->film) cudaFree(ctx->film);
        ctx->film = nullptr;
        ctx->film_w = s->camera.width; ctx->film_h = s->camera.height;
        CK(cudaMalloc((void **)&ctx->film, (size_t)ctx->film_w * ctx->film_h * 4 * sizeof(double)));
    }
    CK(cudaMemset(ctx->film, 0, (size_t)ctx->film_w * ctx->film_h * 4 * sizeof(double)));
    if (!ctx->counters) { CK(cudaMalloc((void **)&ctx->counters, CNT_COUNT * sizeof(unsigned long long))); CK(cudaMemset(ctx->counters, 0, CNT_COUNT * sizeof(unsigned long long))); }
    CK(cudaDeviceSynchronize());
    ctx->have_scene = true;
    return VSPG_OK;
}

}  // extern "C"

#include "vspg_ctx_guiding.inl"
#include "vspg_ctx_render.inl"
#include "vspg_ctx_debug.inl"
#include "vspg_ctx_noalloc.inl"
