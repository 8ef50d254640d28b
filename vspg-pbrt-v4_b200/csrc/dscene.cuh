// Device-resident scene, guiding inputs and wavefront state (all pointers are HBM addresses).
#pragma once
#include "../../include/vspg.h"
#include "dmath.cuh"

namespace vspg {

// Density-grid layouts in HBM.
//  LINEAR   : the reference's SampledGrid order (x fastest), 8 scalar gathers per trilinear query.
//  CELL8    : one 32-byte record per interpolation cell holding its 8 corner densities
//             {d000,d100,d010,d110,d001,d101,d011,d111}; cells cover ix in [-1,nx-1] (zero padded), so
//             a query is exactly one aligned 32-byte sector = the algorithmic 32 B of SURVEY 8(d).
enum { GRID_LINEAR = 0, GRID_CELL8 = 1 };

struct DMedium {
    int type;
    float bmin[3], bmax[3];
    float mInv[4][4];     // render_from_medium inverse (util/transform.h ApplyInverse)
    int identity;         // mInv == I (skips nothing arithmetically; informational)
    float sigma_a[3], sigma_s[3];
    float sigma_scale, g;
    int nx, ny, nz;
    const float *density;      // LINEAR
    const float4 *cells;       // CELL8: 2 x float4 per cell, (nx+1)*(ny+1)*(nz+1) cells
    const float *sa_rgb, *ss_rgb;
    const float4 *cells_rgb;   // rgbgrid, both grids present: 12 x float4 per cell = the 8 rgb corners of sigma_a, then of sigma_s (192 B)
    int mres[3];
    const float *majorant;
    int layout;
    // VSPG_MEDIUM_NANOVDB: `density` / `cells` hold the grid flattened over [imin, imin + (nx,ny,nz)); voxels outside
    // read `bg`.  pIndex = nvInv * (p - nvVec) (NanoVDB Map::applyInverseMapF), the density gets `densityOffset` added.
    int imin[3];
    float nvInv[9], nvVec[3];
    float densityOffset, bg;       // bg = 0 for pbrt's own grids (SampledGrid::Lookup outside the grid)
    // exact shortcuts, set by set_scene when they are bit-neutral (dmedia.cuh): mInv is a pure translation; the bounds'
    // extent along an axis is a power of two, so that x / extent == x * (1 / extent) for every x
    int xlateOnly, extPow2;        // extPow2: bit a set = axis a
    float invExt[3];
    // emission (media.h:243-260, 329-345): Le rgb; grid media scale it by a trilinear lookup of leScale (leN voxels)
    int emissive;
    float Le[3];
    const float *leScale;
    int leN[3];
};

// One surface primitive: a bilinear patch (p00,p10,p01,p11) or a triangle (p0,p1,p2 in p00,p10,p01; p11 unused).
// Quads come first in the array (prim id == quad index, so area lights index it directly), triangles after them,
// then full spheres.
enum { PRIM_QUAD = 0, PRIM_TRIANGLE = 1, PRIM_SPHERE = 2 };   // sphere: centre in p00, radius in p10[0]
struct DQuad {
    float p00[3], p10[3], p01[3], p11[3];
    int material, area_light, medium_inside, medium_outside, has_interface, flip_normal;
    int is_rect;          // BilinearPatch::IsRectangle (shapes.h), precomputed on the host
    float area;
    int kind;             // PRIM_QUAD / PRIM_TRIANGLE / PRIM_SPHERE
};

// Threaded (stackless) BVH node: DFS order, `skip` = next node when this subtree is done/missed.
struct DBvhNode {
    float bmin[3]; int skip;
    float bmax[3]; int prim_begin;   // leaf: first index into prim_ids; interior: -1
    int prim_count; int pad[3];
};

struct DLight {
    int type;
    float L[3];
    float pos[3];
    int shape_index, two_sided, medium;
    float pmf;            // PowerLightSampler: AliasTable bins[i].p
    float q; int alias;   //                    bins[i].q, bins[i].alias (util/sampling.cpp:563-618)
    // spot lights (lights.h:852-866): unscaled intensity and scale apart (I() multiplies SmoothStep * scale first), the
    // rows of the render -> light rotation, the cone
    float I[3], scale, toLight[9], cosStart, cosEnd;
};

// Light-BVH node (lightsamplers.h LightBVHNode), stored decoded -- see host_lightbvh.h.
struct DLightNode {
    float bmin[3], bmax[3], w[3];
    float phi, cosTheta_o, cosTheta_e;
    int twoSided, childOrLight, isLeaf, pad;
};

struct DField {
    int present, usable;
    float bmin[3], bmax[3];
    int n_nodes, n_regions;
    const vspg_kdnode *nodes;
    const vspg_mixture *regions;
    const vspg_neighbours *neigh;
    int stochastic_nn;
};

struct DHGRow { int K; float g; float w[4], mc[4], kappa[4], norm[4]; };

struct DScene {
    vspg_camera cam;
    int n_media; const DMedium *media;
    int n_quads; const DQuad *quads;
    int n_bvh; const DBvhNode *bvh; const int *prim_ids;
    int n_materials; const vspg_material *materials;
    int n_lights; const DLight *lights;
    int n_inf_lights; const int *inf_lights;
    // BVHLightSampler (lightsamplers.h:258-404): tree nodes, per-light bit trail / in-tree flag, lights without bounds
    int n_lbvh; const DLightNode *lbvh; const unsigned *light_trail; const int *light_in_tree;
    int n_unbounded; const int *unbounded_lights;
    float scene_center[3]; float scene_radius;
    DField surf, vol;
    const DHGRow *hg_table;   // 128 rows
    const float *isg_vsp; int isg_w, isg_h;
    const float *tr_rgb; int tr_w, tr_h;
    // storeISGBuffer (vspg_isg_store_*): per pixel {surfSpp, volSpp, surfContrib rgb, volContrib rgb} as the reference's
    // ImageSpaceGuidingBuffer keeps them (float running means), or null; isg_variance = its vspType is EVariance
    float *isg_rec; int isg_variance;
    double *tr_acc;      // storeTrBuffer accumulators {sumR, sumG, sumB, count} per pixel, or null (vspg_tr_store_begin)
    // PixelFilter "gaussian" sampling table (host_util.h build_gaussian_filter); filt_nx == 0 for the box filter
    int filt_nx, filt_ny; const float *filt_f, *filt_cond_cdf, *filt_cond_int, *filt_marg_cdf; float filt_marg_int;
    float bin_lo[3], bin_scale[3];   // kbin.cuh: render-space box the stage queues are binned over, cells per unit length
    vspg_integrator_params ip;
};

// ------------------------------------------------------------------ wavefront state (SoA, one slot per path in flight)
// One float4 stream of the path state.  Slots are visited in queue order, i.e. at random, and a DRAM/L2 sector is 32 B:
// a lone 16-byte element wastes half of every sector it touches (k_surface moved 375 MB per launch for 190 MB of state,
// profiles/r01j).  Streams that are always read or written together therefore share 32-byte records (STRIDE 2, OFF 0/1);
// the accessor keeps the `P.name[slot]` spelling of a plain array.
template <int STRIDE, int OFF>
struct F4S {
    float4 *base;
    __device__ __forceinline__ float4 &operator[](int s) const { return base[(size_t)s * STRIDE + OFF]; }
};
struct RngS {   // the sampler state shares 32-byte records with the hit record (both are read at every surface vertex)
    float4 *base;
    __device__ __forceinline__ ulonglong2 &operator[](int s) const { return reinterpret_cast<ulonglong2 *>(base)[(size_t)s * 2]; }
};
struct LkS {    // the region id chosen by k_lookup rides in the record of sh_ple, which k_lookup writes and the scatter kernels read anyway
    float4 *base;
    __device__ __forceinline__ int &operator[](int s) const { return *reinterpret_cast<int *>(base + (size_t)s * 2 + 1); }
};
struct PathSoA {
    F4S<2, 0> o_dx;      // ray.o xyz, d.x
    F4S<2, 1> d_med;     // d.y d.z, medium(int), packed flags(int)
    F4S<2, 0> beta_ru;   // beta xyz, r_u.x
    F4S<2, 1> ru_rl;     // r_u.y r_u.z r_l.x r_l.y
    F4S<2, 0> rl_L;      // r_l.z, L xyz
    RngS rng;            // sampler PCG32 state, inc
    F4S<2, 1> misc;      // rr_correction, vsp_next, prevCtx.p.x, prevCtx.p.y
    F4S<2, 0> misc2;     // prevCtx.p.z, pixel(int), sampleIndex(int), unused
    F4S<2, 1> hit;       // t, u, v, prim(int)  (prim < 0: no hit)
    float4 *ev;          // medium scatter event: p xyz, g
    float4 *prev_o;      // ray origin spawned at the previous SURFACE vertex (isect.SpawnRay(wi)): Shape::PDF re-intersects the
                         // emitter from there for the emission MIS weight; written by k_shade, read by k_surface on emitter hits only
    // shadow-ray record emitted by the scatter kernels
    F4S<2, 0> sh_o;      // o xyz, d.x
    F4S<2, 1> sh_d;      // d.y d.z, medium(int), flags(int: bit0 delta light, bits 8.. channel)
    F4S<2, 0> sh_c;      // ls.L xyz, beta.x
    F4S<2, 1> sh_rl;     // r_p*p_l xyz, beta.y
    float4 *sh_f;        // f_hat xyz, beta.z
    F4S<2, 0> sh_ru;     // r_p*scatterPDF xyz, nLight.z
    F4S<2, 1> sh_pl;     // pLight xyz, nLight.x
    F4S<2, 0> sh_ple;    // pLightErr xyz, nLight.y
    F4S<2, 1> prev_ns;   // (shares records with misc2) shading normal of the previous vertex (0 for a medium vertex): LightSampleContext::ns for the emission MIS
    LkS lk;              // guiding-cache region chosen by k_lookup for the pending scatter vertex (-1 = none)
    float2 *rs_pre;      // resampling pre-pass (media_sampleTMaj.h:153-164): sum of sigma_maj[ch]*dt, DDA steps(int)
};

// packed flags in d_med.w : bits 0-7 depth, 8-9 channel, 10 specularBounce, 11 anyNonSpecular, 12 lastVertexVolume
struct PathFlags {
    int depth, channel; bool specular, anyNonSpecular, lastVertexVolume;
    bool firstVolume;   // the path's first scattering event was inside a medium (isgbSample.SetSurfaceEvent(false), vspg.cpp:776-783, 950-957)
    VD int pack() const { return (depth & 0xff) | (channel << 8) | (specular << 10) | (anyNonSpecular << 11) | (lastVertexVolume << 12) | (firstVolume << 13); }
    VD static PathFlags unpack(int v) {
        PathFlags f; f.depth = v & 0xff; f.channel = (v >> 8) & 3; f.specular = (v >> 10) & 1; f.anyNonSpecular = (v >> 11) & 1;
        f.lastVertexVolume = (v >> 12) & 1; f.firstVolume = (v >> 13) & 1; return f;
    }
};

struct Queue { int *items; int *count; };

// Per-render arguments of vspg_render (window, sample-index range, sampler seed), uploaded before each render.
struct RenderVars { unsigned long long total_samples; int x0, y0, x1, y1, spp_begin, spp_stride, n_spp, seed, band_rows; };

struct Wavefront {
    PathSoA P;
    int capacity;
    Queue active, next, medium, surface, shade, volscatter, shadow, done, medium_rs, vol8, vol16, vol32, sh8, sh16, sh32;
    unsigned long long *next_sample;   // regeneration cursor
    const RenderVars *rv;              // what changes from one render() to the next; in device memory so that the
                                       // captured wave graph (vspg_ctx_render.inl) stays valid across renders
    double *film;                      // H*W*4
    unsigned long long *counters;      // vspg_counters layout (8 x u64)
    int *cursors;                      // per-kernel dynamic work cursors
};

// CNT_QUERIES / CNT_DDA: issued by the SampleDistance kernels; *_SH: by the shadow-ray kernel; CNT_DDA_PRE: the
// resampling routine's pre-pass (total majorant optical depth), which runs inside k_intersect
enum { CNT_PATHS = 0, CNT_QUERIES = 1, CNT_DDA = 2, CNT_LOOKUPS = 3, CNT_BVH = 4, CNT_SHADOW = 5, CNT_QUERIES_SH = 6, CNT_DDA_SH = 7, CNT_DDA_PRE = 8, CNT_COUNT = 9 };

// ------------------------------------------------------------------ warp-aggregated queue push (ballot + popc prefix)
VD void queue_push(const Queue &q, bool pred, int value) {
    unsigned mask = __ballot_sync(__activemask(), pred);
    if (!pred) return;
    int lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(q.count, __popc(mask));
    base = __shfl_sync(mask, base, leader);
    q.items[base + __popc(mask & ((1u << lane) - 1))] = value;
}

VD void counter_add(unsigned long long *counters, int which, unsigned v) {
    // warp-reduce (REDUX) then one atomic per warp
    unsigned mask = __activemask();
    unsigned s = __reduce_add_sync(mask, v);
    if ((threadIdx.x & 31) == (__ffs(mask) - 1) && s) atomicAdd(&counters[which], (unsigned long long)s);
}

}  // namespace vspg
