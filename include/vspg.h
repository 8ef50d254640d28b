/*
 * vspg.h -- C ABI of the B200-native render pass of GuidedVolPathVSPGIntegrator.
 *
 * This is the drop-in boundary.  Every entry point below
 * replaces a piece of the reference's seam
 *     Integrator::Create(name, params, camera, sampler, aggregate, lights, ...)   src/pbrt/cpu/integrators.cpp:3711-3768
 *     GuidedVolPathVSPGIntegrator::GuidedVolPathVSPGIntegrator(...)              src/pbrt/cpu/guidedvolpathvspgintegrator.cpp:61-198
 *     Integrator::Render()  (ImageTileIntegrator::Render)                        src/pbrt/cpu/integrators.cpp:75-269
 *     Film::AddSample / GetPixelRGB / WriteImage                                 src/pbrt/film.h:251-290
 * and of the OpenPGL C API the integrator consumes (pglFieldLoadFromFile,
 * pglFieldInit{Surface,Volume}SamplingDistribution, ...; openpgl/api/api.cpp:81-945).
 *
 * Plain C: pointers + sizes, int return codes (0 = ok, negative = error; the
 * message is available through vspg_last_error).  Nothing here throws, exits or
 * exposes torch / CUDA types (the stream is passed as void*).
 * All host arrays are borrowed for the duration of the call only; the context
 * copies what it needs into HBM and owns that device memory.
 */
#ifndef VSPG_H
#define VSPG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VSPG_OK 0
#define VSPG_ERR_INVALID (-1)   /* bad argument / inconsistent description          */
#define VSPG_ERR_CUDA (-2)      /* CUDA runtime failure (message has the detail)      */
#define VSPG_ERR_IO (-3)        /* file missing / malformed                           */
#define VSPG_ERR_UNSUPPORTED (-4) /* scene feature outside the hot-path scope */
#define VSPG_ERR_STATE (-5)     /* call order (e.g. render before set_scene)          */

/* ------------------------------------------------------------------ scene POD */

/* Row-major 4x4 and its inverse, as pbrt::Transform stores them (util/transform.h). */
typedef struct vspg_transform {
    float m[4][4];
    float mInv[4][4];
} vspg_transform;

/* cameras.cpp:410 PerspectiveCamera::GenerateRay.  camera_from_raster and
 * render_from_camera are built by the front end exactly as pbrt does
 * (cameras.h ProjectiveCamera ctor; CameraTransform with the default
 * "cameraworld" render space: render = world translated by -cameraPosition). */
typedef struct vspg_camera {
    vspg_transform camera_from_raster;
    vspg_transform render_from_camera;
    float lens_radius;
    float focal_distance;
    int medium;              /* medium the camera sits in, -1 = vacuum */
    int width, height;       /* full film resolution                    */
    float filter_radius[2];  /* box filter radius (filters.h BoxFilter) */
    int filter_type;         /* 0 = box, 1 = gaussian                    */
    float filter_sigma;      /* gaussian only                            */
    float imaging_ratio;     /* film.cpp:233 exposureTime*ISO/100        */
    float max_component_value; /* film.h:257 clamp; +inf = none         */
} vspg_camera;

enum { VSPG_MEDIUM_HOMOGENEOUS = 0, VSPG_MEDIUM_GRID = 1, VSPG_MEDIUM_RGBGRID = 2, VSPG_MEDIUM_NANOVDB = 3 };

/* media.h:221-467.  sigma_a/sigma_s are the RGB spectra ALREADY multiplied by
 * the medium's "scale" for homogeneous/uniformgrid (their ctors call
 * sigma_*_spec.Scale(sigmaScale), media.h:240-242 / media.cpp:249-250); for
 * rgbgrid sigma_scale is applied at lookup time (media.h:427-432). */
typedef struct vspg_medium {
    int type;
    float bounds_min[3], bounds_max[3];
    vspg_transform render_from_medium;
    float sigma_a[3], sigma_s[3];
    float sigma_scale;
    float g;
    int nx, ny, nz;
    const float *density;      /* uniformgrid: nx*ny*nz, x fastest (containers.h:833)   */
    const float *sigma_a_rgb;  /* rgbgrid: nx*ny*nz*3 or NULL (=> 1)                     */
    const float *sigma_s_rgb;  /* rgbgrid: nx*ny*nz*3 or NULL (=> 1)                     */
    int maj_res[3];            /* 16,16,16 for grid/rgbgrid (media.cpp:246,382)          */
    const float *majorant;     /* maj_res product floats; NULL => built by set_scene     */
    /* nanovdb (media.h:657-753, media.cpp:549-734): `density` is the float grid flattened over its index bounding box
     * [index_min, index_min + (nx,ny,nz)) -- voxels outside hold `background` --, sampled at
     * pIndex = index_from_medium * (p - index_origin) (NanoVDB Map::applyInverseMapF) with NanoVDB's trilinear sampler,
     * plus density_offset; bounds = the grid's world bounding box; maj_res = 64^3 and `majorant` is given (the front end
     * builds it as NanoVDBMedium's constructor does, majorantscale included). */
    int index_min[3];
    float index_from_medium[9];
    float index_origin[3];
    float density_offset;
    float background;
    float majorant_scale;      /* "majorantscale" the given majorant grid was built with (informational) */
    /* emission, homogeneous and uniformgrid media (media.h:243-260, 329-345; media.cpp:188-194, 304-329).  Le is the "Le"
     * rgb -- homogeneous: already times "Lescale".  uniformgrid: the emitted radiance at p is s * Le where
     * s = trilinear lookup of le_scale (le_nx * le_ny * le_nz floats, x fastest, same lookup as the density) is positive;
     * le_scale NULL with Le != 0 stands for the reference's 1 x 1 x 1 grid holding 1 when no "Lescale" array is given.
     * The medium is emissive when max(Le) > 0.  Only the delta-tracking branch of SampleDistance adds medium emission
     * (guidedvolpathvspgintegrator.cpp:895-906); the resampling branch ignores it, as the reference does. */
    float Le[3];
    const float *le_scale;
    int le_nx, le_ny, le_nz;
} vspg_medium;

/* One bilinear patch (shapes.h:1350) in render space; vertex order p00,p10,p01,p11. */
typedef struct vspg_quad {
    float p00[3], p10[3], p01[3], p11[3];
    int material;        /* index into materials; -1 = "interface" (null BSDF)            */
    int area_light;      /* index into lights (type DIFFUSE_AREA) or -1                    */
    int medium_inside;   /* MediumInterface (interaction.h:117-121); -1 = vacuum           */
    int medium_outside;
    int has_interface;   /* 0: primitive has no MediumInterface => ray keeps its medium    */
    int flip_normal;     /* reverseOrientation ^ transformSwapsHandedness                 */
} vspg_quad;

/* One triangle (shapes.h:833) in render space. */
typedef struct vspg_triangle {
    float p0[3], p1[3], p2[3];
    int material, area_light, medium_inside, medium_outside, has_interface, flip_normal;
} vspg_triangle;

/* One sphere (shapes.h:107), full sphere only, object space = translate(center). */
typedef struct vspg_sphere {
    float center[3];
    float radius;
    int material, area_light, medium_inside, medium_outside, has_interface, flip_normal;
} vspg_sphere;

enum { VSPG_MATERIAL_DIFFUSE = 0 };
typedef struct vspg_material {
    int type;
    float reflectance[3]; /* DiffuseBxDF R (bxdfs.h:31-82) */
} vspg_material;

enum {
    VSPG_LIGHT_DIFFUSE_AREA = 0, /* lights.h:458 */
    VSPG_LIGHT_POINT = 1,
    VSPG_LIGHT_DISTANT = 2,
    VSPG_LIGHT_UNIFORM_INFINITE = 3,
    VSPG_LIGHT_SPOT = 4          /* lights.h:825-880, lights.cpp:1425-1555 */
};
typedef struct vspg_light {
    int type;
    float L[3];        /* emitted radiance / intensity as the "rgb L" / "rgb I" parameter gives it (RGB mode)  */
    float scale;       /* the light's scale (the "scale" parameter, times power / illuminance normalisation): the
                          reference keeps it apart from the spectrum (lights.h LightBase) and the light samplers'
                          power estimates multiply the two in their own order, so the boundary keeps both   */
    float pos[3];      /* point: position (render space); distant: direction TO the light */
    int shape_kind;    /* area: 0 quad, 1 triangle, 2 sphere                             */
    int shape_index;   /* area: index into that shape array                              */
    int two_sided;     /* area                                                           */
    int medium;        /* point / spot: medium the light sits in; -1 vacuum              */
    /* spot (all zero otherwise): the light's frame and cone as SpotLight keeps them */
    float light_from_render[9]; /* rows of the upper 3x3 of renderFromLight's inverse: SampleLi takes the direction
                                   to the shading point into the light's frame with it (lights.h:857)           */
    float axis[3];              /* Normalize(renderFromLight(0,0,1)): the cone axis of Bounds() (lights.cpp:1459) */
    float cos_falloff_start;    /* cos(Radians(coneangle - conedeltaangle)) (lights.cpp:1439)                   */
    float cos_falloff_end;      /* cos(Radians(coneangle)) (lights.cpp:1438)                                    */
} vspg_light;

enum { VSPG_LIGHTSAMPLER_UNIFORM = 0, VSPG_LIGHTSAMPLER_POWER = 1, VSPG_LIGHTSAMPLER_BVH = 2 };

typedef struct vspg_scene_desc {
    vspg_camera camera;
    int n_media;      const vspg_medium *media;
    int n_quads;      const vspg_quad *quads;
    int n_triangles;  const vspg_triangle *triangles;
    int n_spheres;    const vspg_sphere *spheres;
    int n_materials;  const vspg_material *materials;
    int n_lights;     const vspg_light *lights;
} vspg_scene_desc;

/* ------------------------------------------------------------ guiding inputs */

/* openpgl KDNode (spatial/kdtree/KDTree.h:20-178): splitDim in the top 2 bits
 * (3 = leaf), child / data index in the low 30. */
typedef struct vspg_kdnode {
    float split;
    uint32_t dim_idx;
} vspg_kdnode;

#define VSPG_MAX_LOBES 32
/* One region's ParallaxAwareVonMisesFisherMixture<4,32,true>
 * (directional/vmm/ParallaxAwareVonMisesFisherMixture.h:64-81) flattened in the
 * lobe order of the file (vector k, lane j => lobe 4k+j). */
typedef struct vspg_mixture {
    float weight[VSPG_MAX_LOBES];
    float kappa[VSPG_MAX_LOBES];
    float mu_x[VSPG_MAX_LOBES], mu_y[VSPG_MAX_LOBES], mu_z[VSPG_MAX_LOBES];
    float norm[VSPG_MAX_LOBES];
    float e_minus_2k[VSPG_MAX_LOBES];
    float mean_cos[VSPG_MAX_LOBES];
    float distance[VSPG_MAX_LOBES];
    float vsp_weight[VSPG_MAX_LOBES];
    float pivot[3];
    int num_components;
    int valid;             /* IRegion::valid (spatial/Region.h:80) */
    int pad_[3];
} vspg_mixture;

/* RegionNeighbours<4> (spatial/KNN.h:66-124), rebuilt at load time like
 * Field::deserialize does (field/Field.h:345-348). */
typedef struct vspg_neighbours {
    uint32_t ids[8];       /* ~0u = empty slot */
    float px[8], py[8], pz[8];
    uint32_t size;
    uint32_t pad_[7];
} vspg_neighbours;

typedef struct vspg_field {
    int usable;            /* m_iteration > 0 (field/Field.h:116)                    */
    float bounds_min[3], bounds_max[3]; /* KDTree::m_bounds                          */
    int n_nodes;           const vspg_kdnode *nodes;
    int n_regions;         const vspg_mixture *regions;
    const vspg_neighbours *neighbours;  /* n_regions, or NULL if !stochastic_nn      */
    int stochastic_nn;     /* m_useStochasticNNLookUp                                */
    int is_nn;             /* m_useISNNLookUp (unsupported if set)                   */
} vspg_field;

typedef struct vspg_guiding_desc {
    const vspg_field *surface;  /* NULL = no cache (guiding disabled)                */
    const vspg_field *volume;
    const float *isg_vsp;       /* width*height VSP estimate (layer VSP.Y) or NULL   */
    int isg_width, isg_height;
    const float *tr_rgb;        /* width*height*3 transmittance buffer or NULL       */
    int tr_width, tr_height;
} vspg_guiding_desc;

/* guidedvolpathvspgintegrator.cpp:1260-1322 (same names, same defaults). */
typedef struct vspg_integrator_params {
    int max_depth;              /* maxdepth 5              */
    int min_rr_depth;           /* minrrdepth 1            */
    int use_nee;                /* usenee true             */
    int surface_guiding;        /* surfaceguiding true     */
    int volume_guiding;         /* volumeguiding true      */
    int surface_guiding_ris;    /* surfaceguidingtype "ris" => 1, "mis" => 0 */
    int volume_guiding_ris;     /* volumeguidingtype  "mis" => 0             */
    int vsp_guiding;            /* vspguiding true         */
    int vsp_primary_guiding;    /* vspprimaryguiding true  */
    int vsp_secondary_guiding;  /* vspsecondaryguiding true*/
    float vsp_mis_ratio;        /* vspmisratio 0.5         */
    int vsp_criterion_variance; /* vspcriterion "variance" => 1 (only tags the cache) */
    int vsp_sampling_nds;       /* vspsamplingmethod "resampling" => 0, "nds" => 1    */
    int collision_probability_bias; /* collisionProbabilityBias false (NDS+)          */
    int light_sampler;          /* lightsampler "bvh"      */
    int regularize;             /* regularize false (no effect on diffuse/interface)  */
} vspg_integrator_params;

typedef struct vspg_counters {
    unsigned long long paths;             /* P: camera samples finished                */
    unsigned long long density_queries;   /* Q: Medium::SamplePoint calls              */
    unsigned long long dda_steps;         /* S: DDAMajorantIterator::Next calls        */
    unsigned long long cache_lookups;     /* C: pgl sampling-distribution inits        */
    unsigned long long bvh_nodes;         /* B: BVH nodes visited                      */
    unsigned long long shadow_rays;
    unsigned long long kernel_launches;   /* our kernels launched since create/reset   */
    unsigned long long wave_iterations;
} vspg_counters;

typedef struct vspg_ctx vspg_ctx;

/* ------------------------------------------------------------------ lifecycle */
int vspg_create(vspg_ctx **out, int device_ordinal);
void vspg_destroy(vspg_ctx *ctx);
const char *vspg_last_error(const vspg_ctx *ctx); /* ctx may be NULL: last create error */
const char *vspg_version(void);

void vspg_integrator_defaults(vspg_integrator_params *p);

/* replaces RenderCPU's object construction (cpu/render.cpp:24-160) */
int vspg_set_scene(vspg_ctx *ctx, const vspg_scene_desc *scene);
/* replaces the loadGuidingCache / loadISGBuffer / loadTrBuffer branches (vspg.cpp:117-193) */
int vspg_set_guiding(vspg_ctx *ctx, const vspg_guiding_desc *guiding);
/* replaces GuidedVolPathVSPGIntegrator::Create (vspg.cpp:1260-1322) */
int vspg_set_integrator(vspg_ctx *ctx, const vspg_integrator_params *params);

/* Maximum number of paths resident in HBM at once (wavefront capacity). 0 = default (min(job size, 2^24), ~340 B per path). */
int vspg_set_capacity(vspg_ctx *ctx, int max_paths_in_flight);

/* ------------------------------------------------------------------ rendering */
/* replaces ImageTileIntegrator::Render's wave loop (cpu/integrators.cpp:75-269):
 * renders sample indices [spp_begin, spp_end) with stride spp_stride (sample-pass
 * sharding across ranks: rank r uses spp_begin=r, stride=N) over the pixel window
 * [x0,x1) x [y0,y1), accumulating into the ctx-owned film.  Asynchronous on
 * `cuda_stream` (a cudaStream_t, NULL = default stream) except for the bounded
 * polling of the device-side path counter.  seed = Options->seed. */
int vspg_render(vspg_ctx *ctx, int spp_begin, int spp_end, int spp_stride, int x0, int y0, int x1,
                int y1, int seed, void *cuda_stream);

/* Film (film.h:251-290): 4 doubles per pixel {rgbSum[3], weightSum}. */
int vspg_film_clear(vspg_ctx *ctx);
int vspg_film_device_ptr(vspg_ctx *ctx, double **rgbw, size_t *n_doubles);
int vspg_film_download(vspg_ctx *ctx, double *host_rgbw);           /* H*W*4 doubles     */
int vspg_film_resolve(vspg_ctx *ctx, float *host_rgb);              /* H*W*3 rgb/weight  */
int vspg_counters_get(vspg_ctx *ctx, vspg_counters *out);
int vspg_counters_reset(vspg_ctx *ctx);
/* Per-kernel device time (CUDA events on the render stream), launch counts and the grid traffic counters of the
 * two tracking kernels, accumulated since the last vspg_counters_reset.  Timing is recorded only while enabled
 * (`enable_timing` sets the mode for subsequent vspg_render calls and returns what has been gathered so far). */
enum {
    VSPG_K_GENERATE = 0, VSPG_K_INTERSECT = 1, VSPG_K_MEDIUM = 2, VSPG_K_VOLSCATTER = 3, VSPG_K_SURFACE = 4,
    VSPG_K_SHADE = 5, VSPG_K_SHADOW = 6, VSPG_K_FINISH = 7, VSPG_K_COUNT = 8
};
typedef struct vspg_kernel_stats {
    float ms[VSPG_K_COUNT];                          /* summed device time per kernel                    */
    unsigned long long launches[VSPG_K_COUNT];
    unsigned long long density_queries[VSPG_K_COUNT]; /* Q issued by that kernel (MEDIUM and SHADOW only) */
    unsigned long long dda_steps[VSPG_K_COUNT];       /* S issued by that kernel (MEDIUM and SHADOW only) */
    float ms_total;                                  /* first launch to last kernel end, summed over renders */
} vspg_kernel_stats;
int vspg_kernel_stats_get(vspg_ctx *ctx, int enable_timing, vspg_kernel_stats *out);

/* -------------------------------------------------- host front end (C++ side) */
/* Scene handle produced by the .pbrt front end (parser.cpp / scene.cpp subset).
 * The handle owns every array the descs point at. */
typedef struct vspg_host_scene vspg_host_scene;
int vspg_host_scene_parse(const char *pbrt_path, vspg_host_scene **out, char *err, size_t err_len);
int vspg_host_scene_parse_string(const char *pbrt_text, const char *base_dir, vspg_host_scene **out,
                                 char *err, size_t err_len);
void vspg_host_scene_free(vspg_host_scene *s);
const vspg_scene_desc *vspg_host_scene_desc(const vspg_host_scene *s);
const vspg_guiding_desc *vspg_host_scene_guiding(const vspg_host_scene *s);
const vspg_integrator_params *vspg_host_scene_integrator(const vspg_host_scene *s);
int vspg_host_scene_spp(const vspg_host_scene *s);        /* Sampler "pixelsamples"        */
int vspg_host_scene_seed(const vspg_host_scene *s);       /* Sampler "seed" / Options seed */
const char *vspg_host_scene_outfile(const vspg_host_scene *s); /* Film "filename"           */
/* "storeTrBuffer" true: resolved "trBufferFileName" the caller writes after the render (vspg_tr_store_* +
 * vspg_exr_write_channels, as `vspg-pbrt` does); "" otherwise. */
const char *vspg_host_scene_store_tr_file(const vspg_host_scene *s);
/* "vspguiding" + "vspprimaryguiding" without a loaded ISG buffer: the reference trains the image-space buffer on the fly
 * (vspg.cpp:161-178); train_isg tells the driver to render wave by wave through vspg_isg_store_* (see there), store_isg_file is
 * the "isgBufferFileName" to write afterwards when "storeISGBuffer" is set ("" otherwise). */
int vspg_host_scene_train_isg(const vspg_host_scene *s);
const char *vspg_host_scene_store_isg_file(const vspg_host_scene *s);

/* loaders, usable on their own (replace pglFieldLoadFromFile, ISGB/Tr EXR readers) */
typedef struct vspg_host_field vspg_host_field;
int vspg_host_field_load(const char *pgl_path, vspg_host_field **out, char *err, size_t err_len);
void vspg_host_field_free(vspg_host_field *f);
const vspg_field *vspg_host_field_surface(const vspg_host_field *f);
const vspg_field *vspg_host_field_volume(const vspg_host_field *f);

/* Minimal EXR I/O (util/image.cpp subset: scanline, NONE/ZIP/ZIPS, HALF/FLOAT). */
int vspg_exr_read_channel(const char *path, const char *channel, float **out, int *w, int *h,
                          char *err, size_t err_len); /* free with vspg_free */
int vspg_exr_write_rgb(const char *path, const float *rgb, int w, int h, int as_half);
/* Float ZIP EXR with n named channels, planes[i] = w*h floats of channel names[i] (TrBuffer::Store writes
 * "Transmittance.R/G/B" this way, cpu/trbuffer.h:52-73). */
int vspg_exr_write_channels(const char *path, int w, int h, int n, const char *const *names, const float *const *planes);
void vspg_free(void *p);

/* 16^3 majorant grid exactly as GridMedium / RGBGridMedium ctors build it
 * (media.cpp:218-270, 370-409).  out has maj_res product floats. */
int vspg_build_majorant_grid(const vspg_medium *m, float *out);

/* -------------------------------------------- per-kernel parity hooks (GPU) */
typedef struct vspg_segment {
    float t_min, t_max;
    float sigma_maj[3];
} vspg_segment;

/* DDAMajorantIterator (media.h:140-218) over medium `medium`: for ray i writes up to
 * max_segments segments to out[i*max_segments..] and the count to out_count[i].
 * rays: n*(ox,oy,oz,dx,dy,dz,tMax) floats; tMax is the value AFTER normalisation. */
int vspg_debug_majorant_segments(vspg_ctx *ctx, int medium, int n, const float *rays7,
                                 int max_segments, vspg_segment *out, int *out_count);

/* The same hook for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has no counterpart in the reference.
 * Rays, segments and counts are staged in ONE file-scope device array of fixed size
 * (csrc/vspg_ctx_noalloc.inl, ~5.7 MiB, held by every GPU that loads the library until the process exits) instead of
 * cudaMalloc'd blocks: the call makes no cudaMalloc / cudaFree / allocating library call.  Limits: n <= VSPG_NOALLOC_MAX_RAYS
 * and n * max_segments <= VSPG_NOALLOC_MAX_SEGMENTS, else VSPG_ERR_INVALID with a message.  All calls share the
 * array and it keeps the previous call's contents; the part a call uses is cleared first.  Calls that overlap on one
 * GPU (other threads, other contexts of the same device) race on it: serialise them.  Results are identical to
 * vspg_debug_majorant_segments. */
#define VSPG_NOALLOC_MAX_RAYS 16384
#define VSPG_NOALLOC_MAX_SEGMENTS 262144
int vspg_debug_majorant_segments_noalloc(vspg_ctx *ctx, int medium, int n, const float *rays7,
                                         int max_segments, vspg_segment *out, int *out_count);

/* Medium::SamplePoint (media.h:256-261, 316-348, 417-441) at n render-space points:
 * out7[i] = sigma_a RGB, sigma_s RGB, g. */
int vspg_debug_density(vspg_ctx *ctx, int medium, int n, const float *pos3, float *out7);

/* One SampleDistance call (vspg.cpp:637-1096) up to and including the
 * scatter / pass-through decision, on explicit inputs. */
typedef struct vspg_sd_input {
    float o[3], d[3];
    float t_max;            /* si->tHit */
    int medium;
    int pixel_x, pixel_y;
    int depth;
    int channel;            /* hero channel                                          */
    float vsp;              /* secondary VSP as returned by the guided distribution; ignored at depth 0 */
    float beta[3], r_u[3], r_l[3];
    uint64_t rng_state, rng_inc; /* path sampler (PCG32) state before the call      */
} vspg_sd_input;

typedef struct vspg_sd_output {
    int scattered;          /* a real-scatter vertex was chosen (before phase sampling) */
    int terminated;
    int depth;              /* after the call                                         */
    float p[3];             /* scatter position                                       */
    float sigma_a[3], sigma_s[3]; /* medium properties at p                           */
    float beta[3], r_u[3], r_l[3];
    uint64_t rng_state;     /* path sampler state after the call                      */
    int density_queries;
    int dda_steps;
} vspg_sd_output;
int vspg_debug_sample_distance(vspg_ctx *ctx, int n, const vspg_sd_input *in, vspg_sd_output *out);

/* Field::getRegion (field/Field.h:114-150): leaf id, selected region id, u after draw(). */
int vspg_debug_cache_lookup(vspg_ctx *ctx, int is_surface, int n, const float *pos3, const float *u,
                            int *out_leaf, int *out_region, float *out_u);

/* LightSampler::Sample(ctx, u) and ::PMF(ctx, light) (lightsamplers.h:33-38, 266-359) with the sampler chosen by
 * set_integrator, at n contexts: pos3 = ctx.p(), ns3 = ctx.ns (zero in a medium).  out_light[i] = sampled light or
 * -1, out_pmf[i] its pmf, out_pmf_of[i] = PMF(ctx, light_of[i]). */
int vspg_debug_light_sample(vspg_ctx *ctx, int n, const float *pos3, const float *ns3, const float *u, const int *light_of,
                            int *out_light, float *out_pmf, float *out_pmf_of);

/* storeTrBuffer (cpu/guidedvolpathvspgintegrator.cpp:190-193, 727-728; TrBuffer::AddSample cpu/trbuffer.h:40-45): between
 * begin and end every depth-0 traversal of the resampling routine adds its ratio-tracking transmittance estimate
 * (trRatioEst, RGB) to a per-pixel mean.  download returns the means (h*w*3 floats, 0 where a pixel has no sample)
 * and, if out_count is not NULL, the per-pixel sample counts.  The film resolution is the scene's; begin clears. */
int vspg_tr_store_begin(vspg_ctx *ctx);
int vspg_tr_store_download(vspg_ctx *ctx, float *out_rgb, int *out_count);
int vspg_tr_store_end(vspg_ctx *ctx);

/* storeISGBuffer (cpu/guidedvolpathvspgintegrator.cpp:161-178 config, :230-258 PostProcessWave, :613-622 AddSample;
 * openpgl/imagespace/ImageSpaceGuidingBuffer.h:632-797 update / addSample): the image-space VSP buffer trained on
 * the fly, which is what the reference does when vspguiding + vspprimaryguiding are on and no ISG buffer was loaded.
 *   begin    per-pixel records {surfSpp, volSpp, surfContrib rgb, volContrib rgb} cleared; primary rays read vsp = 0.5
 *            until the first update (the criterion -- contribution or variance -- is the integrator's at this moment:
 *            call after vspg_set_guiding / vspg_set_integrator)
 *   render   ONE sample index per vspg_render call while the store is active (the reference's waves are 1 spp): every
 *            finished path updates its pixel's float running mean exactly as ImageSpaceGuidingBuffer::AddSample does
 *   update   ImageSpaceGuidingBuffer::Update with the denoiser as the identity: VSP per pixel from the records; from
 *            then on primary rays read it.  The reference updates after waves 1, 2, 4, 8, ... and once at the end.
 *   download out_vsp [H*W] (the layer "VSP.Y" of the stored file) and / or out_records8 [H*W*8]
 *   end      back to the loaded ISG buffer (if any)
 * NOT the reference's result where it differs by design: the reference denoises pVol and both contribution buffers
 * with OIDN before forming the ratio; this is the un-denoised estimate (identical to a reference build whose OIDN is the
 * identity, which is what tests/golden pins it against). */
int vspg_isg_store_begin(vspg_ctx *ctx);
int vspg_isg_store_update(vspg_ctx *ctx);
int vspg_isg_store_download(vspg_ctx *ctx, float *out_vsp, float *out_records8);
int vspg_isg_store_end(vspg_ctx *ctx);

/* Integrator::Intersect (cpu/integrators.cpp:341-349 -> BVHAggregate::Intersect, cpu/aggregates.cpp:529-579) at n rays
 * (rays7 = o, d, tMax) followed by the shape's InteractionFromIntersection: out_prim[i] = primitive id counted over
 * quads, then triangles, then spheres (-1 = miss); out10[i] = tHit, p(), pi.Error(), n. */
int vspg_debug_intersect(vspg_ctx *ctx, int n, const float *rays7, int *out_prim, float *out10);

/* Light::SampleLi(ctx, u) (lights.cpp:796-818 -> Shape::Sample(ctx, u), shapes.cpp:1254-1327) followed by
 * Light::PDF_Li(ctx, wi) (lights.cpp:820-824 -> Shape::PDF(ctx, wi), shapes.cpp:1329-1368) along the sampled
 * direction.  ctx9[i] = p(), n, ns of the reference point (n = ns = 0 in a medium); light[i] indexes the lights;
 * u2[i] the 2-D sample.  out12[i] = valid, pdf, wi (3), pLight (3), L (3), PDF_Li(ctx, wi) evaluated at the scene
 * intersection of the ray (ctx, wi) when that hit is the light's own shape, else -1. */
int vspg_debug_sample_li(vspg_ctx *ctx, int n, const float *ctx9, const int *light, const float *u2, float *out12);

/* GuidedPhaseFunction / GuidedBSDF (cpu/guiding.h:57-638) on explicit inputs. */
typedef struct vspg_gs_input {
    float p[3];
    float wo[3];            /* = -ray.d                                              */
    float n[3];             /* surface: shading normal (unused for volume)           */
    float g;                /* volume: HG g; surface: unused                         */
    float reflectance[3];   /* surface: diffuse R                                    */
    float v;                /* cache-lookup sample                                   */
    float u1;               /* surface: Get1D for Sample_f                           */
    float u2[2];            /* Get2D                                                 */
    float wq[3];            /* extra direction at which pdf / VSP are evaluated      */
    uint64_t rng_state, rng_inc; /* sampler, for the extra RIS draws                  */
} vspg_gs_input;
typedef struct vspg_gs_output {
    int ok;                 /* cache init succeeded                                  */
    int region;
    int valid;              /* a sample was produced                                 */
    float wi[3];
    float f[3];             /* surface: f; volume: f[0] = p                          */
    float pdf, scatter_pdf, mis_pdf;
    float pdf_wq;           /* Guided{BSDF,PhaseFunction}::PDF(wo, wq)               */
    float vsp_wq;           /* VolumeScatterProbability(wq)                          */
    uint64_t rng_state;
} vspg_gs_output;
int vspg_debug_guided_sample(vspg_ctx *ctx, int is_surface, int n, const vspg_gs_input *in,
                             vspg_gs_output *out);

/* SampleLd at one path vertex (next-event estimation, reference: cpu/guidedvolpathvspgintegrator.cpp:1136-1252): light
 * pick, Light::SampleLi, guided scatter pdf at the light direction, then the shadow-ray chain through interface
 * surfaces with ratio tracking + Russian roulette and the MIS weighting.  The vertex is described the way the
 * reference's harness (oracle/ref_harness/pbrt_harness.cpp refp_sample_ld) describes it: a surface vertex is the first
 * hit of `ray`; a volume vertex sits at ray.o inside `medium` with incoming direction ray.d.  The sampler is the
 * pixel sampler of (px, py, sample_index) positioned at dimension `dim`; it first yields the cache-lookup draw, then the
 * three light-sampling dimensions.  status: 0 = evaluated, 2 = no (non-interface) surface hit, 3 = bad medium. */
typedef struct vspg_ld_input {
    int is_surface;
    float ray_o[3], ray_d[3];
    int medium;             /* medium the ray travels in (-1 none); for a volume vertex: the vertex's medium */
    int px, py, sample_index, dim, seed;
    float r_p[3];           /* path MIS weight the estimate is formed with (r_u of the caller) */
    int channel;            /* hero channel of the path */
} vspg_ld_input;
typedef struct vspg_ld_output {
    int status;
    float Ld[3];
} vspg_ld_output;
int vspg_debug_sample_ld(vspg_ctx *ctx, int n, const vspg_ld_input *in, vspg_ld_output *out);

/* The path's transcendental functions as the kernels compute them (include/vspg_libm.h, include/vspg_rsqrtss.h; they stand
 * in for the host libm calls of the reference: util/math.h, util/sampling.h, shapes.cpp, embree's rsqrt): out[i] = fn(a[i][, b[i]]),
 * fn = 0 logf, 1 expf, 2 powf(a,b), 3 sinf, 4 cosf, 5 tanf, 6 asinf, 7 acosf, 8 atanf, 9 atan2f(a,b), 10 embree::rsqrt,
 * 11 (float)log((double)a).  The tests compare with the same header compiled for the host (oracle/libportm.so): bit-equal. */
int vspg_debug_libm(vspg_ctx *ctx, int fn, int n, const float *a, const float *b, float *out);

#ifdef __cplusplus
}
#endif
#endif /* VSPG_H */
