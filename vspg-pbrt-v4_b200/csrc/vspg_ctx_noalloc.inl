// vspg_debug_majorant_segments_noalloc: the majorant-segment hook for callers that cannot allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// vspg_debug_majorant_segments stages its rays, segments and counts in three cudaMalloc'd blocks.  This variant
// stages them in ONE file-scope device array of fixed size instead: no cudaMalloc / cudaFree / allocating library call.
// The array lives in every GPU context that loads libvspg.so until the process exits (kNoAllocScratchBytes, ~5.7 MiB).
// All calls share it and it keeps the previous call's contents: the part a call uses is cleared first (segment slots a
// ray does not fill must read as zero, as in the allocating entry point), and calls that overlap on one GPU race.
// (included by vspg_ctx.cu)

namespace {

constexpr int kNoAllocMaxRays = VSPG_NOALLOC_MAX_RAYS;             // include/vspg.h
constexpr int kNoAllocMaxSegments = VSPG_NOALLOC_MAX_SEGMENTS;     // rays x max_segments of one call
constexpr size_t kNoAllocRaysOff = 0;
constexpr size_t kNoAllocCountOff = kNoAllocRaysOff + (size_t)kNoAllocMaxRays * 7 * sizeof(float);
constexpr size_t kNoAllocSegOff = (kNoAllocCountOff + (size_t)kNoAllocMaxRays * sizeof(int) + 255) / 256 * 256;
constexpr size_t kNoAllocScratchBytes = kNoAllocSegOff + (size_t)kNoAllocMaxSegments * sizeof(vspg_segment);

// In host code this name is only a handle for the symbol, not a device address: the kernel gets the address
// cudaGetSymbolAddress returns for the current device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// Device address of the scratch array per device, resolved once per device with that device current.
static std::mutex g_noalloc_mutex;
static unsigned char *g_noalloc_ptr[64] = {};

static cudaError_t noalloc_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        cudaError_t e = cudaSetDevice(device);
        if (e != cudaSuccess) return e;
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = (unsigned char *)p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" {

int vspg_debug_majorant_segments_noalloc(vspg_ctx *ctx, int medium, int n, const float *rays7, int max_segments, vspg_segment *out, int *out_count) {
    if (!ctx || n < 0 || !rays7 || !out || !out_count || max_segments <= 0) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_majorant_segments_noalloc: set_scene has not been called");
    if (medium < 0 || medium >= ctx->sc.n_media) return fail(ctx, VSPG_ERR_INVALID, "debug_majorant_segments_noalloc: medium %d out of range", medium);
    if (n > kNoAllocMaxRays || (long long)n * max_segments > kNoAllocMaxSegments)
        return fail(ctx, VSPG_ERR_INVALID, "debug_majorant_segments_noalloc: %d rays x %d segments exceeds the fixed scratch (at most %d rays and %d segments per call); "
                    "use vspg_debug_majorant_segments or split the call", n, max_segments, kNoAllocMaxRays, kNoAllocMaxSegments);
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    unsigned char *scratch = nullptr;
    CK(noalloc_scratch(ctx->device, &scratch));
    float *dRays = (float *)(scratch + kNoAllocRaysOff);
    int *dCount = (int *)(scratch + kNoAllocCountOff);
    vspg_segment *dSeg = (vspg_segment *)(scratch + kNoAllocSegOff);
    const size_t segBytes = (size_t)n * max_segments * sizeof(vspg_segment);
    CK(cudaMemcpy(dRays, rays7, (size_t)n * 7 * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemset(dSeg, 0, segBytes));                  // the previous call's segments are still in there
    CK(cudaMemset(dCount, 0, (size_t)n * sizeof(int)));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_segments<<<blocks, kBlock>>>(ctx->sc, medium, n, dRays, max_segments, dSeg, dCount);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dSeg, segBytes, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out_count, dCount, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

}  // extern "C"
