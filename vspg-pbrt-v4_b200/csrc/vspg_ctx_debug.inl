// Per-kernel parity hooks of the C ABI (include/vspg.h, "per-kernel parity hooks").  Each one stages the explicit
// inputs in HBM, launches the same device functions the wavefront kernels call, and copies the results back.
// (included by vspg_ctx.cu)

namespace {
struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
    template <typename T> T *as() { return (T *)p; }
};
}  // namespace


extern "C" {

int vspg_debug_majorant_segments(vspg_ctx *ctx, int medium, int n, const float *rays7, int max_segments, vspg_segment *out, int *out_count) {
    if (!ctx || n < 0 || !rays7 || !out || !out_count || max_segments <= 0) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_majorant_segments: set_scene has not been called");
    if (medium < 0 || medium >= ctx->sc.n_media) return fail(ctx, VSPG_ERR_INVALID, "debug_majorant_segments: medium %d out of range", medium);
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf dr, dout, dcnt;
    CK(dr.alloc((size_t)n * 7 * sizeof(float)));
    CK(dout.alloc((size_t)n * max_segments * sizeof(vspg_segment)));
    CK(dcnt.alloc((size_t)n * sizeof(int)));
    CK(cudaMemcpy(dr.p, rays7, (size_t)n * 7 * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemset(dout.p, 0, (size_t)n * max_segments * sizeof(vspg_segment)));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_segments<<<blocks, kBlock>>>(ctx->sc, medium, n, dr.as<float>(), max_segments, dout.as<vspg_segment>(), dcnt.as<int>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout.p, (size_t)n * max_segments * sizeof(vspg_segment), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out_count, dcnt.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_density(vspg_ctx *ctx, int medium, int n, const float *pos3, float *out7) {
    if (!ctx || n < 0 || !pos3 || !out7) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_density: set_scene has not been called");
    if (medium < 0 || medium >= ctx->sc.n_media) return fail(ctx, VSPG_ERR_INVALID, "debug_density: medium %d out of range", medium);
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf dp, dout;
    CK(dp.alloc((size_t)n * 3 * sizeof(float)));
    CK(dout.alloc((size_t)n * 7 * sizeof(float)));
    CK(cudaMemcpy(dp.p, pos3, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_density<<<blocks, kBlock>>>(ctx->sc, medium, n, dp.as<float>(), dout.as<float>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out7, dout.p, (size_t)n * 7 * sizeof(float), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_light_sample(vspg_ctx *ctx, int n, const float *pos3, const float *ns3, const float *u, const int *light_of,
                            int *out_light, float *out_pmf, float *out_pmf_of) {
    if (!ctx || n < 0 || !pos3 || !ns3 || !u || !light_of || !out_light || !out_pmf || !out_pmf_of) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_light_sample: set_scene has not been called");
    for (int i = 0; i < n; i++)
        if (light_of[i] < 0 || light_of[i] >= ctx->sc.n_lights) return fail(ctx, VSPG_ERR_INVALID, "debug_light_sample: item %d: light out of range", i);
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf dp, dn, du, dl, ol, op, oq;
    CK(dp.alloc((size_t)n * 12)); CK(dn.alloc((size_t)n * 12)); CK(du.alloc((size_t)n * 4)); CK(dl.alloc((size_t)n * 4));
    CK(ol.alloc((size_t)n * 4)); CK(op.alloc((size_t)n * 4)); CK(oq.alloc((size_t)n * 4));
    CK(cudaMemcpy(dp.p, pos3, (size_t)n * 12, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dn.p, ns3, (size_t)n * 12, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(du.p, u, (size_t)n * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dl.p, light_of, (size_t)n * 4, cudaMemcpyHostToDevice));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_light_sample<<<blocks, kBlock>>>(ctx->sc, n, dp.as<float>(), dn.as<float>(), du.as<float>(), dl.as<int>(), ol.as<int>(), op.as<float>(), oq.as<float>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out_light, ol.p, (size_t)n * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(out_pmf, op.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out_pmf_of, oq.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_intersect(vspg_ctx *ctx, int n, const float *rays7, int *out_prim, float *out10) {
    if (!ctx || n < 0 || !rays7 || !out_prim || !out10) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_intersect: set_scene has not been called");
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf dr, dp, dout;
    CK(dr.alloc((size_t)n * 28)); CK(dp.alloc((size_t)n * 4)); CK(dout.alloc((size_t)n * 40));
    CK(cudaMemcpy(dr.p, rays7, (size_t)n * 28, cudaMemcpyHostToDevice));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_intersect<<<blocks, kBlock>>>(ctx->sc, n, dr.as<float>(), dp.as<int>(), dout.as<float>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out_prim, dp.p, (size_t)n * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(out10, dout.p, (size_t)n * 40, cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_sample_li(vspg_ctx *ctx, int n, const float *ctx9, const int *light, const float *u2, float *out12) {
    if (!ctx || n < 0 || !ctx9 || !light || !u2 || !out12) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_sample_li: set_scene has not been called");
    for (int i = 0; i < n; i++)
        if (light[i] < 0 || light[i] >= ctx->sc.n_lights) return fail(ctx, VSPG_ERR_INVALID, "debug_sample_li: item %d: light out of range", i);
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf dc, dl, du, dout;
    CK(dc.alloc((size_t)n * 36)); CK(dl.alloc((size_t)n * 4)); CK(du.alloc((size_t)n * 8)); CK(dout.alloc((size_t)n * 48));
    CK(cudaMemcpy(dc.p, ctx9, (size_t)n * 36, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dl.p, light, (size_t)n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(du.p, u2, (size_t)n * 8, cudaMemcpyHostToDevice));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_sample_li<<<blocks, kBlock>>>(ctx->sc, n, dc.as<float>(), dl.as<int>(), du.as<float>(), dout.as<float>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out12, dout.p, (size_t)n * 48, cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_sample_distance(vspg_ctx *ctx, int n, const vspg_sd_input *in, vspg_sd_output *out) {
    if (!ctx || n < 0 || !in || !out) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_sample_distance: set_scene has not been called");
    for (int i = 0; i < n; i++) {
        if (in[i].medium < 0 || in[i].medium >= ctx->sc.n_media) return fail(ctx, VSPG_ERR_INVALID, "debug_sample_distance: item %d: medium out of range", i);
        if (in[i].channel < 0 || in[i].channel > 2) return fail(ctx, VSPG_ERR_INVALID, "debug_sample_distance: item %d: channel out of range", i);
        if (in[i].pixel_x < 0 || in[i].pixel_x >= ctx->film_w || in[i].pixel_y < 0 || in[i].pixel_y >= ctx->film_h)
            return fail(ctx, VSPG_ERR_INVALID, "debug_sample_distance: item %d: pixel outside the film", i);
    }
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf di, dout;
    CK(di.alloc((size_t)n * sizeof(vspg_sd_input)));
    CK(dout.alloc((size_t)n * sizeof(vspg_sd_output)));
    CK(cudaMemcpy(di.p, in, (size_t)n * sizeof(vspg_sd_input), cudaMemcpyHostToDevice));
    CK(cudaMemset(dout.p, 0, (size_t)n * sizeof(vspg_sd_output)));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_sample_distance<<<blocks, kBlock>>>(ctx->sc, n, di.as<vspg_sd_input>(), dout.as<vspg_sd_output>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout.p, (size_t)n * sizeof(vspg_sd_output), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_cache_lookup(vspg_ctx *ctx, int is_surface, int n, const float *pos3, const float *u, int *out_leaf, int *out_region, float *out_u) {
    if (!ctx || n < 0 || !pos3 || !u || !out_leaf || !out_region || !out_u) return VSPG_ERR_INVALID;
    const DField &f = is_surface ? ctx->sc.surf : ctx->sc.vol;
    CK_STATE(f.present, "debug_cache_lookup: no guiding field loaded (set_guiding)");
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf dp, du, dl, dr, dou;
    CK(dp.alloc((size_t)n * 12)); CK(du.alloc((size_t)n * 4)); CK(dl.alloc((size_t)n * 4)); CK(dr.alloc((size_t)n * 4)); CK(dou.alloc((size_t)n * 4));
    CK(cudaMemcpy(dp.p, pos3, (size_t)n * 12, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(du.p, u, (size_t)n * 4, cudaMemcpyHostToDevice));
    int blocks = std::min((n + kBlock - 1) / kBlock, ctx->sm_count * 8);
    k_debug_cache_lookup<<<blocks, kBlock>>>(ctx->sc, is_surface, n, dp.as<float>(), du.as<float>(), dl.as<int>(), dr.as<int>(), dou.as<float>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out_leaf, dl.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out_region, dr.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out_u, dou.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_guided_sample(vspg_ctx *ctx, int is_surface, int n, const vspg_gs_input *in, vspg_gs_output *out) {
    if (!ctx || n < 0 || !in || !out) return VSPG_ERR_INVALID;
    const DField &f = is_surface ? ctx->sc.surf : ctx->sc.vol;
    CK_STATE(f.present, "debug_guided_sample: no guiding field loaded (set_guiding)");
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf di, dout;
    CK(di.alloc((size_t)n * sizeof(vspg_gs_input)));
    CK(dout.alloc((size_t)n * sizeof(vspg_gs_output)));
    CK(cudaMemcpy(di.p, in, (size_t)n * sizeof(vspg_gs_input), cudaMemcpyHostToDevice));
    CK(cudaMemset(dout.p, 0, (size_t)n * sizeof(vspg_gs_output)));
    long long threads = (long long)n * 32;
    int blocks = (int)std::min<long long>((threads + kBlock - 1) / kBlock, (long long)ctx->sm_count * 8);
    k_debug_guided_sample<<<blocks, kBlock>>>(ctx->sc, is_surface, n, di.as<vspg_gs_input>(), dout.as<vspg_gs_output>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout.p, (size_t)n * sizeof(vspg_gs_output), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_sample_ld(vspg_ctx *ctx, int n, const vspg_ld_input *in, vspg_ld_output *out) {
    if (!ctx || n < 0 || !in || !out) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene, "debug_sample_ld: set_scene has not been called");
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf din, dout;
    CK(din.alloc((size_t)n * sizeof(vspg_ld_input)));
    CK(dout.alloc((size_t)n * sizeof(vspg_ld_output)));
    CK(cudaMemcpy(din.p, in, (size_t)n * sizeof(vspg_ld_input), cudaMemcpyHostToDevice));
    int blocks = std::min((n * 32 + kBlock - 1) / kBlock, ctx->sm_count * 8);
    if (ctx->all_gray) k_debug_sample_ld<true><<<blocks, kBlock>>>(ctx->sc, n, din.as<vspg_ld_input>(), dout.as<vspg_ld_output>());
    else k_debug_sample_ld<false><<<blocks, kBlock>>>(ctx->sc, n, din.as<vspg_ld_input>(), dout.as<vspg_ld_output>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout.p, (size_t)n * sizeof(vspg_ld_output), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

int vspg_debug_libm(vspg_ctx *ctx, int fn, int n, const float *a, const float *b, float *out) {
    if (!ctx || n < 0 || !a || !out || fn < 0 || fn > 11) return VSPG_ERR_INVALID;
    if (n == 0) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf da, db, dout;
    CK(da.alloc((size_t)n * sizeof(float))); CK(dout.alloc((size_t)n * sizeof(float)));
    CK(cudaMemcpy(da.p, a, (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
    if (b) { CK(db.alloc((size_t)n * sizeof(float))); CK(cudaMemcpy(db.p, b, (size_t)n * sizeof(float), cudaMemcpyHostToDevice)); }
    k_debug_libm<<<ctx->sm_count * 8, 256>>>(fn, n, da.as<float>(), b ? db.as<float>() : nullptr, dout.as<float>());
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}

}  // extern "C"
