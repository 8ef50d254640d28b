// Wavefront allocation, the render loop (ImageTileIntegrator::Render re-expressed), film and counters.
// (included by vspg_ctx.cu)

enum { Q_A = 0, Q_B = 1, Q_MEDIUM = 2, Q_SURFACE = 3, Q_SHADE = 4, Q_VOLSCATTER = 5, Q_SHADOW = 6, Q_DONE = 7, Q_MEDIUM_RS = 8, Q_VOL8 = 9, Q_VOL16 = 10, Q_VOL32 = 11, Q_SH8 = 12, Q_SH16 = 13, Q_SH32 = 14, Q_COUNT = 15,
       Q_CURSORS = 4, Q_INTS = Q_COUNT + Q_CURSORS };  // + per-kernel dynamic work cursors

// Everything the captured wave graph depends on; compared bytewise against the current values before each replay.
struct WaveGraphKey { DScene sc; Wavefront W; int vgrid, sgrid, sblock, rs_refill, rs_occ, nGridWin, all_gray, any_nvdb, any_emissive, any_grid_medium, sort_mode, prim_kinds, shadow_persist; cudaAccessPolicyWindow win; };

static int ensure_wavefront(vspg_ctx *ctx, unsigned long long total) {
    // Default: up to 16.8 M paths in flight (5.7 GB of the 180 GB).  Measured on C2 (profiles/r01z_summary.md, r02b/c): 2 M ->
    // 131.9, 4 M -> 137.1, 8 M -> 142.9, 16.8 M -> 147.0 Mpaths/s -- fewer, larger and more coherent waves.
    int want = ctx->requested_capacity > 0 ? ctx->requested_capacity : (1 << 24);
    if ((unsigned long long)want > total) want = (int)total;
    want = std::max(want, 1024);
    if (ctx->capacity >= want && (long long)ctx->capacity <= 4ll * want) return VSPG_OK;
    free_all(ctx->wf_allocs);
    ctx->capacity = 0;
    const size_t n = (size_t)want;
    auto alloc = [&](void **p, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) ctx->wf_allocs.push_back(*p);
        return e;
    };
    PathSoA &P = ctx->W.P;
    float4 **f4[] = {&P.ev, &P.prev_o, &P.sh_f};
    for (auto pp : f4) CK(alloc((void **)pp, n * sizeof(float4)));
    // 32-byte records shared by two streams (dscene.cuh F4S)
    float4 **pairs[][2] = {{&P.o_dx.base, &P.d_med.base}, {&P.beta_ru.base, &P.ru_rl.base}, {&P.rl_L.base, &P.misc.base}, {&P.misc2.base, &P.prev_ns.base},
                           {&P.sh_o.base, &P.sh_d.base}, {&P.sh_c.base, &P.sh_rl.base}, {&P.sh_ru.base, &P.sh_pl.base}, {&P.rng.base, &P.hit.base}, {&P.sh_ple.base, &P.lk.base}};
    for (auto &pr : pairs) { CK(alloc((void **)pr[0], 2 * n * sizeof(float4))); *pr[1] = *pr[0]; }
    CK(alloc((void **)&P.rs_pre, n * sizeof(float2)));
    CK(alloc((void **)&ctx->qcounts, Q_INTS * sizeof(int)));
    int *items = nullptr;
    CK(alloc((void **)&items, Q_COUNT * n * sizeof(int)));
    Queue *qs[Q_COUNT] = {&ctx->W.active, &ctx->W.next, &ctx->W.medium, &ctx->W.surface, &ctx->W.shade, &ctx->W.volscatter, &ctx->W.shadow, &ctx->W.done, &ctx->W.medium_rs,
                          &ctx->W.vol8, &ctx->W.vol16, &ctx->W.vol32, &ctx->W.sh8, &ctx->W.sh16, &ctx->W.sh32};
    for (int i = 0; i < Q_COUNT; i++) { qs[i]->items = items + (size_t)i * n; qs[i]->count = ctx->qcounts + i; }
    ctx->phys_q[0] = ctx->W.active; ctx->phys_q[1] = ctx->W.next;
    CK(alloc((void **)&ctx->W.next_sample, sizeof(unsigned long long)));
    CK(alloc((void **)&ctx->bin_keys, n * sizeof(int)));
    CK(alloc((void **)&ctx->sorted_medium, n * sizeof(int)));
    CK(alloc((void **)&ctx->sorted_shadow, n * sizeof(int)));
    CK(alloc((void **)&ctx->bin_hist, kBins * sizeof(int)));
    CK(alloc((void **)&ctx->bin_offsets, kBins * sizeof(int)));
    CK(cudaMemset(ctx->bin_hist, 0, kBins * sizeof(int)));
    CK(alloc((void **)&ctx->d_rv, sizeof(RenderVars)));
    CK(alloc((void **)&ctx->d_iters, sizeof(int)));
    ctx->W.rv = ctx->d_rv;
    if (!ctx->h_pinned) CK(cudaMallocHost((void **)&ctx->h_pinned, 128));
    ctx->capacity = want;
    ctx->W.capacity = want;
    return VSPG_OK;
}

extern "C" {

int vspg_render(vspg_ctx *ctx, int spp_begin, int spp_end, int spp_stride, int x0, int y0, int x1, int y1, int seed, void *cuda_stream) {
    if (!ctx) return VSPG_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, VSPG_ERR_STATE, "render: set_scene has not been called");
    CK(cudaSetDevice(ctx->device));
    if (spp_stride < 1) spp_stride = 1;
    if (x0 < 0 || y0 < 0 || x1 > ctx->film_w || y1 > ctx->film_h || x0 >= x1 || y0 >= y1) return fail(ctx, VSPG_ERR_INVALID, "render: bad pixel window");
    if (spp_begin < 0 || spp_end < spp_begin) return fail(ctx, VSPG_ERR_INVALID, "render: bad sample range");
    const DScene &scRef = ctx->sc;
    // per-pixel inputs are indexed by the film's pixels: a set_scene with another film size after set_guiding must not
    // leave the kernels reading past them (set_guiding checks the same when it runs after set_scene)
    if (scRef.isg_vsp && (scRef.isg_w != scRef.cam.width || scRef.isg_h != scRef.cam.height))
        return fail(ctx, VSPG_ERR_STATE, "render: the ISG buffer is %dx%d but the film is %dx%d (call set_guiding again after set_scene)", scRef.isg_w, scRef.isg_h, scRef.cam.width, scRef.cam.height);
    if (scRef.tr_rgb && (scRef.tr_w != scRef.cam.width || scRef.tr_h != scRef.cam.height))
        return fail(ctx, VSPG_ERR_STATE, "render: the Tr buffer is %dx%d but the film is %dx%d (call set_guiding again after set_scene)", scRef.tr_w, scRef.tr_h, scRef.cam.width, scRef.cam.height);
    if (scRef.ip.vsp_sampling_nds && scRef.ip.collision_probability_bias && !scRef.tr_rgb) {
        // the reference falls back to *training* a Tr buffer here; this build has no training pass
        return fail(ctx, VSPG_ERR_STATE, "render: collisionProbabilityBias needs a loaded transmittance buffer (loadTrBuffer)");
    }
    int nspp = (spp_end - spp_begin + spp_stride - 1) / spp_stride;
    unsigned long long npix = (unsigned long long)(x1 - x0) * (y1 - y0);
    unsigned long long total = npix * (unsigned long long)nspp;
    if (ctx->isg_active && nspp != 1)   // the ISG records are float running means in wave order: one sample per pixel and render
        return fail(ctx, VSPG_ERR_STATE, "render: while the ISG store is active render one sample index per call (got %d); see vspg_isg_store_begin", nspp);
    if (total == 0) return VSPG_OK;
    int rc = ensure_wavefront(ctx, total);
    if (rc != VSPG_OK) return rc;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    Wavefront &W = ctx->W;
    W.film = ctx->film; W.counters = ctx->counters;
    const DScene sc = ctx->sc;
    Queue qa = ctx->phys_q[0], qb = ctx->phys_q[1];   // physical queues 0 and 1; roles swap each iteration
    // what changes from render to render goes through device memory (RenderVars): kernel arguments stay the same, so
    // the captured wave graph below can be replayed
    {
        RenderVars *hv = reinterpret_cast<RenderVars *>(ctx->h_pinned + 4);
        hv->total_samples = total; hv->x0 = x0; hv->y0 = y0; hv->x1 = x1; hv->y1 = y1;
        hv->spp_begin = spp_begin; hv->spp_stride = spp_stride; hv->n_spp = nspp; hv->seed = seed;
        // band-major sample order when the render is larger than the wavefront: bands of ~1/4 of the paths in flight
        hv->band_rows = 0;
        if (ctx->band_mode != 0 && total > (unsigned long long)ctx->capacity) {
            long long rows = ctx->band_mode > 0 ? ctx->band_mode : (long long)(ctx->capacity / 4) / ((long long)(x1 - x0) * nspp);
            hv->band_rows = (int)std::max<long long>(1, std::min<long long>(rows, y1 - y0));
        }
        CK(cudaMemcpyAsync(ctx->d_rv, hv, sizeof(RenderVars), cudaMemcpyHostToDevice, st));
    }
    CK(cudaMemsetAsync(ctx->qcounts, 0, Q_INTS * sizeof(int), st));
    CK(cudaMemsetAsync(W.next_sample, 0, sizeof(unsigned long long), st));
    // lane-group scatter kernels (kernels.cuh GroupIter): persistent grids of exactly the CTAs one wave of the device holds
    const int sblock = ctx->scatter_block;
    if (ctx->scatter_ctas_block != sblock) {
        int ov = 0, os = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ov, k_volscatter<8>, sblock, 0));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&os, k_shade<8, kKindsAll>, sblock, 0));
        ctx->volscatter_ctas = std::max(1, ov); ctx->shade_ctas = std::max(1, os); ctx->scatter_ctas_block = sblock;
        if (const char *e = getenv("VSPG_SCATTER_CTAS")) { int v = atoi(e); if (v >= 1 && v <= 32) ctx->volscatter_ctas = ctx->shade_ctas = v; }
    }
    const int vgrid = ctx->sm_count * ctx->volscatter_ctas, sgrid = ctx->sm_count * ctx->shade_ctas;
    const bool guidedVol = sc.vol.present, guidedSurf = sc.surf.present;
    int nslots = (int)std::min<unsigned long long>((unsigned long long)ctx->capacity, total);
    W.active = qa; W.next = qb;
    const bool timing = ctx->timing;
    if (timing && !ctx->ev[0])
        for (auto &e : ctx->ev) CK(cudaEventCreate(&e));
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    if (timing) { CK(cudaEventCreate(&evStart)); CK(cudaEventCreate(&evStop)); CK(cudaEventRecord(evStart, st)); }
    // Grid-stride kernels get exactly one resident wave: sm_count x (CTAs of kBlock threads that fit one SM).  A fixed 8
    // CTAs/SM left k_surface (5 resident) and k_intersect / k_lookup (6 resident) with a partial second wave.
    if (!ctx->grids_ready) {
        auto occ = [&](auto kernel, int *out) -> cudaError_t {
            int o = 0; cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kernel, kBlock, 0);
            *out = ctx->sm_count * std::max(1, o); return e;
        };
        CK(occ(k_generate, &ctx->g_generate)); CK(occ(k_intersect<kKindsAll>, &ctx->g_intersect)); CK(occ(k_medium, &ctx->g_medium)); CK(occ(k_medium_hom, &ctx->g_medium_hom));
        CK(occ(k_lookup<false, kKindsAll>, &ctx->g_lookup0)); CK(occ(k_lookup<true, kKindsAll>, &ctx->g_lookup1)); CK(occ(k_surface<kKindsAll>, &ctx->g_surface));
        CK(occ(k_volscatter<0>, &ctx->g_volscatter0)); CK(occ(k_shade<0, kKindsAll>, &ctx->g_shade0)); CK(occ(k_finish, &ctx->g_finish));
        CK(occ(k_bin_count, &ctx->g_bin));
        CK(occ(k_shadow<true>, &ctx->g_shadow_gray)); CK(occ(k_shadow<false>, &ctx->g_shadow_rgb));
        ctx->g_surface = ctx->sm_count * 8;   // memory-latency bound with very uneven items: measured better oversubscribed (16.8 vs 18.8 ms)
        ctx->grids_ready = true;
    }
    k_generate<<<ctx->g_generate, kBlock, 0, st>>>(sc, W, nslots);
    ctx->launches++; ctx->launches_kernel[VSPG_K_GENERATE]++;
    if (timing) {
        CK(cudaEventRecord(ctx->ev[0], st)); CK(cudaEventSynchronize(ctx->ev[0]));
        float ms = 0; CK(cudaEventElapsedTime(&ms, evStart, ctx->ev[0])); ctx->ms_kernel[VSPG_K_GENERATE] += ms;
    }
    // The tracking kernels gather from the cell-record grid at random while the path state streams through L2: give the
    // grid a persisting access-policy window (per launch, so it does not depend on which stream the caller passed) and
    // let everything else those kernels touch be streaming.  Opt-in (VSPG_L2_PERSIST=1 or =<MiB>): a net loss on C2, see vspg_create.
    cudaLaunchAttribute gridWin[1];
    int nGridWin = 0;
    if (ctx->cells_bytes && ctx->l2_persist_bytes) {
        gridWin[0].id = cudaLaunchAttributeAccessPolicyWindow;
        cudaAccessPolicyWindow &w = gridWin[0].val.accessPolicyWindow;
        w.base_ptr = (void *)ctx->cells_ptr;
        w.num_bytes = std::min(ctx->cells_bytes, ctx->l2_window_max);
        w.hitRatio = (float)std::min(1.0, (double)ctx->l2_persist_bytes / (double)w.num_bytes);
        w.hitProp = cudaAccessPropertyPersisting; w.missProp = cudaAccessPropertyStreaming;
        nGridWin = 1;
    }
    // One wave iteration = the stage kernels below, in stream order, over the queues of `Wh`; afterwards every counter but
    // the new active queue's (`keep`) is cleared.  `timed` brackets each stage with events (stage-split measurements only).
    const int sortPasses = ctx->bin_ok ? (ctx->sort_mode & 1) + ((ctx->sort_mode >> 1) & 1) : 0;
    const int perIter = 9 + (guidedVol ? 3 : 0) + (guidedSurf ? 3 : 0) + 3 * sortPasses;
    auto enqueue_iteration = [&](const Wavefront &Wh0, int keep, cudaStream_t s, bool timed) -> int {
        auto launch_tracking = [&](auto kernel, int nblocks, auto... args) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(nblocks); cfg.blockDim = dim3(kBlock); cfg.dynamicSmemBytes = 0; cfg.stream = s;
            cfg.attrs = gridWin; cfg.numAttrs = nGridWin;
            return cudaLaunchKernelEx(&cfg, kernel, args...);
        };
#define VSPG_STAGE(K, launch) do { launch; if (timed) { ctx->launches_kernel[K]++; CK(cudaEventRecord(ctx->ev[K], s)); } } while (0)
        if (timed) CK(cudaEventRecord(ctx->ev[0], s));
        const bool quadsOnly = ctx->prim_kinds == kKindsQuads;   // kernels without the triangle / sphere branches (dshapes.cuh KINDS)
        if (quadsOnly) VSPG_STAGE(VSPG_K_INTERSECT, (k_intersect<kKindsQuads><<<ctx->g_intersect, kBlock, 0, s>>>(sc, Wh0)));
        else VSPG_STAGE(VSPG_K_INTERSECT, (k_intersect<kKindsAll><<<ctx->g_intersect, kBlock, 0, s>>>(sc, Wh0)));
        const int sortMode = ctx->bin_ok ? ctx->sort_mode : 0;
        auto sort_queue = [&](const Queue &q, int which, int *sorted) {
            k_bin_count<<<ctx->g_bin, kBlock, 0, s>>>(sc, Wh0, q, which, ctx->bin_keys, ctx->bin_hist);
            k_bin_scan<<<1, 1024, 0, s>>>(ctx->bin_hist, ctx->bin_offsets);
            k_bin_scatter<<<ctx->g_bin, kBlock, 0, s>>>(q, ctx->bin_keys, ctx->bin_offsets, sorted);
        };
        Wavefront Wh = Wh0;
        if (sortMode & 1) { sort_queue(Wh0.medium_rs, 0, ctx->sorted_medium); Wh.medium_rs.items = ctx->sorted_medium; }
        if (ctx->all_gray && ctx->any_nvdb) CK(launch_tracking(k_medium_rs<true, 4, true>, ctx->sm_count * 4, sc, Wh, ctx->qcounts + Q_COUNT, ctx->rs_refill));
        else if (ctx->all_gray && ctx->rs_occ == 5) CK(launch_tracking(k_medium_rs<true, VSPG_RS_MINB>, ctx->sm_count * VSPG_RS_MINB, sc, Wh, ctx->qcounts + Q_COUNT, ctx->rs_refill));
        else if (ctx->all_gray) CK(launch_tracking(k_medium_rs<true, 4>, ctx->sm_count * 4, sc, Wh, ctx->qcounts + Q_COUNT, ctx->rs_refill));
        else CK(launch_tracking(k_medium_rs<false, 4>, ctx->sm_count * 4, sc, Wh, ctx->qcounts + Q_COUNT, ctx->rs_refill));
        // full-occupancy builds hide the gathers of the walk over a grid; homogeneous-only scenes have none (kernels.cuh)
        if (ctx->any_emissive) { VSPG_STAGE(VSPG_K_MEDIUM, (k_medium_emit<<<ctx->g_medium, kBlock, 0, s>>>(sc, Wh))); }
        else if (ctx->any_grid_medium) { VSPG_STAGE(VSPG_K_MEDIUM, (k_medium<<<ctx->g_medium, kBlock, 0, s>>>(sc, Wh))); }
        else { VSPG_STAGE(VSPG_K_MEDIUM, (k_medium_hom<<<ctx->g_medium_hom, kBlock, 0, s>>>(sc, Wh))); }
        if (guidedVol) {
            if (quadsOnly) k_lookup<false, kKindsQuads><<<ctx->g_lookup0, kBlock, 0, s>>>(sc, Wh, Wh.volscatter, Wh.vol8, Wh.vol16, Wh.vol32);
            else k_lookup<false, kKindsAll><<<ctx->g_lookup0, kBlock, 0, s>>>(sc, Wh, Wh.volscatter, Wh.vol8, Wh.vol16, Wh.vol32);
            k_volscatter<8><<<vgrid, sblock, 0, s>>>(sc, Wh, Wh.vol8);
            k_volscatter<16><<<vgrid, sblock, 0, s>>>(sc, Wh, Wh.vol16);
            VSPG_STAGE(VSPG_K_VOLSCATTER, (k_volscatter<32><<<vgrid, sblock, 0, s>>>(sc, Wh, Wh.vol32)));
        } else VSPG_STAGE(VSPG_K_VOLSCATTER, (k_volscatter<0><<<ctx->g_volscatter0, kBlock, 0, s>>>(sc, Wh, Wh.volscatter)));
        if (quadsOnly) VSPG_STAGE(VSPG_K_SURFACE, (k_surface<kKindsQuads><<<ctx->g_surface, kBlock, 0, s>>>(sc, Wh)));
        else VSPG_STAGE(VSPG_K_SURFACE, (k_surface<kKindsAll><<<ctx->g_surface, kBlock, 0, s>>>(sc, Wh)));
        if (guidedSurf) {
            if (quadsOnly) {
                k_lookup<true, kKindsQuads><<<ctx->g_lookup1, kBlock, 0, s>>>(sc, Wh, Wh.shade, Wh.sh8, Wh.sh16, Wh.sh32);
                k_shade<8, kKindsQuads><<<sgrid, sblock, 0, s>>>(sc, Wh, Wh.sh8);
                k_shade<16, kKindsQuads><<<sgrid, sblock, 0, s>>>(sc, Wh, Wh.sh16);
                VSPG_STAGE(VSPG_K_SHADE, (k_shade<32, kKindsQuads><<<sgrid, sblock, 0, s>>>(sc, Wh, Wh.sh32)));
            } else {
                k_lookup<true, kKindsAll><<<ctx->g_lookup1, kBlock, 0, s>>>(sc, Wh, Wh.shade, Wh.sh8, Wh.sh16, Wh.sh32);
                k_shade<8, kKindsAll><<<sgrid, sblock, 0, s>>>(sc, Wh, Wh.sh8);
                k_shade<16, kKindsAll><<<sgrid, sblock, 0, s>>>(sc, Wh, Wh.sh16);
                VSPG_STAGE(VSPG_K_SHADE, (k_shade<32, kKindsAll><<<sgrid, sblock, 0, s>>>(sc, Wh, Wh.sh32)));
            }
        } else VSPG_STAGE(VSPG_K_SHADE, (k_shade<0><<<ctx->g_shade0, kBlock, 0, s>>>(sc, Wh, Wh.shade)));
        if (sortMode & 2) { sort_queue(Wh0.shadow, 1, ctx->sorted_shadow); Wh.shadow.items = ctx->sorted_shadow; }
        if (ctx->shadow_persist) {   // persistent lanes; work cursor = the second dynamic cursor (cleared with the queue counters)
            int *cur = ctx->qcounts + Q_COUNT + 1;
            const int g = ctx->sm_count * VSPG_SHADOWP_MINB;
            if (ctx->all_gray && quadsOnly) VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow_p<true, kKindsQuads>, g, sc, Wh, cur, ctx->shadow_refill)));
            else if (ctx->all_gray) VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow_p<true, kKindsAll>, g, sc, Wh, cur, ctx->shadow_refill)));
            else if (quadsOnly) VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow_p<false, kKindsQuads>, g, sc, Wh, cur, ctx->shadow_refill)));
            else VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow_p<false, kKindsAll>, g, sc, Wh, cur, ctx->shadow_refill)));
        } else
        if (ctx->all_gray && quadsOnly) VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow<true, kKindsQuads>, ctx->g_shadow_gray, sc, Wh)));
        else if (ctx->all_gray) VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow<true, kKindsAll>, ctx->g_shadow_gray, sc, Wh)));
        else if (quadsOnly) VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow<false, kKindsQuads>, ctx->g_shadow_rgb, sc, Wh)));
        else VSPG_STAGE(VSPG_K_SHADOW, CK(launch_tracking(k_shadow<false, kKindsAll>, ctx->g_shadow_rgb, sc, Wh)));
        VSPG_STAGE(VSPG_K_FINISH, (k_finish<<<ctx->g_finish, kBlock, 0, s>>>(sc, Wh)));
        k_reset_counts<<<1, 32, 0, s>>>(ctx->qcounts, Q_INTS, keep);
#undef VSPG_STAGE
        return VSPG_OK;
    };
    Wavefront Wab = W, Wba = W;            // iteration with (active, next) = (A, B) and the one with the roles swapped
    Wba.active = qb; Wba.next = qa;
    unsigned iterations = 0;
    if (timing || !ctx->use_graph) {
        // host-driven loop: one stream sync + counter read per iteration (stage-split measurements, VSPG_GRAPH=0)
        for (int cur = 0;; cur ^= 1) {
            int rc2 = enqueue_iteration(cur ? Wba : Wab, cur ^ 1, st, timing);
            if (rc2 != VSPG_OK) return rc2;
            iterations++;
            CK(cudaMemcpyAsync(ctx->h_pinned, ctx->phys_q[cur ^ 1].count, sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (timing)
                for (int k = VSPG_K_INTERSECT; k < VSPG_K_COUNT; k++) {
                    float ms = 0; CK(cudaEventElapsedTime(&ms, ctx->ev[k == VSPG_K_INTERSECT ? 0 : k - 1], ctx->ev[k]));
                    ctx->ms_kernel[k] += ms;
                }
            if (ctx->h_pinned[0] == 0) break;
        }
    } else {
        // Device-driven loop: a CUDA graph whose WHILE node repeats {iteration A->B, iteration B->A} until the active queue
        // is empty -- no host round trip between waves.  The graph only depends on kernel arguments (scene, queues, grids),
        // all of which are compared below; it is rebuilt when any of them changed.
        WaveGraphKey key;
        memset(&key, 0, sizeof key);
        key.sc = sc; key.W = Wab; key.vgrid = vgrid; key.sgrid = sgrid; key.sblock = sblock; key.rs_refill = ctx->rs_refill; key.rs_occ = ctx->rs_occ; key.all_gray = ctx->all_gray; key.any_emissive = ctx->any_emissive; key.any_grid_medium = ctx->any_grid_medium; key.shadow_persist = ctx->shadow_persist * 100 + ctx->shadow_refill; key.prim_kinds = ctx->prim_kinds; key.sort_mode = ctx->bin_ok ? ctx->sort_mode : 0; key.any_nvdb = ctx->any_nvdb;
        key.nGridWin = nGridWin; if (nGridWin) key.win = gridWin[0].val.accessPolicyWindow;
        if (!ctx->wave_exec || memcmp(&key, ctx->wave_key, sizeof key) != 0) {
            if (ctx->wave_exec) { cudaGraphExecDestroy(ctx->wave_exec); ctx->wave_exec = nullptr; }
            if (ctx->wave_graph) { cudaGraphDestroy(ctx->wave_graph); ctx->wave_graph = nullptr; }
            if (!ctx->cap_stream) CK(cudaStreamCreateWithFlags(&ctx->cap_stream, cudaStreamNonBlocking));
            CK(cudaGraphCreate(&ctx->wave_graph, 0));
            cudaGraphConditionalHandle handle;
            CK(cudaGraphConditionalHandleCreate(&handle, ctx->wave_graph, 1, cudaGraphCondAssignDefault));
            cudaGraphNodeParams np = {};
            np.type = cudaGraphNodeTypeConditional;
            np.conditional.handle = handle; np.conditional.type = cudaGraphCondTypeWhile; np.conditional.size = 1;
            cudaGraphNode_t whileNode;
            CK(cudaGraphAddNode(&whileNode, ctx->wave_graph, nullptr, 0, &np));
            cudaGraph_t body = np.conditional.phGraph_out[0];
            CK(cudaStreamBeginCaptureToGraph(ctx->cap_stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
            int rc2 = enqueue_iteration(Wab, 1, ctx->cap_stream, false);
            if (rc2 == VSPG_OK) rc2 = enqueue_iteration(Wba, 0, ctx->cap_stream, false);
            if (rc2 == VSPG_OK) k_wave_condition<<<1, 1, 0, ctx->cap_stream>>>(handle, qa.count, ctx->d_iters);
            cudaGraph_t captured = nullptr;
            cudaError_t ce = cudaStreamEndCapture(ctx->cap_stream, &captured);
            if (rc2 != VSPG_OK) return rc2;
            CK(ce);
            CK(cudaGraphInstantiate(&ctx->wave_exec, ctx->wave_graph, 0));
            if (!ctx->wave_key) ctx->wave_key = (WaveGraphKey *)malloc(sizeof(WaveGraphKey));
            memcpy(ctx->wave_key, &key, sizeof key);
            ctx->graph_builds++;
            if (getenv("VSPG_VERBOSE")) fprintf(stderr, "vspg: wave graph built (#%u)\n", ctx->graph_builds);
        }
        CK(cudaMemsetAsync(ctx->d_iters, 0, sizeof(int), st));
        CK(cudaGraphLaunch(ctx->wave_exec, st));
        CK(cudaMemcpyAsync(ctx->h_pinned, ctx->d_iters, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        iterations = 2u * (unsigned)ctx->h_pinned[0];
    }
    ctx->iterations += iterations;
    ctx->launches += (unsigned long long)iterations * perIter;
    if (!timing) {   // launch bookkeeping per stage (the timed path counts inside VSPG_STAGE)
        const unsigned it = iterations;
        ctx->launches_kernel[VSPG_K_INTERSECT] += it; ctx->launches_kernel[VSPG_K_MEDIUM] += it; ctx->launches_kernel[VSPG_K_VOLSCATTER] += it;
        ctx->launches_kernel[VSPG_K_SURFACE] += it; ctx->launches_kernel[VSPG_K_SHADE] += it; ctx->launches_kernel[VSPG_K_SHADOW] += it; ctx->launches_kernel[VSPG_K_FINISH] += it;
    }
    if (timing) {
        CK(cudaEventRecord(evStop, st)); CK(cudaEventSynchronize(evStop));
        float ms = 0; CK(cudaEventElapsedTime(&ms, evStart, evStop));
        ctx->ms_total += ms;
        cudaEventDestroy(evStart); cudaEventDestroy(evStop);
    }
    CK(cudaGetLastError());
    return VSPG_OK;
}

int vspg_film_clear(vspg_ctx *ctx) {
    if (!ctx || !ctx->film) return VSPG_ERR_STATE;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemset(ctx->film, 0, (size_t)ctx->film_w * ctx->film_h * 4 * sizeof(double)));
    return VSPG_OK;
}
int vspg_tr_store_begin(vspg_ctx *ctx) {
    if (!ctx) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene && ctx->film, "tr_store_begin: set_scene has not been called");
    CK(cudaSetDevice(ctx->device));
    const size_t bytes = (size_t)ctx->film_w * ctx->film_h * 4 * sizeof(double);
    if (!ctx->tr_acc || ctx->tr_acc_pixels != (size_t)ctx->film_w * ctx->film_h) {
        if (ctx->tr_acc) cudaFree(ctx->tr_acc);
        ctx->tr_acc = nullptr;
        CK(cudaMalloc((void **)&ctx->tr_acc, bytes));
        ctx->tr_acc_pixels = (size_t)ctx->film_w * ctx->film_h;
    }
    CK(cudaMemset(ctx->tr_acc, 0, bytes));
    ctx->sc.tr_acc = ctx->tr_acc;
    return VSPG_OK;
}
int vspg_tr_store_download(vspg_ctx *ctx, float *out_rgb, int *out_count) {
    if (!ctx || !out_rgb) return VSPG_ERR_INVALID;
    CK_STATE(ctx->sc.tr_acc != nullptr, "tr_store_download: tr_store_begin has not been called");
    CK(cudaSetDevice(ctx->device));
    const size_t n = ctx->tr_acc_pixels;
    std::vector<double> acc(4 * n);
    CK(cudaMemcpy(acc.data(), ctx->tr_acc, acc.size() * sizeof(double), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) {
        double c = acc[4 * i + 3];
        for (int k = 0; k < 3; k++) out_rgb[3 * i + k] = c > 0 ? (float)(acc[4 * i + k] / c) : 0.f;
        if (out_count) out_count[i] = (int)c;
    }
    return VSPG_OK;
}
int vspg_tr_store_end(vspg_ctx *ctx) {
    if (!ctx) return VSPG_ERR_INVALID;
    ctx->sc.tr_acc = nullptr;
    return VSPG_OK;
}
// ---- storeISGBuffer: the image-space VSP buffer trained on the fly, as the reference does when vspguiding and
// vspprimaryguiding are on and no ISG buffer was loaded (vspg.cpp:161-178, 230-258, 613-622)
int vspg_isg_store_begin(vspg_ctx *ctx) {
    if (!ctx) return VSPG_ERR_INVALID;
    CK_STATE(ctx->have_scene && ctx->film, "isg_store_begin: set_scene has not been called");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = (size_t)ctx->film_w * ctx->film_h;
    if (!ctx->isg_rec || ctx->isg_pixels != npix) {
        if (ctx->isg_rec) cudaFree(ctx->isg_rec);
        if (ctx->isg_trained) cudaFree(ctx->isg_trained);
        ctx->isg_rec = nullptr; ctx->isg_trained = nullptr;
        CK(cudaMalloc((void **)&ctx->isg_rec, npix * 8 * sizeof(float)));
        CK(cudaMalloc((void **)&ctx->isg_trained, npix * sizeof(float)));
        ctx->isg_pixels = npix;
    }
    CK(cudaMemset(ctx->isg_rec, 0, npix * 8 * sizeof(float)));
    ctx->isg_saved_vsp = ctx->sc.isg_vsp; ctx->isg_saved_w = ctx->sc.isg_w; ctx->isg_saved_h = ctx->sc.isg_h;
    ctx->sc.isg_vsp = nullptr;          // not ready: primary rays use vsp = 0.5 (vspg.cpp:1100-1105) until the first update
    ctx->sc.isg_w = ctx->film_w; ctx->sc.isg_h = ctx->film_h;
    ctx->sc.isg_rec = ctx->isg_rec;
    ctx->sc.isg_variance = ctx->sc.ip.vsp_criterion_variance ? 1 : 0;
    ctx->isg_active = true;
    return VSPG_OK;
}
int vspg_isg_store_update(vspg_ctx *ctx) {
    if (!ctx) return VSPG_ERR_INVALID;
    CK_STATE(ctx->isg_active, "isg_store_update: isg_store_begin has not been called");
    CK(cudaSetDevice(ctx->device));
    k_isg_update<<<ctx->sm_count * 4, 256>>>(ctx->isg_rec, (int)ctx->isg_pixels, ctx->sc.isg_variance, ctx->isg_trained);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    ctx->sc.isg_vsp = ctx->isg_trained;   // ready: the following waves read the estimate
    return VSPG_OK;
}
int vspg_isg_store_download(vspg_ctx *ctx, float *out_vsp, float *out_records8) {
    if (!ctx) return VSPG_ERR_INVALID;
    CK_STATE(ctx->isg_active, "isg_store_download: isg_store_begin has not been called");
    CK(cudaSetDevice(ctx->device));
    if (out_vsp) {
        CK_STATE(ctx->sc.isg_vsp == ctx->isg_trained, "isg_store_download: no isg_store_update yet");
        CK(cudaMemcpy(out_vsp, ctx->isg_trained, ctx->isg_pixels * sizeof(float), cudaMemcpyDeviceToHost));
    }
    if (out_records8) CK(cudaMemcpy(out_records8, ctx->isg_rec, ctx->isg_pixels * 8 * sizeof(float), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}
int vspg_isg_store_end(vspg_ctx *ctx) {
    if (!ctx) return VSPG_ERR_INVALID;
    if (!ctx->isg_active) return VSPG_OK;
    ctx->sc.isg_rec = nullptr;
    ctx->sc.isg_vsp = ctx->isg_saved_vsp; ctx->sc.isg_w = ctx->isg_saved_w; ctx->sc.isg_h = ctx->isg_saved_h;
    ctx->isg_active = false;
    return VSPG_OK;
}
int vspg_film_device_ptr(vspg_ctx *ctx, double **rgbw, size_t *n_doubles) {
    if (!ctx || !ctx->film || !rgbw) return VSPG_ERR_STATE;
    *rgbw = ctx->film;
    if (n_doubles) *n_doubles = (size_t)ctx->film_w * ctx->film_h * 4;
    return VSPG_OK;
}
int vspg_film_download(vspg_ctx *ctx, double *host_rgbw) {
    if (!ctx || !ctx->film || !host_rgbw) return VSPG_ERR_STATE;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(host_rgbw, ctx->film, (size_t)ctx->film_w * ctx->film_h * 4 * sizeof(double), cudaMemcpyDeviceToHost));
    return VSPG_OK;
}
int vspg_film_resolve(vspg_ctx *ctx, float *host_rgb) {  // film.h:269-290 GetPixelRGB: rgbSum / weightSum
    if (!ctx || !ctx->film || !host_rgb) return VSPG_ERR_STATE;
    size_t n = (size_t)ctx->film_w * ctx->film_h;
    std::vector<double> tmp(n * 4);
    int rc = vspg_film_download(ctx, tmp.data());
    if (rc != VSPG_OK) return rc;
    for (size_t i = 0; i < n; i++) {
        double w = tmp[4 * i + 3];
        for (int k = 0; k < 3; k++) host_rgb[3 * i + k] = w != 0 ? (float)(tmp[4 * i + k] / w) : 0.f;
    }
    return VSPG_OK;
}
int vspg_counters_get(vspg_ctx *ctx, vspg_counters *out) {
    if (!ctx || !out) return VSPG_ERR_INVALID;
    memset(out, 0, sizeof *out);
    if (!ctx->counters) return VSPG_OK;
    CK(cudaSetDevice(ctx->device));
    unsigned long long h[CNT_COUNT];
    CK(cudaMemcpy(h, ctx->counters, sizeof h, cudaMemcpyDeviceToHost));
    out->paths = h[CNT_PATHS]; out->density_queries = h[CNT_QUERIES] + h[CNT_QUERIES_SH]; out->dda_steps = h[CNT_DDA] + h[CNT_DDA_SH] + h[CNT_DDA_PRE]; out->cache_lookups = h[CNT_LOOKUPS];
    out->bvh_nodes = h[CNT_BVH]; out->shadow_rays = h[CNT_SHADOW]; out->kernel_launches = ctx->launches; out->wave_iterations = ctx->iterations;
    return VSPG_OK;
}
int vspg_counters_reset(vspg_ctx *ctx) {
    if (!ctx) return VSPG_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    if (ctx->counters) CK(cudaMemset(ctx->counters, 0, CNT_COUNT * sizeof(unsigned long long)));
    ctx->launches = 0; ctx->iterations = 0; ctx->ms_total = 0;
    for (int k = 0; k < VSPG_K_COUNT; k++) { ctx->ms_kernel[k] = 0; ctx->launches_kernel[k] = 0; }
    return VSPG_OK;
}
int vspg_kernel_stats_get(vspg_ctx *ctx, int enable_timing, vspg_kernel_stats *out) {
    if (!ctx) return VSPG_ERR_INVALID;
    ctx->timing = enable_timing != 0;
    if (!out) return VSPG_OK;
    memset(out, 0, sizeof *out);
    for (int k = 0; k < VSPG_K_COUNT; k++) { out->ms[k] = (float)ctx->ms_kernel[k]; out->launches[k] = ctx->launches_kernel[k]; }
    out->ms_total = (float)ctx->ms_total;
    if (ctx->counters) {
        CK(cudaSetDevice(ctx->device));
        unsigned long long h[CNT_COUNT];
        CK(cudaMemcpy(h, ctx->counters, sizeof h, cudaMemcpyDeviceToHost));
        out->density_queries[VSPG_K_MEDIUM] = h[CNT_QUERIES]; out->dda_steps[VSPG_K_MEDIUM] = h[CNT_DDA];
        out->density_queries[VSPG_K_SHADOW] = h[CNT_QUERIES_SH]; out->dda_steps[VSPG_K_SHADOW] = h[CNT_DDA_SH];
        out->dda_steps[VSPG_K_INTERSECT] = h[CNT_DDA_PRE];
    }
    return VSPG_OK;
}

}  // extern "C"
