// vspg_set_guiding / vspg_set_integrator / vspg_set_capacity  (included by vspg_ctx.cu)

static int upload_field(vspg_ctx *ctx, const vspg_field *f, DField *d) {
    memset(d, 0, sizeof *d);
    if (!f) return VSPG_OK;
    if (f->is_nn) return fail(ctx, VSPG_ERR_UNSUPPORTED, "set_guiding: importance-sampled NN lookup (m_useISNNLookUp) is not supported");
    if (f->n_nodes <= 0 || f->n_regions <= 0 || !f->nodes || !f->regions) return fail(ctx, VSPG_ERR_INVALID, "set_guiding: empty field");
    for (int i = 0; i < f->n_nodes; i++) {
        uint32_t v = f->nodes[i].dim_idx, dim = v >> 30, idx = (v << 2) >> 2;
        if (dim == 3 ? idx >= (uint32_t)f->n_regions : idx + 1 >= (uint32_t)f->n_nodes)
            return fail(ctx, VSPG_ERR_INVALID, "set_guiding: KD node %d points outside the tree", i);
    }
    for (int i = 0; i < f->n_regions; i++)
        if (f->regions[i].num_components < 0 || f->regions[i].num_components > VSPG_MAX_LOBES)
            return fail(ctx, VSPG_ERR_INVALID, "set_guiding: region %d has %d lobes", i, f->regions[i].num_components);
    d->usable = f->usable; d->n_nodes = f->n_nodes; d->n_regions = f->n_regions; d->stochastic_nn = f->stochastic_nn;
    for (int a = 0; a < 3; a++) { d->bmin[a] = f->bounds_min[a]; d->bmax[a] = f->bounds_max[a]; }
    CK(upload(ctx, ctx->guide_allocs, f->nodes, (size_t)f->n_nodes, &d->nodes));
    CK(upload(ctx, ctx->guide_allocs, f->regions, (size_t)f->n_regions, &d->regions));
    if (f->neighbours) {
        for (int i = 0; i < f->n_regions; i++)
            for (int s = 0; s < 8; s++)
                if (f->neighbours[i].ids[s] != ~0u && f->neighbours[i].ids[s] >= (uint32_t)f->n_regions)
                    return fail(ctx, VSPG_ERR_INVALID, "set_guiding: neighbour table of region %d is out of range", i);
        CK(upload(ctx, ctx->guide_allocs, f->neighbours, (size_t)f->n_regions, &d->neigh));
    }
    d->present = 1;   // only once every table is on the device: a failed upload leaves the field absent, not half-present
    return VSPG_OK;
}

extern "C" {

int vspg_set_guiding(vspg_ctx *ctx, const vspg_guiding_desc *g) {
    if (!ctx) return VSPG_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    dev_release(ctx, ctx->guide_allocs);
    struct Trim { vspg_ctx *c; ~Trim() { dev_trim(c); } } trimOnExit{ctx};
    DScene &sc = ctx->sc;
    memset(&sc.surf, 0, sizeof sc.surf); memset(&sc.vol, 0, sizeof sc.vol);
    sc.hg_table = nullptr; sc.isg_vsp = nullptr; sc.tr_rgb = nullptr; sc.isg_w = sc.isg_h = sc.tr_w = sc.tr_h = 0;
    if (!g) return VSPG_OK;
    int rc;
    if ((rc = upload_field(ctx, g->surface, &sc.surf)) != VSPG_OK) return rc;
    if ((rc = upload_field(ctx, g->volume, &sc.vol)) != VSPG_OK) return rc;
    if (sc.vol.present) {
        // directional/vmm/VMMPhaseFunctions.cpp: 128 single-lobe-HG -> vMF-mixture fits (data table)
        std::vector<DHGRow> rows(128);
        for (int i = 0; i < 128; i++) {
            float r[18]; memcpy(r, vspg_hg_vmm_table_bits + 18 * i, sizeof r);
            rows[i].K = (int)r[0]; rows[i].g = r[1];
            for (int j = 0; j < 4; j++) { rows[i].w[j] = r[2 + j]; rows[i].mc[j] = r[6 + j]; rows[i].kappa[j] = r[10 + j]; rows[i].norm[j] = r[14 + j]; }
            if (rows[i].K < 0 || rows[i].K > 3) return fail(ctx, VSPG_ERR_INVALID, "HG table row %d has K=%d", i, rows[i].K);
        }
        CK(upload(ctx, ctx->guide_allocs, rows.data(), rows.size(), &sc.hg_table));
    }
    if (g->isg_vsp) {
        if (g->isg_width != sc.cam.width || g->isg_height != sc.cam.height)
            return fail(ctx, VSPG_ERR_INVALID, "set_guiding: ISG buffer is %dx%d but the film is %dx%d (call set_scene first)", g->isg_width, g->isg_height, sc.cam.width, sc.cam.height);
        CK(upload(ctx, ctx->guide_allocs, g->isg_vsp, (size_t)g->isg_width * g->isg_height, &sc.isg_vsp));
        sc.isg_w = g->isg_width; sc.isg_h = g->isg_height;
    }
    if (g->tr_rgb) {
        if (g->tr_width != sc.cam.width || g->tr_height != sc.cam.height)
            return fail(ctx, VSPG_ERR_INVALID, "set_guiding: Tr buffer is %dx%d but the film is %dx%d", g->tr_width, g->tr_height, sc.cam.width, sc.cam.height);
        CK(upload(ctx, ctx->guide_allocs, g->tr_rgb, (size_t)g->tr_width * g->tr_height * 3, &sc.tr_rgb));
        sc.tr_w = g->tr_width; sc.tr_h = g->tr_height;
    }
    return VSPG_OK;
}

int vspg_set_integrator(vspg_ctx *ctx, const vspg_integrator_params *p) {
    if (!ctx || !p) return VSPG_ERR_INVALID;
    if (p->max_depth < 0 || p->max_depth > 255) return fail(ctx, VSPG_ERR_INVALID, "set_integrator: maxdepth must be in [0,255]");
    if (!(p->vsp_mis_ratio >= 0.f && p->vsp_mis_ratio <= 1.f)) return fail(ctx, VSPG_ERR_INVALID, "set_integrator: vspmisratio must be in [0,1]");
    if (p->light_sampler != VSPG_LIGHTSAMPLER_UNIFORM && p->light_sampler != VSPG_LIGHTSAMPLER_POWER && p->light_sampler != VSPG_LIGHTSAMPLER_BVH) return fail(ctx, VSPG_ERR_INVALID, "set_integrator: bad lightsampler");
    ctx->sc.ip = *p;
    return VSPG_OK;
}

int vspg_set_capacity(vspg_ctx *ctx, int max_paths_in_flight) {
    if (!ctx || max_paths_in_flight < 0) return VSPG_ERR_INVALID;
    ctx->requested_capacity = max_paths_in_flight;
    return VSPG_OK;
}

}  // extern "C"
