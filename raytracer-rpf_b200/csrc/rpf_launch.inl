// rpf_launch.inl -- host side of one channel layout: shared-memory plans and kernel launches of an ApplyRPFFilter pass
// (rpf.cpp:497-733). RE-INCLUDED by rpf_api.cu once per layout of rpf_layouts.cuh with RPF_NS / RPF_NR / RPF_NF / RPF_FULL
// set; everything lands in that layout's namespace next to its kernels. rpf_api.cu dispatches through LayoutOps.

namespace RPF_NS {

#if RPF_FULL
static bool plan_smem(long long ncap, int spp, int limit, SmemLayout &L) {
    if (ncap > rpf_fused_max_candidates()) return false;
    L.ncap = static_cast<int>(ncap);
    L.nstride = L.ncap | 1;
    const int nb = bins_for(L.ncap);
    // words per lane of a joint histogram (4 u8 cells per word); also large enough for the exact path's
    // contiguous u16 table of nb*nb+nb entries spread over the warp's 32 columns
    L.wpl = (nb * nb + 3) / 4;
    while (L.wpl * 128 < 2 * (nb * nb + nb)) ++L.wpl;
    int off = 0;
    L.off_idx = off; off = align16(off + 4 * L.ncap);
    L.off_bins = off; off = align16(off + 20 * L.ncap);
    L.off_gtab = off; off = align16(off + 8 * (L.ncap + 1));
    L.off_small = off;
    L.off_big = off;
    const long long raw = 1ll * L.nstride * (NC * 8 + 16 * 4);
    const long long wgt = 18ll * spp * 8;
    for (int mw = 4; mw >= 1; mw >>= 1) {
        // with four histogram warps the last one only holds the 19 marginals (nb cells per lane)
        const long long hist = mw == 4 ? (3ll * L.wpl + (nb + 3) / 4) * 128 : 1ll * mw * L.wpl * 128;
        long long big = raw > wgt ? raw : wgt;
        if (hist > big) big = hist;
        if (off + big <= limit) {
            L.mi_warps = mw;
            L.total = static_cast<int>(off + big);
            return true;
        }
    }
    return false;
}

#endif

// Shared-memory plan of the dependency kernel for ncap = box*box*spp candidates. Two CTAs of 512 threads per SM when
// one full-size histogram (the rare pixel that accepts every candidate) fits in half an SM's shared memory, else one CTA
// of 1024 threads (box 55 at 8 spp: 355 vs 419 ms with two CTAs). Whatever is left
// after the fixed tables goes to the histogram region (more slots = more warp teams); the bin bytes stay in shared
// memory only if eight full-size histograms still fit beside them, otherwise they go to an L2-resident per-CTA scratch.
static bool plan_general(long long ncap, int spp, int static_smem, GeneralPlan &L) {
    if (ncap > kLargeMaxCandidates) return false;
    L = GeneralPlan{};
    L.ncap = static_cast<int>(ncap);
    L.mw = ((L.ncap + 127) / 128) * 4;
    // Beyond 65,535 candidates this kernel still serves every pixel whose ACCEPTED count N stays within 65,535 (u16
    // cells, u8 bins): candidate numbers then go to a 32-bit list in HBM and the per-sample sizes are capped at 65,535.
    L.idx_global = ncap > kGeneralMaxCandidates;
    const int neff = L.idx_global ? kGeneralMaxCandidates : L.ncap;
    L.nq = (neff + 3) / 4;
    const int nb = bins_for(neff);
    L.nbs = (nb + 3) & ~3;
    L.hw = (((nb * nb + 1) / 2) + 3) & ~3;
    L.off_gsm = 0;
    L.off_marg = KB_GSM * 8;
    L.off_region = align16(L.off_marg + NCH * L.nbs * 4);
    L.kc_smem = spp * ZS * 8;
    const int idx_bytes = L.idx_global ? 0 : align16(2 * L.ncap);
    const int bins_bytes = NCH * L.nq * 4;
    const int hist_bytes = L.hw * 4;
    auto atleast = [&](int n) { const int h = n * hist_bytes; return h > idx_bytes ? h : idx_bytes; };
    const int limit[2] = {(228 * 1024) / 2 - static_smem - 1024, 227 * 1024 - static_smem - 1024};
    for (int v = 0; v < 2; ++v) {
        const int avail = (limit[v] - L.off_region) & ~15;
        if (avail < atleast(1)) continue;
        L.nt = v == 0 ? 512 : 1024;
        L.bins_in_smem = bins_bytes + atleast(8) <= avail;
        // neighbourhoods whose CAPACITY does not fit (boxes 35 / 55) still mostly accept few enough samples: per-pixel choice
        L.bins_dyn = !L.bins_in_smem && !L.idx_global && !(getenv("RPF_BINS_DYN") && atoi(getenv("RPF_BINS_DYN")) == 0);
        const int region = L.bins_in_smem ? avail - bins_bytes : avail;
        L.off_bins = L.bins_in_smem ? L.off_region + region : 0;
        L.region_words = region / 4;
        L.kb_smem = L.off_region + avail;
        return true;
    }
    return false;
}

// Plan of dependency_large_kernel (rpf_large.cuh): one CTA of 1024 threads per SM; shared memory = small-count table,
// 19 marginals, and everything else as the histogram region (u32 cells).
#if RPF_FULL
// Plan of dependency_large_kernel (rpf_large.cuh): one CTA of 1024 threads per SM; shared memory = small-count table,
// 19 marginals, and everything else as the histogram region (u32 cells).
static bool plan_large(long long ncap, int spp, int static_smem, GeneralPlan &L) {
    if (ncap > kLargeMaxCandidates) return false;
    L = GeneralPlan{};
    L.large = 1;
    L.ncap = static_cast<int>(ncap);
    L.mw = ((L.ncap + 127) / 128) * 4;
    L.nq = (L.ncap + 3) / 4;
    const int nb = bins_for(L.ncap);
    L.nbs = (nb + 3) & ~3;
    L.off_gsm = 0;
    L.off_marg = KB_GSM * 8;
    L.off_region = align16(L.off_marg + NCH * L.nbs * 4);
    L.kc_smem = spp * 18 * 8;
    const int limit = 227 * 1024 - static_smem - 1024;
    const int avail = (limit - L.off_region) & ~15;
    if (avail < nb * 4 * 4) return false;
    L.nt = KL_THREADS;
    L.region_words = avail / 4;
    L.kb_smem = L.off_region + avail;
    return true;
}
#endif

static int ensure_pstat(rpf_ctx *ctx, const FrameView &f, int y0, int y1) {
    (void)y0; (void)y1;
    if (ctx->pstat_y1 > ctx->pstat_y0) return RPF_OK;
    // Features never change between iterations (rpf.cpp:715 writes colours only): computed once per upload, for EVERY
    // row of the context (0.16 ms at 720p) -- callers may filter any rows in any order (rpf_filter_rows), and a cached
    // interval merged across a gap would leave rows in between uninitialised.
    const long long n = 1ll * ctx->cfg.height * ctx->cfg.width * NF;  // one thread per (pixel, feature)
    prof_begin(ctx, RPF_K_PIXEL_STATS);
    pixel_stats_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, ctx->stream>>>(f, 0, ctx->cfg.height, ctx->d_pstat, ctx->min_sigma,
                                                                                         ctx->pos_sigmas, ctx->d_pbound);
    prof_end(ctx);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    ctx->pstat_y0 = 0;
    ctx->pstat_y1 = ctx->cfg.height;
    return RPF_OK;
}

// Rows [y0,y1) and, when y3 > y2, rows [y2,y3) as well (y2 >= y1): both ranges go through ONE set of launches as a chunk
// with a gap (chunk_row, rpf_general.cuh).
static int launch_general(rpf_ctx *ctx, int box, double seed, int y0, int y1, int y2 = 0, int y3 = 0) {
    GeneralParams p{};
    const long long ncap = 1ll * box * box * ctx->cfg.spp;
    // Up to 65,535 candidates: dependency_kernel. Beyond: dependency_kernel for the pixels that accept at most 65,535
    // samples and dependency_large_kernel (plan pl_large) for the rest. RPF_FLAG_FORCE_LARGE: the large kernel only.
    GeneralPlan pl_large{};
    bool use_small = !ctx->force_large && plan_general(ncap, ctx->cfg.spp, ctx->kb_static_smem, p.pl);
#if RPF_FULL
    const bool use_large = (!use_small || p.pl.idx_global) && plan_large(ncap, ctx->cfg.spp, ctx->kb_static_smem, pl_large);
#else
    const bool use_large = false;  // beyond 65,535 candidates: reference layout only
    if (use_small && p.pl.idx_global) use_small = false;
#endif
    if (use_small && p.pl.idx_global && !use_large) use_small = false;
    if (!use_small && !use_large)
        return fail(ctx, RPF_ERR_UNSUPPORTED, "neighbourhood of box %d x %d spp = %lld candidates exceeds this build (max %d)", box,
                    ctx->cfg.spp, ncap, kLargeMaxCandidates);
    if (!use_small) p.pl = pl_large;
#if RPF_FULL
    if ((KW_THREADS * KW_ZS + ctx->cfg.spp * 18) * 8 > 100 * 1024) return fail(ctx, RPF_ERR_UNSUPPORTED, "spp %d too large for the blend kernel", ctx->cfg.spp);
#else
    if (ctx->cfg.spp >= ctx->blend_wide_min)
        return fail(ctx, RPF_ERR_UNSUPPORTED, "%d samples per pixel: the wide blend kernel exists for the reference channel layout only (this context: %d random "
                    "parameters, %d features)", ctx->cfg.spp, NR, NF);
#endif
    int rc = RPF_OK;
    if (use_large) rc = ensure_hql(ctx, pl_large.ncap);
    if (rc == RPF_OK && use_small) rc = ensure_gtab(ctx, p.pl.idx_global ? kGeneralMaxCandidates : p.pl.ncap);
    if (rc != RPF_OK) return rc;
    p.f = FrameView{ctx->d_planes, ctx->S, ctx->cfg.width, ctx->cfg.height, ctx->cfg.spp};
    rc = ensure_pstat(ctx, p.f, y0, y1);
    if (rc != RPF_OK) return rc;
    const int W = ctx->cfg.width;
    const size_t per_pixel = static_cast<size_t>(p.pl.mw) * 4 + REC_DOUBLES * 8;
    const int n1 = y1 - y0, n2 = y3 > y2 ? y3 - y2 : 0, nv = n1 + n2;  // rows of the first range, of the second, together
    long long rows = static_cast<long long>(ctx->scratch_budget / (per_pixel * W));
    if (rows > 8) rows &= ~7ll;  // chunks start on a multiple of the membership tile height, so the gap does too
    if (rows < 1) rows = 1;
    if (rows > nv) rows = nv;
    const long long chunk_px = rows * W;
    auto grid_for = [&](GeneralPlan &L) -> int {
        int occ = 0;
        cudaError_t e;
#if RPF_FULL
        if (L.large) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, dependency_large_kernel, KL_THREADS, L.kb_smem);
        else
#endif
        if (L.nt == 512) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, dependency_kernel<512>, 512, L.kb_smem);
        else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, dependency_kernel<1024>, 1024, L.kb_smem);
        if (e != cudaSuccess || occ < 1) return 0;
        L.kb_grid = static_cast<int>(chunk_px < 1ll * occ * ctx->num_sms ? chunk_px : 1ll * occ * ctx->num_sms);
        return L.kb_grid;
    };
    if (!grid_for(p.pl) || (use_small && use_large && !grid_for(pl_large)))
        return fail(ctx, RPF_ERR_CUDA, "dependency kernel does not fit on an SM with %d bytes of shared memory", p.pl.kb_smem);
    if ((rc = ensure_buffer(ctx, &ctx->d_mask, &ctx->mask_bytes, static_cast<size_t>(chunk_px) * p.pl.mw * 4)) != RPF_OK) return rc;
    if ((rc = ensure_buffer(ctx, &ctx->d_rec, &ctx->rec_bytes, static_cast<size_t>(chunk_px) * REC_DOUBLES * 8)) != RPF_OK) return rc;
    // big_scratch per persistent CTA: the large kernel keeps 4 B candidate number + 19 u16 bins per candidate (rows padded
    // to 4); dependency_kernel beyond 65,535 candidates keeps its 32-bit candidate list there
    const GeneralPlan &PL = use_large ? (use_small ? pl_large : p.pl) : p.pl;
    size_t big_stride = 0;
    int big_grid = 0;
    if (use_large) { big_stride = static_cast<size_t>(PL.nq) * 168; big_grid = PL.kb_grid; }
    if (use_small && p.pl.idx_global) {
        const size_t st = (static_cast<size_t>(p.pl.ncap) * 4 + 15) & ~static_cast<size_t>(15);
        if (st > big_stride) big_stride = st;
        if (p.pl.kb_grid > big_grid) big_grid = p.pl.kb_grid;
    }
    p.big_stride = big_stride;
    if (big_stride && (rc = ensure_buffer(ctx, &ctx->d_big, &ctx->big_bytes, static_cast<size_t>(big_grid) * big_stride)) != RPF_OK) return rc;
    if (use_small && !p.pl.bins_in_smem &&
        (rc = ensure_buffer(ctx, &ctx->d_bins, &ctx->bins_bytes, static_cast<size_t>(p.pl.kb_grid) * NCH * p.pl.nq * 4)) != RPF_OK)
        return rc;
    if (!ctx->d_rec16) {
        CU(ctx, cudaMalloc(&ctx->d_rec16, ctx->S * NQ4 * sizeof(float4)));
        CU(ctx, cudaMalloc(&ctx->d_col4, ctx->S * sizeof(double4)));
        CU(ctx, cudaMalloc(&ctx->d_rec19, ctx->S * REC19 * sizeof(double)));
    }
    if (!ctx->rec16_valid) {
        pack_features_kernel<<<static_cast<unsigned>((NQ4 * ctx->S + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_planes, ctx->S, ctx->d_rec16);
        pack_rec19_kernel<<<static_cast<unsigned>((ctx->S * REC19 + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_planes, ctx->S, ctx->d_rec19);
        ctx->launches += 2;
        ctx->rec16_valid = true;
    }
    if (!ctx->col4_valid) {  // once per pass, however many row ranges the pass is launched in (band edges first, then the interior)
        pack_colors_kernel<<<static_cast<unsigned>((ctx->S + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_col[ctx->cur], ctx->S, ctx->d_col4, ctx->d_rec19);
        ctx->launches++;
        CU(ctx, cudaGetLastError());
        ctx->col4_valid = true;
    }
    p.rec16 = ctx->d_rec16;
    p.col4 = ctx->d_col4;
    p.rec19 = ctx->d_rec19;
    p.col_in = ctx->d_col[ctx->cur];
    p.col_out = ctx->d_col[ctx->cur ^ 1];
    p.pstat = ctx->d_pstat;
    p.gq = ctx->d_gq;
    p.hq = p.pl.large ? ctx->d_hql : ctx->d_hq;
    p.large_min_n = 0;
    p.big_scratch = ctx->d_big;
    p.box = box;
    p.cd = make_cand_decode(box, ctx->cfg.spp);
    p.flags = ctx->cfg.flags;
    p.seed = seed;
    p.w_eps = ctx->w_eps;
    p.mask = ctx->d_mask;
    p.rec = ctx->d_rec;
    p.bins_scratch = ctx->d_bins;
    p.work_counter = ctx->d_work;
    p.dbg_n = ctx->d_dbg_n;
    p.dbg_alpha = ctx->d_dbg_alpha;
    p.dbg_beta = ctx->d_dbg_beta;
    p.dbg_wrc = ctx->d_dbg_wrc;
    p.counters = ctx->d_counters;
    p.probe = DebugProbe{ctx->d_probe_mi ? ctx->probe_pixel : -1, ctx->d_probe_mi, ctx->d_probe_marg};
    for (int v0 = 0; v0 < nv; v0 += static_cast<int>(rows)) {  // v0: first row of the chunk, counted over both ranges
        p.nrows = v0 + rows <= nv ? static_cast<int>(rows) : nv - v0;
        p.row0 = v0 < n1 ? y0 + v0 : y2 + (v0 - n1);
        const bool straddles = n2 > 0 && v0 < n1 && v0 + p.nrows > n1;
        p.gap_at = straddles ? n1 - v0 : 0x7fffffff;
        p.gap = straddles ? y2 - y1 : 0;
        const unsigned npx = static_cast<unsigned>(p.nrows * W);
        CU(ctx, cudaMemsetAsync(ctx->d_work, 0, 2 * sizeof(unsigned int), ctx->stream));
        {
            MembershipParams m{};
            m.f = p.f;
            m.pbound = ctx->d_pbound;
            m.rec16 = ctx->d_rec16;
            m.mask = ctx->d_mask;
            m.box = box;
            m.row0 = p.row0;
            m.nrows = p.nrows;
            m.gap_at = p.gap_at;
            m.gap = p.gap;
            m.mw = p.pl.mw;
            m.sub = ctx->sub;
            membership_tile(box, m.tx, m.ty);
            if (const char *e = getenv("RPF_MEMB_TILE")) {  // tuning: "TXxTY", TY a divisor of 8 (chunks and gaps start on multiples of 8 rows)
                int tx = 0, ty = 0;
                if (sscanf(e, "%dx%d", &tx, &ty) == 2 && tx > 0 && ty > 0 && 8 % ty == 0 && STRIP_W % tx == 0 && tx * ty <= KM_MAX_THREADS && (tx * ty) % 32 == 0) {
                    m.tx = tx;
                    m.ty = ty;
                }
            }
            m.tiles_x = (W + m.tx - 1) / m.tx;
            m.col_rows = m.ty + box - 1;
            const size_t row_bytes = static_cast<size_t>(membership_row_floats(ctx->cfg.spp)) * sizeof(float);
            m.rb_rows = m.col_rows;
            const size_t memb_kb = getenv("RPF_MEMB_SMEM_KB") ? static_cast<size_t>(atoi(getenv("RPF_MEMB_SMEM_KB"))) : 32;  // per CTA: six CTAs per SM
            while (m.rb_rows > 1 && m.rb_rows * row_bytes > memb_kb * 1024) m.rb_rows = (m.rb_rows + 1) / 2;
            const unsigned tiles = static_cast<unsigned>(m.tiles_x) * static_cast<unsigned>((p.nrows + m.ty - 1) / m.ty);
            prof_begin(ctx, RPF_K_MEMBERSHIP);
            if (m.sub.keep_q24 < RPF_SUBSAMPLE_ALL) membership_kernel<true><<<tiles, m.tx * m.ty, m.rb_rows * row_bytes, ctx->stream>>>(m);
            else membership_kernel<false><<<tiles, m.tx * m.ty, m.rb_rows * row_bytes, ctx->stream>>>(m);
            prof_end(ctx);
        }
        prof_begin(ctx, RPF_K_ORDERED_STATS);
        ordered_stats_kernel<<<(npx + KO_WARPS - 1) / KO_WARPS, KO_WARPS * 32, 0, ctx->stream>>>(p);
        prof_end(ctx);
        const unsigned gridb = npx < static_cast<unsigned>(p.pl.kb_grid) ? npx : static_cast<unsigned>(p.pl.kb_grid);
        prof_begin(ctx, RPF_K_DEPENDENCY);
#if RPF_FULL
        if (p.pl.large) dependency_large_kernel<<<gridb, KL_THREADS, p.pl.kb_smem, ctx->stream>>>(p);
        else
#endif
        if (p.pl.nt == 512) dependency_kernel<512><<<gridb, 512, p.pl.kb_smem, ctx->stream>>>(p);
        else dependency_kernel<1024><<<gridb, 1024, p.pl.kb_smem, ctx->stream>>>(p);
#if RPF_FULL
        if (use_small && use_large) {  // the pixels dependency_kernel left: more than 65,535 accepted samples
            GeneralParams q = p;
            q.pl = pl_large;
            q.hq = ctx->d_hql;
            q.large_min_n = kGeneralMaxCandidates + 1;
            q.work_counter = ctx->d_work + 1;
            const unsigned gridl = npx < static_cast<unsigned>(pl_large.kb_grid) ? npx : static_cast<unsigned>(pl_large.kb_grid);
            dependency_large_kernel<<<gridl, KL_THREADS, pl_large.kb_smem, ctx->stream>>>(q);
            ctx->launches++;
        }
#endif
        prof_end(ctx);
        prof_begin(ctx, RPF_K_BLEND);
        // few samples per pixel: a thread per neighbour with the own samples' accumulators in registers; many: own samples
        // across threads, neighbours staged through shared memory (rpf_blend_wide.cuh)
#if RPF_FULL
        if (ctx->cfg.spp >= ctx->blend_wide_min) blend_wide_kernel<<<npx, KW_THREADS, (KW_THREADS * KW_ZS + ctx->cfg.spp * 18) * 8, ctx->stream>>>(p);
        else
#endif
        blend_kernel<<<npx, KC_THREADS, p.pl.kc_smem, ctx->stream>>>(p);
        prof_end(ctx);
        ctx->launches += 4;
        CU(ctx, cudaGetLastError());
    }
    ctx->pixels_filtered += 1ll * nv * W;
    return RPF_OK;
}

static int launch_rows(rpf_ctx *ctx, int box, double seed, int y0, int y1, int y2, int y3) {
    if (!ctx->uploaded) return fail(ctx, RPF_ERR_STATE, "rpf_filter: no samples uploaded");
    if (box < 1 || (box & 1) == 0) return fail(ctx, RPF_ERR_INVALID, "box size %d must be odd and positive (rpf.cpp:561)", box);
    if (y0 < 0 || y1 > ctx->cfg.height || y0 >= y1) return fail(ctx, RPF_ERR_INVALID, "bad row range [%d,%d)", y0, y1);
    if (y3 > y2 && (y2 < y1 || y3 > ctx->cfg.height)) return fail(ctx, RPF_ERR_INVALID, "bad second row range [%d,%d)", y2, y3);
    if (!(seed > 0.0)) return fail(ctx, RPF_ERR_INVALID, "sigma_fc_seed must be positive");
    if (y3 > y2 && y2 == y1) { y1 = y3; y2 = y3 = 0; }  // adjacent ranges are one range
#if !RPF_FULL
    return launch_general(ctx, box, seed, y0, y1, y2, y3);  // wider layouts: the general path only
#else
    FilterParams p{};
    // the fused shared-memory-resident kernel serves small neighbourhoods; everything else takes the general path
    if (ctx->force_general || !plan_smem(1ll * box * box * ctx->cfg.spp, ctx->cfg.spp, ctx->max_dyn_smem, p.sm))
        return launch_general(ctx, box, seed, y0, y1, y2, y3);
    if (y3 > y2) {  // the fused kernel takes one contiguous range per launch
        const int rc2 = launch_rows(ctx, box, seed, y2, y3, 0, 0);
        if (rc2 != RPF_OK) return rc2;
    }
    int rc = ensure_gtab(ctx, p.sm.ncap);
    if (rc != RPF_OK) return rc;
    p.f = FrameView{ctx->d_planes, ctx->S, ctx->cfg.width, ctx->cfg.height, ctx->cfg.spp};
    rc = ensure_pstat(ctx, p.f, y0, y1);
    if (rc != RPF_OK) return rc;
    p.col_in = ctx->d_col[ctx->cur];
    p.col_out = ctx->d_col[ctx->cur ^ 1];
    p.pstat = ctx->d_pstat;
    p.gtab = ctx->d_gtab;
    p.gpre = ctx->d_gpre;
    p.box = box;
    p.y0 = y0;
    p.y1 = y1;
    p.flags = ctx->cfg.flags;
    p.seed = seed;
    p.w_eps = ctx->w_eps;
    p.dbg_n = ctx->d_dbg_n;
    p.dbg_alpha = ctx->d_dbg_alpha;
    p.dbg_beta = ctx->d_dbg_beta;
    p.dbg_wrc = ctx->d_dbg_wrc;
    p.counters = ctx->d_counters;
    p.probe = DebugProbe{ctx->d_probe_mi ? ctx->probe_pixel : -1, ctx->d_probe_mi, ctx->d_probe_marg};
    p.sub = ctx->sub;
    const unsigned grid = static_cast<unsigned>((y1 - y0) * ctx->cfg.width);
    prof_begin(ctx, RPF_K_FUSED);
    if (p.sub.keep_q24 < RPF_SUBSAMPLE_ALL) rpf_pixel_kernel<true><<<grid, THREADS, p.sm.total, ctx->stream>>>(p);
    else rpf_pixel_kernel<false><<<grid, THREADS, p.sm.total, ctx->stream>>>(p);
    prof_end(ctx);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    ctx->pixels_filtered += grid;
    return RPF_OK;
#endif
}

// One-time per-context setup of this layout's kernels: opt-in shared-memory sizes.
static int configure(rpf_ctx *ctx) {
    cudaFuncAttributes fa;
    CU(ctx, cudaFuncGetAttributes(&fa, dependency_kernel<1024>));
    ctx->kb_static_smem = static_cast<int>(fa.sharedSizeBytes);
    CU(ctx, cudaFuncSetAttribute(dependency_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - ctx->kb_static_smem - 1024));
    CU(ctx, cudaFuncSetAttribute(dependency_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (228 * 1024) / 2 - ctx->kb_static_smem - 1024));
    CU(ctx, cudaFuncSetAttribute(membership_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(ctx, cudaFuncSetAttribute(membership_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#if RPF_FULL
    CU(ctx, cudaFuncSetAttribute(rpf_pixel_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    CU(ctx, cudaFuncSetAttribute(rpf_pixel_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    CU(ctx, cudaFuncSetAttribute(blend_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CU(ctx, cudaFuncSetAttribute(dependency_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - ctx->kb_static_smem - 1024));
#endif
    return RPF_OK;
}

static const LayoutOps kOps = {NR, NF, NCH, N_JOINT, REC19, &launch_rows, &configure};

}  // namespace RPF_NS
