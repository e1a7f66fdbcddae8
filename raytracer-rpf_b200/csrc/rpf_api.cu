// rpf_api.cu -- the C ABI of include/rpf_b200.h on top of the sm_100a kernels.
// Replaces the call site RPFIntegrator::Render -> ApplyRPFFilter (reference src/custom/rpf.cpp:767-775).
#include "../../include/rpf_b200.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "rpf_layouts.cuh"  // every kernel, once per channel layout

using namespace rpf_types;

struct rpf_ctx;
// One channel layout = one compiled copy of the kernels (rpf_layouts.cuh) and of their launch code (rpf_launch.inl).
struct LayoutOps {
    int nr, nf, nch, n_joint, rec19;
    int (*launch_rows)(rpf_ctx *, int box, double seed, int y0, int y1, int y2, int y3);  // rows [y0,y1) and, if y3 > y2, [y2,y3)
    int (*configure)(rpf_ctx *);
};

struct rpf_ctx {
    rpf_config cfg{};
    const LayoutOps *ops = nullptr;  // the context's channel layout (rpf_config.n_random / n_features)
    size_t S = 0;
    int npix = 0;
    float *d_planes = nullptr;      // [nch][S]
    double *d_col[2] = {nullptr, nullptr};  // [3][S] each: current / next
    int cur = 0;
    double *d_pstat = nullptr;      // [npix][2 nf]
    float *d_pbound = nullptr;      // [npix][2 nf]: the same test as float intervals (membership_kernel)
    int pstat_y0 = 0, pstat_y1 = 0; // rows with valid per-pixel statistics
    double *d_gtab = nullptr, *d_gpre = nullptr, *d_hln = nullptr;
    unsigned long long *d_gq = nullptr, *d_hq = nullptr;
    unsigned long long *d_hql = nullptr;  // n ln n in units of 2^-38 for the large-neighbourhood kernel (rpf_large.cuh)
    int gtab_n = 0, hql_n = 0;
    uint8_t *d_big = nullptr;
    size_t big_bytes = 0;
    bool force_large = false;
    SubsampleParams sub{0ull, RPF_SUBSAMPLE_ALL, 0};
    double w_eps = 0.0, min_sigma = 0.0, pos_sigmas = 0.0;  // paper mode (rpf_set_paper_mode); all 0 = reference
    int blend_wide_min = 16;  // spp from which blend_wide_kernel replaces blend_kernel (RPF_BLEND_WIDE_MIN overrides)
    // general path (rpf_general.cuh): per-chunk scratch in HBM
    uint32_t *d_mask = nullptr;
    double *d_rec = nullptr;
    uint8_t *d_bins = nullptr;
    float4 *d_rec16 = nullptr;     // gather-friendly copy of the 16 non-colour planes (general path), valid per upload
    double4 *d_col4 = nullptr;     // current colours as {r,g,b,0}, rebuilt per general pass
    double *d_rec19 = nullptr;     // all 19 channels as doubles per sample (ordered_stats_kernel); colours rebuilt per pass
    bool rec16_valid = false;
    unsigned int *d_work = nullptr;
    size_t mask_bytes = 0, rec_bytes = 0, bins_bytes = 0;
    int kb_static_smem = 0, num_sms = 0;
    bool force_general = false;
    // optional per-kernel timing: an event pair around every launch, resolved at the next synchronising call
    bool profiling = false;
    struct Span { cudaEvent_t a, b; int k; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> free_events;
    double k_ms[RPF_NUM_KERNELS] = {};
    long long k_launches[RPF_NUM_KERNELS] = {};
    size_t scratch_budget = 2048ull << 20;
    unsigned long long *d_counters = nullptr;  // 4
    int *d_dbg_n = nullptr;
    double *d_dbg_alpha = nullptr, *d_dbg_beta = nullptr, *d_dbg_wrc = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool uploaded = false;
    bool timing_pending = false;
    long long pixels_filtered = 0, launches = 0;
    double last_ms = 0.0;
    int max_dyn_smem = 0;
    std::string err;
    long long probe_pixel = -1;     // rpf_set_debug_pixel
    double *d_probe_mi = nullptr;   // [98]
    unsigned *d_probe_marg = nullptr;  // [nch][PROBE_BINS]
    bool col4_valid = false;        // col4 / the colour slots of rec19 hold the CURRENT colours (general path, once per pass)
    // ---- row bands (rpf_bands.inl) ----
    std::vector<rpf_ctx *> shards;  // n_gpus > 1: this context is the parent of one child per device and owns no device memory
    std::vector<int> shard_own0;    // [n+1] first owned global row of every band
    std::vector<int> shard_buf0, shard_buf1;  // global rows a band's buffer holds
    int halo_rows = 0;
    rpf_band band{};                // one band per process (rpf_band_attach): world > 1
    void *nccl_comm = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_edge = nullptr, ev_copy = nullptr, ev_start = nullptr;
    unsigned long long halo_bytes = 0;  // colour bytes this context sent to neighbouring bands since reset_stats
};

static thread_local std::string g_create_err;

static int fail(rpf_ctx *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_create_err = buf;
    return code;
}

#define CU(ctx, call)                                                                         \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(ctx, RPF_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_));   \
    } while (0)

static int bins_for(int n) {
    int b = static_cast<int>(std::sqrt(static_cast<double>(n)));
    return b < 1 ? 1 : b;
}
static int align16(int v) { return (v + 15) & ~15; }

// Shared-memory plan for one pixel CTA. Returns false if the neighbourhood cannot fit.
// Largest box*box*spp the fused kernel takes. It CAN run up to 1,279 candidates (u8 histogram cells with at most 4
// remembered wraps, N >> 8 <= 4), but one CTA of 128 threads per pixel stops paying well before that: measured at 8 spp on
// B200, fused / general = 90 / 101 ms at box 7 (392 candidates), 188 / 121 ms at box 9 (648), 433 / 145 ms at box 11
// (968). RPF_FUSED_MAX overrides (tests keep the fused kernel covered up to its limit with it).
static constexpr int kFusedHardMax = 1279, kFusedDefaultMax = 512;
extern "C" int rpf_fused_max_candidates(void) {
    int v = kFusedDefaultMax;
    if (const char *e = getenv("RPF_FUSED_MAX")) v = atoi(e);
    return v < 0 ? 0 : (v > kFusedHardMax ? kFusedHardMax : v);
}

static const int kSmemLimit = 227 * 1024 - 6 * 1024;  // minus static PixelSmall and the per-CTA reserve

static const int kLargeMaxCandidates = 1 << 20;  // 2^-38 fixed-point entropy sums stay below 2^64 (rpf_large.cuh)

extern "C" int rpf_max_neighbourhood(void) { return kLargeMaxCandidates; }

static void band_release(rpf_ctx *ctx);  // rpf_bands.inl state: NCCL communicator, copy stream, events

extern "C" const char *rpf_last_error(const rpf_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" void rpf_destroy(rpf_ctx *ctx) {
    if (!ctx) return;
    if (!ctx->shards.empty()) {
        for (rpf_ctx *c : ctx->shards) rpf_destroy(c);
        delete ctx;
        return;
    }
    cudaSetDevice(ctx->cfg.device);
    band_release(ctx);
    cudaFree(ctx->d_planes);
    cudaFree(ctx->d_col[0]);
    cudaFree(ctx->d_col[1]);
    cudaFree(ctx->d_pstat);
    cudaFree(ctx->d_pbound);
    cudaFree(ctx->d_gtab);
    cudaFree(ctx->d_gpre);
    cudaFree(ctx->d_hln);
    cudaFree(ctx->d_gq);
    cudaFree(ctx->d_hq);
    cudaFree(ctx->d_hql);
    cudaFree(ctx->d_big);
    cudaFree(ctx->d_mask);
    cudaFree(ctx->d_rec);
    cudaFree(ctx->d_bins);
    cudaFree(ctx->d_rec16);
    cudaFree(ctx->d_col4);
    cudaFree(ctx->d_rec19);
    cudaFree(ctx->d_work);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_dbg_n);
    cudaFree(ctx->d_dbg_alpha);
    cudaFree(ctx->d_dbg_beta);
    cudaFree(ctx->d_dbg_wrc);
    cudaFree(ctx->d_probe_mi);
    cudaFree(ctx->d_probe_marg);
    for (auto &sp : ctx->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto e : ctx->free_events) cudaEventDestroy(e);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

static void prof_begin(rpf_ctx *ctx, int k) {
    if (!ctx->profiling) return;
    rpf_ctx::Span sp{nullptr, nullptr, k};
    for (cudaEvent_t *e : {&sp.a, &sp.b}) {
        if (!ctx->free_events.empty()) { *e = ctx->free_events.back(); ctx->free_events.pop_back(); }
        else cudaEventCreate(e);
    }
    cudaEventRecord(sp.a, ctx->stream);
    ctx->spans.push_back(sp);
}
static void prof_end(rpf_ctx *ctx) {
    if (ctx->profiling) cudaEventRecord(ctx->spans.back().b, ctx->stream);
}
static void prof_collect(rpf_ctx *ctx) {  // the stream must be idle
    for (auto &sp : ctx->spans) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) {
            ctx->k_ms[sp.k] += ms;
            ctx->k_launches[sp.k]++;
        }
        ctx->free_events.push_back(sp.a);
        ctx->free_events.push_back(sp.b);
    }
    ctx->spans.clear();
}

static int reset_counters(rpf_ctx *ctx) {
    const unsigned long long init[4] = {0ull, 0ull, ~0ull, 0ull};
    CU(ctx, cudaMemcpyAsync(ctx->d_counters, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->pixels_filtered = 0;
    return RPF_OK;
}

// g[k] = (k+1)ln(k+1) - k ln k in extended precision, and its sequentially summed prefix.
static int ensure_gtab(rpf_ctx *ctx, int n) {
    if (n + 2 <= ctx->gtab_n) return RPF_OK;
    const int m = n + 2;
    std::vector<double> g(m), pre(m + 1), hln(m + 1);
    const int mq = m > rpf::KB_GSM ? m : rpf::KB_GSM;
    std::vector<unsigned long long> gq(mq, 0ull), hq(mq + 1, 0ull);  // fixed point, unit 2^-43 (rpf_general.cuh)
    for (int k = 0; k < m; ++k) {
        const long double a = k + 1, b = k;
        g[k] = static_cast<double>(a * logl(a) - (k ? b * logl(b) : 0.0L));
        hln[k] = k ? static_cast<double>(b * logl(b)) : 0.0;
    }
    for (int k = 0; k < mq; ++k) {
        const long double a = k + 1, b = k;
        gq[k] = static_cast<unsigned long long>(llroundl((a * logl(a) - (k ? b * logl(b) : 0.0L)) * 8796093022208.0L));
        hq[k + 1] = hq[k] + gq[k];
    }
    hln[m] = static_cast<double>(static_cast<long double>(m) * logl(static_cast<long double>(m)));
    double acc = 0.0;
    pre[0] = 0.0;
    for (int k = 0; k < m; ++k) {
        acc += g[k];
        pre[k + 1] = acc;
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));  // kernels in flight may still read the old tables
    cudaFree(ctx->d_gtab);
    cudaFree(ctx->d_gpre);
    cudaFree(ctx->d_hln);
    cudaFree(ctx->d_gq);
    cudaFree(ctx->d_hq);
    ctx->d_gtab = ctx->d_gpre = ctx->d_hln = nullptr;
    ctx->d_gq = ctx->d_hq = nullptr;
    ctx->gtab_n = 0;
    CU(ctx, cudaMalloc(&ctx->d_gtab, (m > rpf::KB_GSM ? m : rpf::KB_GSM) * sizeof(double)));
    CU(ctx, cudaMalloc(&ctx->d_gpre, (m + 1) * sizeof(double)));
    CU(ctx, cudaMalloc(&ctx->d_hln, (m + 1) * sizeof(double)));
    CU(ctx, cudaMemset(ctx->d_gtab, 0, (m > rpf::KB_GSM ? m : rpf::KB_GSM) * sizeof(double)));
    CU(ctx, cudaMemcpy(ctx->d_gtab, g.data(), m * sizeof(double), cudaMemcpyHostToDevice));
    CU(ctx, cudaMemcpy(ctx->d_gpre, pre.data(), (m + 1) * sizeof(double), cudaMemcpyHostToDevice));
    CU(ctx, cudaMemcpy(ctx->d_hln, hln.data(), (m + 1) * sizeof(double), cudaMemcpyHostToDevice));
    CU(ctx, cudaMalloc(&ctx->d_gq, mq * sizeof(unsigned long long)));
    CU(ctx, cudaMalloc(&ctx->d_hq, (mq + 1) * sizeof(unsigned long long)));
    CU(ctx, cudaMemcpy(ctx->d_gq, gq.data(), mq * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    CU(ctx, cudaMemcpy(ctx->d_hq, hq.data(), (mq + 1) * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    ctx->gtab_n = m;
    return RPF_OK;
}

static int create_sharded(rpf_ctx **out, const rpf_config *cfg, int ndev);
static const LayoutOps *find_layout(int n_random, int n_features);

extern "C" int rpf_create(rpf_ctx **out, const rpf_config *cfg) {
    if (!out || !cfg) return fail(nullptr, RPF_ERR_INVALID, "rpf_create: null argument");
    *out = nullptr;
    if (cfg->width <= 0 || cfg->height <= 0 || cfg->spp <= 0)
        return fail(nullptr, RPF_ERR_INVALID, "rpf_create: width/height/spp must be positive");
    if (1ull * cfg->width * cfg->height * cfg->spp >= (1ull << 32))
        return fail(nullptr, RPF_ERR_INVALID, "rpf_create: more than 2^32 samples per context");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, RPF_ERR_NO_DEVICE, "rpf_create: no CUDA device (this library has no CPU fallback)");
    if (cfg->n_gpus > 1) return create_sharded(out, cfg, ndev);
    if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, RPF_ERR_INVALID, "rpf_create: bad device ordinal");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess || prop.major != 10)
        return fail(nullptr, RPF_ERR_NO_DEVICE, "rpf_create: device %d is not an sm_100 part (kernels are built for sm_100a only)",
                    cfg->device);
    const LayoutOps *ops = find_layout(cfg->n_random, cfg->n_features);
    if (!ops)
        return fail(nullptr, RPF_ERR_UNSUPPORTED, "rpf_create: no kernels for %d random parameters and %d features in this build (layouts: "
                    "2/12 = the reference's record, 5/12, 5/15; rpf_layouts.cuh)", cfg->n_random, cfg->n_features);
    rpf_ctx *ctx = new rpf_ctx;
    ctx->cfg = *cfg;
    ctx->ops = ops;
    ctx->cfg.flags &= ~(RPF_FLAG_FORCE_GENERAL | RPF_FLAG_FORCE_LARGE);
    if (ctx->cfg.flags == 0) ctx->cfg.flags = RPF_FLAGS_DEFINED;
    ctx->S = 1ull * cfg->width * cfg->height * cfg->spp;
    ctx->npix = cfg->width * cfg->height;
    ctx->max_dyn_smem = kSmemLimit;
#define CUC(call)                                                                                        \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            fail(nullptr, RPF_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_));                 \
            rpf_destroy(ctx);                                                                            \
            return RPF_ERR_CUDA;                                                                         \
        }                                                                                                \
    } while (0)
    CUC(cudaSetDevice(cfg->device));
    CUC(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CUC(cudaEventCreate(&ctx->ev0));
    CUC(cudaEventCreate(&ctx->ev1));
    CUC(cudaMalloc(&ctx->d_planes, ops->nch * ctx->S * sizeof(float)));
    CUC(cudaMalloc(&ctx->d_col[0], 3 * ctx->S * sizeof(double)));
    CUC(cudaMalloc(&ctx->d_col[1], 3 * ctx->S * sizeof(double)));
    CUC(cudaMalloc(&ctx->d_pstat, 2ull * ops->nf * ctx->npix * sizeof(double)));
    CUC(cudaMalloc(&ctx->d_pbound, 2ull * ops->nf * ctx->npix * sizeof(float)));
    CUC(cudaMalloc(&ctx->d_counters, 4 * sizeof(unsigned long long)));
    if (cfg->debug) {
        CUC(cudaMalloc(&ctx->d_dbg_n, ctx->npix * sizeof(int)));
        CUC(cudaMalloc(&ctx->d_dbg_alpha, 3ull * ctx->npix * sizeof(double)));
        CUC(cudaMalloc(&ctx->d_dbg_beta, 1ull * ops->nf * ctx->npix * sizeof(double)));
        CUC(cudaMalloc(&ctx->d_dbg_wrc, 1ull * ctx->npix * sizeof(double)));
        CUC(cudaMemset(ctx->d_dbg_n, 0, ctx->npix * sizeof(int)));
    }
    {
        if (ops->configure(ctx) != RPF_OK) {
            g_create_err = ctx->err;
            rpf_destroy(ctx);
            return RPF_ERR_CUDA;
        }
        ctx->num_sms = prop.multiProcessorCount;
        CUC(cudaMalloc(&ctx->d_work, 2 * sizeof(unsigned int)));
        ctx->force_large = (cfg->flags & RPF_FLAG_FORCE_LARGE) || (getenv("RPF_FORCE_LARGE") && atoi(getenv("RPF_FORCE_LARGE")));
        ctx->force_general = ctx->force_large || (cfg->flags & RPF_FLAG_FORCE_GENERAL) ||
                             (getenv("RPF_FORCE_GENERAL") && atoi(getenv("RPF_FORCE_GENERAL")));
        if (const char *bw = getenv("RPF_BLEND_WIDE_MIN")) if (atoi(bw) > 0) ctx->blend_wide_min = atoi(bw);
        if (const char *mb = getenv("RPF_SCRATCH_MB")) if (atoll(mb) > 0) ctx->scratch_budget = static_cast<size_t>(atoll(mb)) << 20;
    }
#undef CUC
    int rc = reset_counters(ctx);
    if (rc != RPF_OK) {
        g_create_err = ctx->err;
        rpf_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return RPF_OK;
}

static int after_upload(rpf_ctx *ctx) {
    const size_t n = 3 * ctx->S;
    rpf::colors_to_f64_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_planes, ctx->S,
                                                                                          ctx->d_col[ctx->cur]);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    ctx->pstat_y0 = ctx->pstat_y1 = 0;
    ctx->rec16_valid = false;
    ctx->col4_valid = false;
    ctx->uploaded = true;
    return RPF_OK;
}

extern "C" int rpf_upload(rpf_ctx *ctx, const float *const planes[RPF_NUM_CHANNELS]) {
    if (!ctx || !planes) return fail(ctx, RPF_ERR_INVALID, "rpf_upload: null argument");
    if (!ctx->shards.empty()) {  // every band takes its rows (own + halo) of every plane: y is the slow index, a band is contiguous
        const size_t row = static_cast<size_t>(ctx->cfg.width) * ctx->cfg.spp;
        for (size_t g = 0; g < ctx->shards.size(); ++g) {
            const float *sub[32];
            for (int c = 0; c < ctx->ops->nch; ++c) {
                if (!planes[c]) return fail(ctx, RPF_ERR_INVALID, "rpf_upload: plane %d is null", c);
                sub[c] = planes[c] + ctx->shard_buf0[g] * row;
            }
            const int rc = rpf_upload(ctx->shards[g], sub);
            if (rc != RPF_OK) { ctx->err = ctx->shards[g]->err; return rc; }
        }
        ctx->uploaded = true;
        return RPF_OK;
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    for (int c = 0; c < ctx->ops->nch; ++c) {
        if (!planes[c]) return fail(ctx, RPF_ERR_INVALID, "rpf_upload: plane %d is null", c);
        CU(ctx, cudaMemcpyAsync(ctx->d_planes + c * ctx->S, planes[c], ctx->S * sizeof(float), cudaMemcpyHostToDevice,
                                ctx->stream));
    }
    return after_upload(ctx);
}

extern "C" int rpf_upload_device(rpf_ctx *ctx, const float *const d_planes[RPF_NUM_CHANNELS]) {
    if (!ctx || !d_planes) return fail(ctx, RPF_ERR_INVALID, "rpf_upload_device: null argument");
    if (!ctx->shards.empty()) return fail(ctx, RPF_ERR_UNSUPPORTED, "rpf_upload_device: not available on a context sharded over several devices");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    for (int c = 0; c < ctx->ops->nch; ++c) {
        if (!d_planes[c]) return fail(ctx, RPF_ERR_INVALID, "rpf_upload_device: plane %d is null", c);
        CU(ctx, cudaMemcpyAsync(ctx->d_planes + c * ctx->S, d_planes[c], ctx->S * sizeof(float),
                                cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return after_upload(ctx);
}

// Working colours <- planes[2..4] again (what rpf_upload leaves): a frame that stays resident can be filtered afresh
// without another upload. Stream ordered.
extern "C" int rpf_reset_colors(rpf_ctx *ctx) {
    if (!ctx) return RPF_ERR_INVALID;
    if (!ctx->uploaded) return fail(ctx, RPF_ERR_STATE, "rpf_reset_colors: no samples uploaded");
    if (!ctx->shards.empty()) {
        for (rpf_ctx *c : ctx->shards) {
            const int rc = rpf_reset_colors(c);
            if (rc != RPF_OK) { ctx->err = c->err; return rc; }
        }
        return RPF_OK;
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    const size_t n = 3 * ctx->S;
    rpf::colors_to_f64_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_planes, ctx->S, ctx->d_col[ctx->cur]);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    ctx->col4_valid = false;
    return RPF_OK;
}

extern "C" int rpf_upload_colors_f64(rpf_ctx *ctx, const double *rgb) {
    if (!ctx || !rgb) return fail(ctx, RPF_ERR_INVALID, "rpf_upload_colors_f64: null argument");
    if (!ctx->uploaded) return fail(ctx, RPF_ERR_STATE, "rpf_upload_colors_f64: call rpf_upload first");
    if (!ctx->shards.empty()) {
        const size_t row = static_cast<size_t>(ctx->cfg.width) * ctx->cfg.spp, S = row * ctx->cfg.height;
        for (size_t g = 0; g < ctx->shards.size(); ++g) {
            rpf_ctx *c = ctx->shards[g];
            CU(ctx, cudaSetDevice(c->cfg.device));
            for (int k = 0; k < 3; ++k)
                CU(ctx, cudaMemcpyAsync(c->d_col[c->cur] + k * c->S, rgb + k * S + ctx->shard_buf0[g] * row, c->S * sizeof(double),
                                        cudaMemcpyHostToDevice, c->stream));
            c->col4_valid = false;
        }
        return RPF_OK;
    }
    ctx->col4_valid = false;
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    CU(ctx, cudaMemcpyAsync(ctx->d_col[ctx->cur], rgb, 3 * ctx->S * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    return RPF_OK;
}


// ---- general path (rpf_general.cuh) ----------------------------------------------------------------------
static const int kGeneralMaxCandidates = 65535;  // u16 candidate numbers and histogram cells, u8 bins (nb <= 255)

// n ln n in units of 2^-38 for counts up to n (rpf_large.cuh)
static int ensure_hql(rpf_ctx *ctx, int n) {
    if (n + 2 <= ctx->hql_n) return RPF_OK;
    const int m = (n + 2 > rpf::KB_GSM ? n + 2 : rpf::KB_GSM) + 1;
    std::vector<unsigned long long> h(m, 0ull);
    for (int k = 2; k < m; ++k) {
        const long double a = k;
        h[k] = static_cast<unsigned long long>(llroundl(a * logl(a) * 274877906944.0L));
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->d_hql);
    ctx->d_hql = nullptr;
    ctx->hql_n = 0;
    CU(ctx, cudaMalloc(&ctx->d_hql, m * sizeof(unsigned long long)));
    CU(ctx, cudaMemcpy(ctx->d_hql, h.data(), m * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    ctx->hql_n = m;
    return RPF_OK;
}

template <class T>
static int ensure_buffer(rpf_ctx *ctx, T **ptr, size_t *have, size_t need) {
    if (need <= *have) return RPF_OK;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(*ptr);
    *ptr = nullptr;
    *have = 0;
    CU(ctx, cudaMalloc(ptr, need));
    *have = need;
    return RPF_OK;
}

// ---- the launch code of every layout ------------------------------------------------------------------------------------------
#define RPF_NS rpf
#define RPF_NR 2
#define RPF_NF 12
#define RPF_FULL 1
#include "rpf_launch.inl"
#undef RPF_NS
#undef RPF_NR
#undef RPF_NF
#undef RPF_FULL
#define RPF_NS rpf_r5f12
#define RPF_NR 5
#define RPF_NF 12
#define RPF_FULL 0
#include "rpf_launch.inl"
#undef RPF_NS
#undef RPF_NR
#undef RPF_NF
#undef RPF_FULL
#define RPF_NS rpf_r5f15
#define RPF_NR 5
#define RPF_NF 15
#define RPF_FULL 0
#include "rpf_launch.inl"
#undef RPF_NS
#undef RPF_NR
#undef RPF_NF
#undef RPF_FULL

static const LayoutOps *const kLayouts[] = {&rpf::kOps, &rpf_r5f12::kOps, &rpf_r5f15::kOps};

static const LayoutOps *find_layout(int n_random, int n_features) {
    if (n_random == 0) n_random = 2;      // 0 = the reference's record (sd.h:46)
    if (n_features == 0) n_features = 12;
    for (const LayoutOps *o : kLayouts)
        if (o->nr == n_random && o->nf == n_features) return o;
    return nullptr;
}

static int launch_rows(rpf_ctx *ctx, int box, double seed, int y0, int y1, int y2 = 0, int y3 = 0) {
    return ctx->ops->launch_rows(ctx, box, seed, y0, y1, y2, y3);
}

extern "C" int rpf_num_channels(const rpf_ctx *ctx) { return ctx && ctx->ops ? ctx->ops->nch : RPF_NUM_CHANNELS; }
extern "C" int rpf_num_pairs(const rpf_ctx *ctx) { return ctx && ctx->ops ? ctx->ops->n_joint : 96; }
extern "C" int rpf_layout_supported(int n_random, int n_features) { return find_layout(n_random, n_features) != nullptr; }

static int finish_timing(rpf_ctx *ctx) {
    if (ctx->timing_pending) {
        CU(ctx, cudaEventSynchronize(ctx->ev1));
        float ms = 0.f;
        CU(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        ctx->last_ms = ms;
        ctx->timing_pending = false;
    }
    return RPF_OK;
}

#include "rpf_bands.inl"

#define NOT_SHARDED(ctx, name)                                                                                         \
    if (!(ctx)->shards.empty())                                                                                        \
        return fail(ctx, RPF_ERR_UNSUPPORTED, name ": not available on a context sharded over several devices (n_gpus > 1)")

extern "C" int rpf_filter_rows(rpf_ctx *ctx, int box_size, double sigma_fc_seed, int y0, int y1) {
    if (!ctx) return RPF_ERR_INVALID;
    NOT_SHARDED(ctx, "rpf_filter_rows");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    int rc = finish_timing(ctx);
    if (rc != RPF_OK) return rc;
    CU(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    rc = launch_rows(ctx, box_size, sigma_fc_seed, y0, y1);
    if (rc != RPF_OK) return rc;
    CU(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->timing_pending = true;
    return RPF_OK;
}

// Two row ranges [y0,y1) and [y2,y3), y2 >= y1, through ONE set of launches (what the band schedule does with the rows next to
// the two edges of a band: rpf_bands.inl).
extern "C" int rpf_filter_rows2(rpf_ctx *ctx, int box_size, double sigma_fc_seed, int y0, int y1, int y2, int y3) {
    if (!ctx) return RPF_ERR_INVALID;
    NOT_SHARDED(ctx, "rpf_filter_rows2");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    int rc = finish_timing(ctx);
    if (rc != RPF_OK) return rc;
    CU(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    rc = launch_rows(ctx, box_size, sigma_fc_seed, y0, y1, y2, y3);
    if (rc != RPF_OK) return rc;
    CU(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->timing_pending = true;
    return RPF_OK;
}

extern "C" int rpf_set_subsampling(rpf_ctx *ctx, double keep_probability, unsigned long long rng_seed, int global_row0) {
    if (!ctx || !(keep_probability > 0.0)) return fail(ctx, RPF_ERR_INVALID, "rpf_set_subsampling: keep probability must be in (0, 1]");
    for (size_t g = 0; g < ctx->shards.size(); ++g)  // the RNG is keyed by GLOBAL coordinates: every band draws the same stream
        rpf_set_subsampling(ctx->shards[g], keep_probability, rng_seed, global_row0 + ctx->shard_buf0[g]);
    ctx->sub.keep_q24 = keep_probability >= 1.0 ? RPF_SUBSAMPLE_ALL : static_cast<unsigned>(keep_probability * (1 << 24));
    ctx->sub.seed = rng_seed;
    ctx->sub.row_g0 = global_row0;
    return RPF_OK;
}

extern "C" int rpf_set_paper_mode(rpf_ctx *ctx, const rpf_paper_opts *o) {
    if (!ctx) return RPF_ERR_INVALID;
    const rpf_paper_opts none{};
    if (!o) o = &none;
    if (!(o->w_epsilon >= 0.0) || !(o->min_sigma >= 0.0) || !(o->world_pos_sigmas >= 0.0))
        return fail(ctx, RPF_ERR_INVALID, "rpf_set_paper_mode: options must be non-negative numbers");
    for (rpf_ctx *c : ctx->shards) rpf_set_paper_mode(c, o);
    ctx->w_eps = o->w_epsilon;
    if (ctx->min_sigma != o->min_sigma || ctx->pos_sigmas != o->world_pos_sigmas) {
        ctx->min_sigma = o->min_sigma;
        ctx->pos_sigmas = o->world_pos_sigmas;
        ctx->pstat_y0 = ctx->pstat_y1 = 0;  // the per-pixel thresholds are rebuilt by the next pass
    }
    return RPF_OK;
}

extern "C" int rpf_hdr_clamp(rpf_ctx *ctx, double k_sigmas, int y0, int y1) {
    if (!ctx) return RPF_ERR_INVALID;
    if (!ctx->uploaded) return fail(ctx, RPF_ERR_STATE, "rpf_hdr_clamp: no samples uploaded");
    if (!(k_sigmas > 0.0)) return fail(ctx, RPF_ERR_INVALID, "rpf_hdr_clamp: k_sigmas must be positive");
    if (y1 <= 0) y1 = ctx->cfg.height;
    if (y0 < 0 || y1 > ctx->cfg.height || y0 >= y1) return fail(ctx, RPF_ERR_INVALID, "bad row range [%d,%d)", y0, y1);
    if (!ctx->shards.empty()) {  // local to a pixel: every band clamps the rows it owns
        for (size_t g = 0; g < ctx->shards.size(); ++g) {
            const int a = std::max(y0, ctx->shard_own0[g]), b = std::min(y1, ctx->shard_own0[g + 1]);
            if (a >= b) continue;
            const int rc = rpf_hdr_clamp(ctx->shards[g], k_sigmas, a - ctx->shard_buf0[g], b - ctx->shard_buf0[g]);
            if (rc != RPF_OK) { ctx->err = ctx->shards[g]->err; return rc; }
        }
        return RPF_OK;
    }
    ctx->col4_valid = false;
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    const long long n = 1ll * (y1 - y0) * ctx->cfg.width;
    rpf::hdr_clamp_kernel<<<static_cast<unsigned>((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_col[ctx->cur], ctx->S, ctx->cfg.width,
                                                                                      ctx->cfg.spp, y0, y1, k_sigmas);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    return RPF_OK;
}

extern "C" int rpf_swap_colors(rpf_ctx *ctx) {
    if (!ctx) return RPF_ERR_INVALID;
    NOT_SHARDED(ctx, "rpf_swap_colors");
    ctx->cur ^= 1;
    ctx->col4_valid = false;
    return RPF_OK;
}

// The loop of rpf.cpp:767-775. seeds == nullptr: seed_all for every pass.
static int run_filter(rpf_ctx *ctx, const int *box_sizes, const double *seeds, double seed_all, int n_iter) {
    if (!ctx->shards.empty()) {
        if (!ctx->uploaded) return fail(ctx, RPF_ERR_STATE, "rpf_filter: no samples uploaded");
        for (rpf_ctx *c : ctx->shards) {
            CU(ctx, cudaSetDevice(c->cfg.device));
            int rc = finish_timing(c);
            if (rc != RPF_OK) { ctx->err = c->err; return rc; }
            CU(ctx, cudaEventRecord(c->ev0, c->stream));
        }
        const int rc = sharded_filter(ctx, box_sizes, seeds, seed_all, n_iter);
        if (rc != RPF_OK) return rc;
        for (rpf_ctx *c : ctx->shards) {
            CU(ctx, cudaSetDevice(c->cfg.device));
            CU(ctx, cudaEventRecord(c->ev1, c->stream));
            c->timing_pending = true;
        }
        return RPF_OK;
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    int rc = finish_timing(ctx);
    if (rc != RPF_OK) return rc;
    CU(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if (ctx->band.world > 1) {
        if ((rc = attached_filter(ctx, box_sizes, seeds, seed_all, n_iter)) != RPF_OK) return rc;
    } else {
        for (int i = 0; i < n_iter; ++i) {
            rc = launch_rows(ctx, box_sizes[i], seeds ? seeds[i] : seed_all, 0, ctx->cfg.height);
            if (rc != RPF_OK) return rc;
            ctx->cur ^= 1;  // rpf.cpp:732
            ctx->col4_valid = false;
        }
    }
    CU(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->timing_pending = true;
    return RPF_OK;
}

extern "C" int rpf_filter(rpf_ctx *ctx, const int *box_sizes, int n_iter, double sigma_fc_seed) {
    if (!ctx || (n_iter > 0 && !box_sizes) || n_iter < 0) return fail(ctx, RPF_ERR_INVALID, "rpf_filter: bad arguments");
    return run_filter(ctx, box_sizes, nullptr, sigma_fc_seed, n_iter);
}

extern "C" int rpf_filter_sigmas(rpf_ctx *ctx, const int *box_sizes, const double *sigma_seeds, int n_iter) {
    if (!ctx || (n_iter > 0 && (!box_sizes || !sigma_seeds)) || n_iter < 0) return fail(ctx, RPF_ERR_INVALID, "rpf_filter_sigmas: bad arguments");
    return run_filter(ctx, box_sizes, sigma_seeds, 0.0, n_iter);
}

extern "C" int rpf_color_device_ptr(rpf_ctx *ctx, int which, double **out) {
    if (!ctx || !out || which < 0 || which > 1) return fail(ctx, RPF_ERR_INVALID, "rpf_color_device_ptr: bad arguments");
    NOT_SHARDED(ctx, "rpf_color_device_ptr");
    *out = ctx->d_col[ctx->cur ^ which];
    return RPF_OK;
}

extern "C" int rpf_stream(rpf_ctx *ctx, void **out) {
    if (!ctx || !out) return RPF_ERR_INVALID;
    NOT_SHARDED(ctx, "rpf_stream");
    *out = ctx->stream;
    return RPF_OK;
}

extern "C" int rpf_synchronize(rpf_ctx *ctx) {
    if (!ctx) return RPF_ERR_INVALID;
    if (!ctx->shards.empty()) {
        ctx->last_ms = 0.0;
        for (rpf_ctx *c : ctx->shards) {
            const int rc = rpf_synchronize(c);
            if (rc != RPF_OK) { ctx->err = c->err; return rc; }
            ctx->last_ms = std::max(ctx->last_ms, c->last_ms);  // the bands run side by side: the slowest one is the step
        }
        return RPF_OK;
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->copy_stream) CU(ctx, cudaStreamSynchronize(ctx->copy_stream));
    prof_collect(ctx);
    return finish_timing(ctx);
}

extern "C" int rpf_set_profiling(rpf_ctx *ctx, int on) {
    if (!ctx) return RPF_ERR_INVALID;
    int rc = rpf_synchronize(ctx);
    ctx->profiling = on != 0;
    for (rpf_ctx *c : ctx->shards) c->profiling = on != 0;
    return rc;
}

extern "C" int rpf_get_kernel_times(rpf_ctx *ctx, rpf_kernel_times_t *out) {
    if (!ctx || !out) return RPF_ERR_INVALID;
    int rc = rpf_synchronize(ctx);
    if (rc != RPF_OK) return rc;
    for (int k = 0; k < RPF_NUM_KERNELS; ++k) { out->ms[k] = ctx->k_ms[k]; out->launches[k] = ctx->k_launches[k]; }
    for (rpf_ctx *c : ctx->shards)  // summed over the bands (device-milliseconds, not elapsed time)
        for (int k = 0; k < RPF_NUM_KERNELS; ++k) { out->ms[k] += c->k_ms[k]; out->launches[k] += c->k_launches[k]; }
    return RPF_OK;
}

extern "C" int rpf_download_colors_f64(rpf_ctx *ctx, double *rgb) {
    if (!ctx || !rgb) return fail(ctx, RPF_ERR_INVALID, "rpf_download_colors_f64: null argument");
    if (!ctx->shards.empty()) {  // every band returns the rows it owns
        const size_t row = static_cast<size_t>(ctx->cfg.width) * ctx->cfg.spp, S = row * ctx->cfg.height;
        for (size_t g = 0; g < ctx->shards.size(); ++g) {
            rpf_ctx *c = ctx->shards[g];
            CU(ctx, cudaSetDevice(c->cfg.device));
            const size_t lo = static_cast<size_t>(ctx->shard_own0[g] - ctx->shard_buf0[g]) * row, n = static_cast<size_t>(ctx->shard_own0[g + 1] - ctx->shard_own0[g]) * row;
            for (int k = 0; k < 3; ++k)
                CU(ctx, cudaMemcpyAsync(rgb + k * S + ctx->shard_own0[g] * row, c->d_col[c->cur] + k * c->S + lo, n * sizeof(double),
                                        cudaMemcpyDeviceToHost, c->stream));
        }
        return rpf_synchronize(ctx);
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    CU(ctx, cudaMemcpyAsync(rgb, ctx->d_col[ctx->cur], 3 * ctx->S * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    return rpf_synchronize(ctx);
}

extern "C" int rpf_download_colors(rpf_ctx *ctx, float *r, float *g, float *b) {
    if (!ctx) return RPF_ERR_INVALID;
    if (!ctx->shards.empty()) {
        const size_t row = static_cast<size_t>(ctx->cfg.width) * ctx->cfg.spp;
        float *dst[3] = {r, g, b};
        for (size_t s = 0; s < ctx->shards.size(); ++s) {
            rpf_ctx *c = ctx->shards[s];
            CU(ctx, cudaSetDevice(c->cfg.device));
            float *tmp = reinterpret_cast<float *>(c->d_col[c->cur ^ 1]);
            const size_t n3 = 3 * c->S;
            rpf::colors_to_f32_kernel<<<static_cast<unsigned>((n3 + 255) / 256), 256, 0, c->stream>>>(c->d_col[c->cur], n3, tmp);
            c->launches++;
            CU(ctx, cudaGetLastError());
            const size_t lo = static_cast<size_t>(ctx->shard_own0[s] - ctx->shard_buf0[s]) * row, n = static_cast<size_t>(ctx->shard_own0[s + 1] - ctx->shard_own0[s]) * row;
            for (int k = 0; k < 3; ++k)
                if (dst[k])
                    CU(ctx, cudaMemcpyAsync(dst[k] + ctx->shard_own0[s] * row, tmp + k * c->S + lo, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
        }
        return rpf_synchronize(ctx);
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    // the "next" colour buffer is dead between passes: use its first 3*S floats as conversion scratch
    float *tmp = reinterpret_cast<float *>(ctx->d_col[ctx->cur ^ 1]);
    const size_t n = 3 * ctx->S;
    rpf::colors_to_f32_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_col[ctx->cur], n, tmp);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    float *dst[3] = {r, g, b};
    for (int k = 0; k < 3; ++k)
        if (dst[k])
            CU(ctx, cudaMemcpyAsync(dst[k], tmp + k * ctx->S, ctx->S * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    return rpf_synchronize(ctx);
}

extern "C" int rpf_download_debug(rpf_ctx *ctx, int *nbh, double *alpha, double *beta, double *wrc) {
    if (!ctx) return RPF_ERR_INVALID;
    if (!ctx->shards.empty()) {
        const size_t W = ctx->cfg.width;
        for (size_t g = 0; g < ctx->shards.size(); ++g) {
            rpf_ctx *c = ctx->shards[g];
            if (!c->d_dbg_n) return fail(ctx, RPF_ERR_STATE, "rpf_download_debug: context created without cfg.debug");
            CU(ctx, cudaSetDevice(c->cfg.device));
            const size_t lo = static_cast<size_t>(ctx->shard_own0[g] - ctx->shard_buf0[g]) * W, go = static_cast<size_t>(ctx->shard_own0[g]) * W,
                         n = static_cast<size_t>(ctx->shard_own0[g + 1] - ctx->shard_own0[g]) * W;
            if (nbh) CU(ctx, cudaMemcpyAsync(nbh + go, c->d_dbg_n + lo, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
            if (alpha) CU(ctx, cudaMemcpyAsync(alpha + 3 * go, c->d_dbg_alpha + 3 * lo, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
            if (beta) CU(ctx, cudaMemcpyAsync(beta + ctx->ops->nf * go, c->d_dbg_beta + ctx->ops->nf * lo, ctx->ops->nf * n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
            if (wrc) CU(ctx, cudaMemcpyAsync(wrc + go, c->d_dbg_wrc + lo, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        }
        return rpf_synchronize(ctx);
    }
    if (!ctx->d_dbg_n) return fail(ctx, RPF_ERR_STATE, "rpf_download_debug: context created without cfg.debug");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    const size_t n = ctx->npix;
    if (nbh) CU(ctx, cudaMemcpyAsync(nbh, ctx->d_dbg_n, n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (alpha) CU(ctx, cudaMemcpyAsync(alpha, ctx->d_dbg_alpha, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (beta) CU(ctx, cudaMemcpyAsync(beta, ctx->d_dbg_beta, ctx->ops->nf * n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (wrc) CU(ctx, cudaMemcpyAsync(wrc, ctx->d_dbg_wrc, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    return rpf_synchronize(ctx);
}

// Component-level probe ("histogram counts exact", MI known answers): one pixel's 96 mutual informations
// and 19 marginal histograms, written by whichever dependency path filters that pixel in the next passes.
extern "C" int rpf_set_debug_pixel(rpf_ctx *ctx, int x, int y) {
    if (!ctx) return RPF_ERR_INVALID;
    NOT_SHARDED(ctx, "rpf_set_debug_pixel");
    if (x < 0 || y < 0) { ctx->probe_pixel = -1; return RPF_OK; }
    if (x >= ctx->cfg.width || y >= ctx->cfg.height) return fail(ctx, RPF_ERR_INVALID, "rpf_set_debug_pixel: (%d,%d) outside the context", x, y);
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    if (!ctx->d_probe_mi) {
        CU(ctx, cudaMalloc(&ctx->d_probe_mi, (ctx->ops->n_joint + 2) * sizeof(double)));
        CU(ctx, cudaMalloc(&ctx->d_probe_marg, ctx->ops->nch * PROBE_BINS * sizeof(unsigned)));
    }
    CU(ctx, cudaMemsetAsync(ctx->d_probe_mi, 0, (ctx->ops->n_joint + 2) * sizeof(double), ctx->stream));
    CU(ctx, cudaMemsetAsync(ctx->d_probe_marg, 0, ctx->ops->nch * PROBE_BINS * sizeof(unsigned), ctx->stream));
    ctx->probe_pixel = static_cast<long long>(y) * ctx->cfg.width + x;
    return RPF_OK;
}

extern "C" int rpf_download_debug_mi(rpf_ctx *ctx, double *mi, int *n, int *bins, unsigned *marginals, int bins_cap) {
    if (!ctx) return RPF_ERR_INVALID;
    NOT_SHARDED(ctx, "rpf_download_debug_mi");
    if (!ctx->d_probe_mi || ctx->probe_pixel < 0) return fail(ctx, RPF_ERR_STATE, "rpf_download_debug_mi: no pixel selected (rpf_set_debug_pixel)");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    int rc = rpf_synchronize(ctx);
    if (rc != RPF_OK) return rc;
    const int nj = ctx->ops->n_joint;
    std::vector<double> h(nj + 2);
    CU(ctx, cudaMemcpy(h.data(), ctx->d_probe_mi, h.size() * sizeof(double), cudaMemcpyDeviceToHost));
    if (mi) memcpy(mi, h.data(), nj * sizeof(double));
    const int nb = static_cast<int>(h[nj + 1]);
    if (n) *n = static_cast<int>(h[nj]);
    if (bins) *bins = nb;
    if (marginals) {
        if (bins_cap < nb) return fail(ctx, RPF_ERR_INVALID, "rpf_download_debug_mi: %d bins, room for %d", nb, bins_cap);
        std::vector<unsigned> m(ctx->ops->nch * PROBE_BINS);
        CU(ctx, cudaMemcpy(m.data(), ctx->d_probe_marg, m.size() * sizeof(unsigned), cudaMemcpyDeviceToHost));
        for (int c = 0; c < ctx->ops->nch; ++c)
            for (int i = 0; i < bins_cap; ++i) marginals[c * bins_cap + i] = i < nb ? m[c * PROBE_BINS + i] : 0u;
    }
    return RPF_OK;
}

extern "C" int rpf_reconstruct_box(rpf_ctx *ctx, float *rgb) {
    if (!ctx || !rgb) return fail(ctx, RPF_ERR_INVALID, "rpf_reconstruct_box: null argument");
    NOT_SHARDED(ctx, "rpf_reconstruct_box");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    float *tmp = reinterpret_cast<float *>(ctx->d_col[ctx->cur ^ 1]);
    const int n = ctx->npix * 3;
    rpf::reconstruct_box_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_col[ctx->cur], ctx->S, ctx->npix, ctx->cfg.spp, tmp);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    CU(ctx, cudaMemcpyAsync(rgb, tmp, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    return rpf_synchronize(ctx);
}

extern "C" int rpf_reconstruct_film(rpf_ctx *ctx, const rpf_film_desc *film, const float *ray_weight, float *rgb) {
    if (!ctx || !film || !rgb) return fail(ctx, RPF_ERR_INVALID, "rpf_reconstruct_film: null argument");
    NOT_SHARDED(ctx, "rpf_reconstruct_film");
    if (!ctx->uploaded) return fail(ctx, RPF_ERR_STATE, "rpf_reconstruct_film: no samples uploaded");
    const long long pw = 1ll * film->pixel_x1 - film->pixel_x0, ph = 1ll * film->pixel_y1 - film->pixel_y0;
    if (pw <= 0 || ph <= 0 || pw * ph > (1ll << 30) || !(film->radius_x > 0.f) || !(film->radius_y > 0.f))
        return fail(ctx, RPF_ERR_INVALID, "rpf_reconstruct_film: empty pixel bounds or non-positive filter radius");
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    // scratch in the dead "next" colour buffer (3*S doubles): [ray weights S floats][table 256 floats][rgb 3*pw*ph floats]
    const size_t need = (ctx->S + 256 + 3ull * pw * ph) * sizeof(float);
    float *scratch = reinterpret_cast<float *>(ctx->d_col[ctx->cur ^ 1]);
    float *own = nullptr;
    if (need > 3 * ctx->S * sizeof(double)) {
        CU(ctx, cudaMalloc(&own, need));
        scratch = own;
    }
    rpf::FilmParams p{};
    p.planes = ctx->d_planes;
    p.col = ctx->d_col[ctx->cur];
    p.S = ctx->S;
    p.W = ctx->cfg.width; p.H = ctx->cfg.height; p.P = ctx->cfg.spp;
    p.sx0 = film->sample_x0; p.sy0 = film->sample_y0;
    p.px0 = film->pixel_x0; p.py0 = film->pixel_y0; p.px1 = film->pixel_x1; p.py1 = film->pixel_y1;
    p.rx = film->radius_x; p.ry = film->radius_y; p.scale = film->scale; p.max_lum = film->max_sample_luminance;
    float *d_table = scratch + ctx->S;
    p.table = d_table;
    p.rgb = d_table + 256;
    cudaError_t e = cudaMemcpyAsync(d_table, film->filter_table, 256 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && ray_weight) {
        e = cudaMemcpyAsync(scratch, ray_weight, ctx->S * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
        p.ray_weight = scratch;
    }
    if (e == cudaSuccess) {
        const unsigned n = static_cast<unsigned>(pw * ph);
        rpf::film_reconstruct_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(p);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb, p.rgb, 3ull * pw * ph * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    int rc = e == cudaSuccess ? rpf_synchronize(ctx) : fail(ctx, RPF_ERR_CUDA, "rpf_reconstruct_film: %s", cudaGetErrorString(e));
    if (own) { cudaStreamSynchronize(ctx->stream); cudaFree(own); }
    return rc;
}

extern "C" int rpf_get_stats(rpf_ctx *ctx, rpf_stats_t *out) {
    if (!ctx || !out) return RPF_ERR_INVALID;
    int rc = rpf_synchronize(ctx);
    if (rc != RPF_OK) return rc;
    if (!ctx->shards.empty()) {
        rpf_stats_t t{};
        t.nbh_min = 0x7fffffff;
        for (rpf_ctx *c : ctx->shards) {
            rpf_stats_t u{};
            if ((rc = rpf_get_stats(c, &u)) != RPF_OK) { ctx->err = c->err; return rc; }
            t.pixels_filtered += u.pixels_filtered; t.nbh_sum += u.nbh_sum; t.nan_channels += u.nan_channels;
            t.kernel_launches += u.kernel_launches; t.halo_bytes += u.halo_bytes;
            if (u.pixels_filtered) { t.nbh_min = std::min(t.nbh_min, u.nbh_min); t.nbh_max = std::max(t.nbh_max, u.nbh_max); }
            t.last_filter_ms = std::max(t.last_filter_ms, u.last_filter_ms);
        }
        if (!t.pixels_filtered) t.nbh_min = 0;
        t.halo_bytes += static_cast<long long>(ctx->halo_bytes);
        *out = t;
        return RPF_OK;
    }
    unsigned long long c[4];
    CU(ctx, cudaMemcpy(c, ctx->d_counters, sizeof c, cudaMemcpyDeviceToHost));
    out->pixels_filtered = ctx->pixels_filtered;
    out->nbh_sum = static_cast<long long>(c[0]);
    out->nan_channels = static_cast<long long>(c[1]);
    out->nbh_min = ctx->pixels_filtered ? static_cast<int>(c[2]) : 0;
    out->nbh_max = static_cast<int>(c[3]);
    out->kernel_launches = ctx->launches;
    out->last_filter_ms = ctx->last_ms;
    out->halo_bytes = static_cast<long long>(ctx->halo_bytes);
    return RPF_OK;
}

extern "C" int rpf_reset_stats(rpf_ctx *ctx) {
    if (!ctx) return RPF_ERR_INVALID;
    ctx->halo_bytes = 0;
    if (!ctx->shards.empty()) {
        for (rpf_ctx *c : ctx->shards) {
            const int rc = rpf_reset_stats(c);
            if (rc != RPF_OK) { ctx->err = c->err; return rc; }
        }
        return RPF_OK;
    }
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    int rc = rpf_synchronize(ctx);
    if (rc != RPF_OK) return rc;
    ctx->launches = 0;
    for (int k = 0; k < RPF_NUM_KERNELS; ++k) { ctx->k_ms[k] = 0.0; ctx->k_launches[k] = 0; }
    return reset_counters(ctx);
}

// ---- row bands: creation, attachment (rpf_bands.inl has the schedules) -------------------------------------------------
static void band_release(rpf_ctx *ctx) {
    if (ctx->nccl_comm) {
        if (NcclApi *nc = nccl_api()) nc->CommDestroy(static_cast<ncclComm_t>(ctx->nccl_comm));
        ctx->nccl_comm = nullptr;
    }
    if (ctx->ev_edge) cudaEventDestroy(ctx->ev_edge);
    if (ctx->ev_copy) cudaEventDestroy(ctx->ev_copy);
    if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    ctx->ev_edge = ctx->ev_copy = ctx->ev_start = nullptr;
    ctx->copy_stream = nullptr;
}

static int create_sharded(rpf_ctx **out, const rpf_config *cfg, int ndev) {
    const int G = cfg->n_gpus;
    if (G > RPF_MAX_GPUS) return fail(nullptr, RPF_ERR_INVALID, "rpf_create: n_gpus %d above RPF_MAX_GPUS", G);
    for (int g = 0; g < G; ++g) {
        if (cfg->device_ids[g] < 0 || cfg->device_ids[g] >= ndev)
            return fail(nullptr, RPF_ERR_INVALID, "rpf_create: device_ids[%d] = %d, the process sees %d device(s)", g, cfg->device_ids[g], ndev);
        for (int h = 0; h < g; ++h)
            if (cfg->device_ids[h] == cfg->device_ids[g]) return fail(nullptr, RPF_ERR_INVALID, "rpf_create: device %d listed twice", cfg->device_ids[g]);
    }
    const int max_box = cfg->max_box > 0 ? cfg->max_box : 55;
    if ((max_box & 1) == 0) return fail(nullptr, RPF_ERR_INVALID, "rpf_create: max_box must be odd");
    const LayoutOps *ops = find_layout(cfg->n_random, cfg->n_features);
    if (!ops) return fail(nullptr, RPF_ERR_UNSUPPORTED, "rpf_create: no kernels for %d random parameters and %d features in this build", cfg->n_random, cfg->n_features);
    rpf_ctx *par = new rpf_ctx;
    par->cfg = *cfg;
    par->ops = ops;
    par->S = 1ull * cfg->width * cfg->height * cfg->spp;
    par->npix = cfg->width * cfg->height;
    par->halo_rows = b_delta(max_box);
    par->shard_own0.resize(G + 1);
    const int plan_boxes[2] = {max_box, max_box};  // any schedule of boxes up to max_box: every band owns b_delta(max_box) rows
    if (band_plan(cfg->height, G, plan_boxes, 2, par->shard_own0.data()) != RPF_OK) {
        delete par;
        return fail(nullptr, RPF_ERR_INVALID, "rpf_create: %d rows cannot be split into %d bands of at least %d rows (box %d)", cfg->height, G,
                    b_delta(max_box), max_box);
    }
    for (int g = 0; g < G; ++g) {
        par->shard_buf0.push_back(std::max(0, par->shard_own0[g] - par->halo_rows));
        par->shard_buf1.push_back(std::min(cfg->height, par->shard_own0[g + 1] + par->halo_rows));
        rpf_config sub = *cfg;
        sub.n_gpus = 0;
        sub.device = cfg->device_ids[g];
        sub.height = par->shard_buf1[g] - par->shard_buf0[g];
        rpf_ctx *c = nullptr;
        int rc = rpf_create(&c, &sub);
        if (rc == RPF_OK && (rc = ensure_band_streams(c)) != RPF_OK) { g_create_err = c->err; rpf_destroy(c); }
        if (rc != RPF_OK) {
            for (rpf_ctx *d : par->shards) rpf_destroy(d);
            delete par;
            return rc;
        }
        c->sub.row_g0 = par->shard_buf0[g];
        par->shards.push_back(c);
    }
    // halos travel with cudaMemcpyPeerAsync; with peer access they go straight over NVLink, without it through the host
    for (int g = 0; g + 1 < G; ++g) {
        int can = 0;
        const int a = cfg->device_ids[g], b = cfg->device_ids[g + 1];
        if (cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) {
            cudaSetDevice(a); cudaDeviceEnablePeerAccess(b, 0);
            cudaSetDevice(b); cudaDeviceEnablePeerAccess(a, 0);
            cudaGetLastError();  // "already enabled" is fine
        }
    }
    *out = par;
    return RPF_OK;
}

extern "C" int rpf_band_plan(int height, int n_bands, const int *box_sizes, int n_iter, int *own0) {
    return band_plan(height, n_bands, box_sizes, n_iter, own0);
}

extern "C" int rpf_band_plan_weighted(int height, int n_bands, const int *box_sizes, int n_iter, const double *row_weight, int *own0) {
    return band_plan(height, n_bands, box_sizes, n_iter, own0, row_weight);
}

extern "C" int rpf_band_unique_id(void *out128) {
    if (!out128) return RPF_ERR_INVALID;
    NcclApi *nc = nccl_api();
    if (!nc) return fail(nullptr, RPF_ERR_UNSUPPORTED, "rpf_band_unique_id: NCCL is not available (%s)", nccl_api() ? "" : "dlopen libnccl.so.2 failed");
    ncclUniqueId id;
    if (nc->GetUniqueId(&id) != ncclSuccess) return fail(nullptr, RPF_ERR_CUDA, "ncclGetUniqueId failed");
    static_assert(sizeof id == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, sizeof id);
    return RPF_OK;
}

extern "C" int rpf_band_attach(rpf_ctx *ctx, const rpf_band *band, const void *nccl_unique_id) {
    if (!ctx || !band) return fail(ctx, RPF_ERR_INVALID, "rpf_band_attach: null argument");
    NOT_SHARDED(ctx, "rpf_band_attach");
    if (band->world < 1 || band->rank < 0 || band->rank >= band->world || band->own0 < band->buf0 || band->own1 <= band->own0 ||
        band->own1 > band->buf0 + ctx->cfg.height || band->buf0 < 0 || band->buf0 + ctx->cfg.height > band->global_height)
        return fail(ctx, RPF_ERR_INVALID, "rpf_band_attach: inconsistent band (rows [%d,%d) of a buffer holding [%d,%d), frame of %d rows)", band->own0,
                    band->own1, band->buf0, band->buf0 + ctx->cfg.height, band->global_height);
    band_release(ctx);
    ctx->band = *band;
    ctx->sub.row_g0 = band->buf0;
    if (band->world == 1) return RPF_OK;
    if (!nccl_unique_id) return fail(ctx, RPF_ERR_INVALID, "rpf_band_attach: null NCCL unique id");
    NcclApi *nc = nccl_api();
    if (!nc) { ctx->band.world = 1; return fail(ctx, RPF_ERR_UNSUPPORTED, "rpf_band_attach: NCCL is not available (dlopen libnccl.so.2 failed)"); }
    int rc = ensure_band_streams(ctx);
    if (rc != RPF_OK) return rc;
    ncclUniqueId id;
    memcpy(&id, nccl_unique_id, sizeof id);
    ncclComm_t comm = nullptr;
    CU(ctx, cudaSetDevice(ctx->cfg.device));
    NC(ctx, nc->CommInitRank(&comm, band->world, id, band->rank));
    ctx->nccl_comm = comm;
    return RPF_OK;
}

extern "C" int rpf_host_alloc(void **out, size_t bytes) {
    if (!out) return RPF_ERR_INVALID;
    *out = nullptr;
    const cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);
    return e == cudaSuccess ? RPF_OK : fail(nullptr, RPF_ERR_CUDA, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
}

extern "C" void rpf_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

// =====================================================================================================
// rpf_filter_pass_noalloc: one pass of the fused kernel with every byte of device scratch in file-scope arrays, for callers
// that may not allocate (CUDA graph capture, tight serving loops). No counterpart in the reference.
// Shares nothing with the context-based path above: same kernels (pixel_stats_kernel,
// rpf_pixel_kernel<false>), own scratch.
//
// The arrays below exist once per GPU that loads this library, until the process exits (12 MiB + 21 KiB). Every call on a
// device uses the SAME arrays: calls that overlap on the GPU (two streams) race; successive calls on one stream, or calls
// ordered by events, are fine. Nothing is kept between calls except the two constant tables.
static constexpr int kNoAllocMaxPixels = RPF_NOALLOC_MAX_PIXELS;
static constexpr int kNoAllocTable = kFusedHardMax + 2;  // g[k], k <= candidates + 1
static __device__ __align__(256) double g_noalloc_pstat[static_cast<size_t>(kNoAllocMaxPixels) * 2 * rpf::NF];  // per-pixel mean[12], 3 sigma[12]
static __device__ __align__(256) double g_noalloc_gtab[kNoAllocTable];       // (k+1) ln (k+1) - k ln k
static __device__ __align__(256) double g_noalloc_gpre[kNoAllocTable + 1];   // its sequential prefix sums
static __device__ __align__(256) unsigned long long g_noalloc_counters[4];   // sum N, NaN channels, min N, max N of the last call

// device code names the array directly; only the host needs cudaGetSymbolAddress
__global__ void noalloc_reset_kernel() {
    if (threadIdx.x < 4) g_noalloc_counters[threadIdx.x] = threadIdx.x == 2 ? ~0ull : 0ull;
}

namespace {
constexpr int kNoAllocMaxDevices = 64;
struct NoAllocDevice {
    bool ready = false;
    double *pstat = nullptr, *gtab = nullptr, *gpre = nullptr;  // DEVICE addresses of the arrays above on this device
    unsigned long long *counters = nullptr;
};
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
std::mutex g_noalloc_mutex;
double g_noalloc_host_g[kNoAllocTable], g_noalloc_host_pre[kNoAllocTable + 1];  // static: no heap either
}  // namespace

// Once per device, with that device current: the arrays' device addresses (in host code the array names are host-side
// stand-ins, not device addresses), the constant tables, the kernel's shared-memory opt-in. None of it allocates.
static int noalloc_setup(int device, NoAllocDevice **out) {
    if (device < 0 || device >= kNoAllocMaxDevices) return fail(nullptr, RPF_ERR_INVALID, "rpf_filter_pass_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    NoAllocDevice &d = g_noalloc_dev[device];
    *out = &d;
    if (d.ready) return RPF_OK;
    cudaDeviceProp prop;
    CU(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(nullptr, RPF_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    CU(nullptr, cudaGetSymbolAddress(reinterpret_cast<void **>(&d.pstat), g_noalloc_pstat));
    CU(nullptr, cudaGetSymbolAddress(reinterpret_cast<void **>(&d.gtab), g_noalloc_gtab));
    CU(nullptr, cudaGetSymbolAddress(reinterpret_cast<void **>(&d.gpre), g_noalloc_gpre));
    CU(nullptr, cudaGetSymbolAddress(reinterpret_cast<void **>(&d.counters), g_noalloc_counters));
    // the same tables as ensure_gtab (extended precision, sequentially summed prefix)
    double acc = 0.0;
    g_noalloc_host_pre[0] = 0.0;
    for (int k = 0; k < kNoAllocTable; ++k) {
        const long double a = k + 1, b = k;
        g_noalloc_host_g[k] = static_cast<double>(a * logl(a) - (k ? b * logl(b) : 0.0L));
        acc += g_noalloc_host_g[k];
        g_noalloc_host_pre[k + 1] = acc;
    }
    CU(nullptr, cudaMemcpy(d.gtab, g_noalloc_host_g, sizeof g_noalloc_host_g, cudaMemcpyHostToDevice));
    CU(nullptr, cudaMemcpy(d.gpre, g_noalloc_host_pre, sizeof g_noalloc_host_pre, cudaMemcpyHostToDevice));
    CU(nullptr, cudaFuncSetAttribute(rpf::rpf_pixel_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    d.ready = true;
    return RPF_OK;
}

extern "C" int rpf_filter_pass_noalloc(const float *d_planes, const double *d_colors_in, double *d_colors_out, int width, int height,
                                       int spp, int box, double sigma_fc_seed, int flags, void *cuda_stream) {
    if (!d_planes || !d_colors_in || !d_colors_out) return fail(nullptr, RPF_ERR_INVALID, "rpf_filter_pass_noalloc: null buffer");
    if (d_colors_in == d_colors_out) return fail(nullptr, RPF_ERR_INVALID, "rpf_filter_pass_noalloc: a pass reads its neighbours' colours; in and out must differ");
    if (width <= 0 || height <= 0 || spp <= 0) return fail(nullptr, RPF_ERR_INVALID, "rpf_filter_pass_noalloc: bad frame %dx%dx%d", width, height, spp);
    if (box < 1 || !(box & 1)) return fail(nullptr, RPF_ERR_INVALID, "rpf_filter_pass_noalloc: box size %d must be odd and positive", box);
    if (!(sigma_fc_seed > 0.0)) return fail(nullptr, RPF_ERR_INVALID, "sigma_fc_seed must be positive");
    if (1ll * width * height > kNoAllocMaxPixels)
        return fail(nullptr, RPF_ERR_UNSUPPORTED, "rpf_filter_pass_noalloc: %d x %d = %lld pixels exceed the static scratch of %d pixels (RPF_NOALLOC_MAX_PIXELS); "
                    "filter in tiles or use rpf_create / rpf_filter", width, height, 1ll * width * height, kNoAllocMaxPixels);
    rpf::FilterParams p{};
    const long long ncap = 1ll * box * box * spp;
    if (!rpf::plan_smem(ncap, spp, kSmemLimit, p.sm))
        return fail(nullptr, RPF_ERR_UNSUPPORTED, "rpf_filter_pass_noalloc: box %d x %d spp = %lld candidates exceed the fused kernel's %d (rpf_fused_max_candidates); "
                    "larger neighbourhoods need the scratch of rpf_create / rpf_filter", box, spp, ncap, rpf_fused_max_candidates());
    int device = 0;
    CU(nullptr, cudaGetDevice(&device));
    cudaPointerAttributes at;
    CU(nullptr, cudaPointerGetAttributes(&at, d_planes));
    if (at.type != cudaMemoryTypeDevice || at.device != device)
        return fail(nullptr, RPF_ERR_INVALID, "rpf_filter_pass_noalloc: d_planes is not memory of the current device %d", device);
    NoAllocDevice *d = nullptr;
    int rc = noalloc_setup(device, &d);
    if (rc != RPF_OK) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const size_t S = static_cast<size_t>(width) * height * spp;
    p.f = rpf::FrameView{d_planes, S, width, height, spp};
    p.col_in = d_colors_in;
    p.col_out = d_colors_out;
    p.pstat = d->pstat;
    p.gtab = d->gtab;
    p.gpre = d->gpre;
    p.counters = d->counters;
    p.box = box;
    p.y0 = 0;
    p.y1 = height;
    p.flags = flags == 0 ? RPF_FLAGS_DEFINED : flags;  // as rpf_config.flags
    p.seed = sigma_fc_seed;
    p.probe = rpf::DebugProbe{-1, nullptr, nullptr};
    p.sub = rpf::SubsampleParams{0ull, RPF_SUBSAMPLE_ALL, 0};
    // the scratch keeps the last call's contents: the counters start over, the per-pixel statistics are rewritten for
    // every pixel this call reads (all width x height of them)
    noalloc_reset_kernel<<<1, 32, 0, stream>>>();
    const long long n = 1ll * width * height * rpf::NF;
    rpf::pixel_stats_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(p.f, 0, height, d->pstat, 0.0, 0.0);
    rpf::rpf_pixel_kernel<false><<<static_cast<unsigned>(width * height), rpf::THREADS, p.sm.total, stream>>>(p);
    CU(nullptr, cudaGetLastError());
    return RPF_OK;
}

#ifdef RPF_PHASE_TIMERS
// development build only: cycles per phase of dependency_kernel since the last call (thread 0 of every CTA; [6] = team 0
// waiting at the barrier after its fill, [7] = team 0's sweep)
extern "C" int rpf_debug_phase_cycles(rpf_ctx *ctx, unsigned long long *out8) {
    if (!ctx || !out8) return RPF_ERR_INVALID;
    int rc = rpf_synchronize(ctx);
    if (rc != RPF_OK) return rc;
    CU(ctx, cudaMemcpyFromSymbol(out8, rpf::g_phase_cycles, 8 * sizeof(unsigned long long)));
    unsigned long long z[8] = {};
    CU(ctx, cudaMemcpyToSymbol(rpf::g_phase_cycles, z, sizeof z));
    return RPF_OK;
}
#endif
