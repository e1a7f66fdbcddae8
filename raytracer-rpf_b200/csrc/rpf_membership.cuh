// rpf_membership.cuh -- first stage of the general path, two kernels:
//
//   membership_kernel     3-sigma membership of every candidate of every pixel (rpf.cpp:556-586, ops.h:100-107) ->
//                         per-pixel acceptance bit mask in HBM, bit t = candidate t in reference enumeration order
//                         (xn ascending, yn ascending, sample: rpf.cpp:562-573). A CTA owns a TX x TY tile of pixels,
//                         one THREAD per pixel with that pixel's test in registers as one float interval per feature
//                         (three_sigma_interval, rpf_kernels.cuh: exactly the reference's double-precision test); the
//                         union of the tile's boxes streams through shared memory one column of neighbour pixels at a time,
//                         so a sample's 12 features are fetched from L2 once per tile instead of once per pixel.
//   ordered_stats_kernel  neighbourhood mean / sigma / min / max of all 19 channels (rpf.cpp:596-601). The sums must
//                         follow the reference's order (ops.h:111-144 sums sequentially; the rounding of mean and sigma
//                         decides histogram bin boundaries later), so one WARP per pixel walks the mask in order with
//                         one lane per channel and two dependent FP64 add chains per lane; 64 warps per SM hide the
//                         add latency.
// RE-INCLUDABLE (once per channel layout, rpf_layouts.cuh).

namespace RPF_NS {

constexpr int KM_MAX_THREADS = 128;
#ifndef RPF_KM_MINB
#define RPF_KM_MINB 6  // CTAs per SM the register allocation aims at
#endif

// Tile of pixels per CTA for a box size. A thread is idle while the column being streamed lies outside its own box: 8 columns
// of pixels keep 55 of the 62 columns of a box-55 tile's union busy (16 x 8: 55 of 70; 56.4 -> 54.2 ms at box 55, 26.7 -> 23.3
// at box 35, profiles/r2/ab_v37.txt), small boxes waste less on the halo with four rows.
__host__ __device__ inline void membership_tile(int box, int &tx, int &ty) {
    tx = 8;
    ty = box >= 13 ? 8 : 4;
}

struct MembershipParams {
    FrameView f;
    const float *pbound;   // [W*H][2 NF] per-pixel acceptance interval of every feature: lo[NF], hi[NF] (pixel_stats_kernel)
    const float4 *rec16;   // [S][NQ4] gather records: a pixel's samples are contiguous, a sample's features adjacent
    uint32_t *mask;        // [chunk pixels][mw]
    int box, row0, nrows, mw;
    int gap_at, gap;       // chunk_row() (rpf_general.cuh); gap_at is a multiple of the tile height: no tile straddles the gap
    int tx, ty, tiles_x;   // tile shape and tiles per row
    int col_rows;          // ty + box - 1: neighbour-pixel rows per column of the tile's union
    int rb_rows;           // rows staged in shared memory at a time (col_rows unless spp is very large)
    SubsampleParams sub;
};

constexpr int KM_NFP = (NF + 3) & ~3;         // floats per staged sample (whole float4s)
constexpr int KM_SLOT0 = NP + NR;             // slot of feature 0 in the gather record (channel_slot(CH_FEAT))
constexpr bool KM_VEC = KM_SLOT0 % 4 == 0 && NF % 4 == 0;  // the features are whole float4s of the record (reference layout)
__host__ __device__ inline int membership_row_floats(int spp) { return spp * KM_NFP + 4; }  // + 4: the 2-4 rows a warp reads at once
                                                                                             // fall into different banks, 16-byte aligned

// a candidate is outside iff some feature lies outside the pixel's [lo, hi]: features 0..3, then the rest
__device__ __forceinline__ bool outside_head(const float4 a, const float (&lo)[NF], const float (&hi)[NF]) {
    bool out = a.x < lo[0] || a.x > hi[0];
    if (NF > 1) out = out || a.y < lo[1 % NF] || a.y > hi[1 % NF];
    if (NF > 2) out = out || a.z < lo[2 % NF] || a.z > hi[2 % NF];
    if (NF > 3) out = out || a.w < lo[3 % NF] || a.w > hi[3 % NF];
    return out;
}
__device__ __forceinline__ bool outside_tail(const float4 *cp, bool out, const float (&lo)[NF], const float (&hi)[NF]) {
#pragma unroll
    for (int q = 1; q < KM_NFP / 4; ++q) {
        const float4 b = cp[q];
        const float v[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * q + j < NF) out = out || v[j] < lo[(4 * q + j) % NF] || v[j] > hi[(4 * q + j) % NF];
    }
    return out;
}

template <bool SUBSAMPLED>  // neighbour subsampling (include/rpf_subsample.h) compiled out of the default path
__global__ void __launch_bounds__(KM_MAX_THREADS, RPF_KM_MINB) membership_kernel(const MembershipParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    float *col = reinterpret_cast<float *>(smem);  // [rb_rows][P][KM_NFP] features of one column of neighbour pixels, as floats
    const int RS = membership_row_floats(p.f.P);
    constexpr int QS = KM_NFP / 4;  // float4s per staged sample
    const int tid = threadIdx.x;
    const int W = p.f.W, H = p.f.H, P = p.f.P, B = p.box, bd = (B - 1) / 2;
    // tiles in the same column strips as the other kernels of the path (strip_order, rpf_general.cuh): a strip's window of
    // feature rows stays L2-resident from one row of tiles to the next
    const int tiles_y = (p.nrows + p.ty - 1) / p.ty;
    const int tl = strip_order_n(blockIdx.x, p.tiles_x, tiles_y, STRIP_W / p.tx);
    const int tile_x = tl % p.tiles_x, tile_y = tl / p.tiles_x;
    const int x0 = tile_x * p.tx, y0 = chunk_row(tile_y * p.ty, p.row0, p.gap_at, p.gap);
    const int lx = tid % p.tx, ly = tid / p.tx;
    const int px = x0 + lx, py = y0 + ly;
    const bool live = px < W && tile_y * p.ty + ly < p.nrows;

    // the pixel's test in registers: a candidate fails iff some feature lies outside [lo, hi] (three_sigma_interval, rpf_kernels.cuh)
    float lo[NF], hi[NF];
    if (live) {
        const float *pb = p.pbound + (static_cast<size_t>(py) * W + px) * (2 * NF);
#pragma unroll
        for (int k = 0; k < NF; ++k) { lo[k] = pb[k]; hi[k] = pb[NF + k]; }
    } else {
#pragma unroll
        for (int k = 0; k < NF; ++k) { lo[k] = 0.f; hi[k] = 0.f; }
    }
    const uint64_t ckey = SUBSAMPLED ? rpf_subsample_centre_key(p.sub.seed, B, px, py + p.sub.row_g0) : 0ull;
    uint32_t *mrow = p.mask + (static_cast<size_t>(tile_y * p.ty + ly) * W + px) * p.mw;
    unsigned long long buf = 0ull;
    int nbits = 0, wi = 0;
    const int warp = tid >> 5, lane = tid & 31, nwarps = blockDim.x >> 5;

    const int nrows = p.col_rows;  // rows of neighbour pixels a column holds for this tile: yn = y0 - bd + r
    for (int xn = x0 - bd; xn <= x0 + p.tx - 1 + bd; ++xn) {
        // a column is staged in blocks of rb_rows rows (one block unless spp is very large); ascending blocks keep
        // every pixel's candidates in enumeration order
        for (int r0 = 0; r0 < nrows; r0 += p.rb_rows) {
            const int rbn = min(p.rb_rows, nrows - r0);
            __syncthreads();  // previous block fully consumed
            if (xn >= 0 && xn < W) {
                // one warp per row of the column: the P samples of a neighbour pixel are P consecutive gather records (64 P bytes)
                for (int r = warp; r < rbn; r += nwarps) {
                    const int yn = y0 - bd + r0 + r;
                    const bool in = yn >= 0 && yn < H;
                    const size_t g0 = (static_cast<size_t>(in ? yn : 0) * W + xn) * P;
                    if (KM_VEC) {
                        constexpr int Q = NF / 4;
                        float4 *dst = reinterpret_cast<float4 *>(col + r * RS);
                        const float4 *src = p.rec16 + g0 * NQ4 + KM_SLOT0 / 4;
                        for (int e = lane; e < P * Q; e += 32) {
                            const int s = e / Q, q = e - s * Q;
                            dst[e] = in ? __ldg(src + s * NQ4 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                    } else {
                        float *dst = col + r * RS;
                        const float *src = reinterpret_cast<const float *>(p.rec16 + g0 * NQ4) + KM_SLOT0;
                        for (int e = lane; e < P * KM_NFP; e += 32) {
                            const int s = e / KM_NFP, k = e - s * KM_NFP;
                            dst[e] = in && k < NF ? __ldg(src + s * (4 * NQ4) + k) : 0.f;
                        }
                    }
                }
            }
            __syncthreads();
            // this pixel's box covers columns px-bd .. px+bd and rows ly .. ly+B-1 of the column
            if (live && xn >= px - bd && xn <= px + bd) {
                const bool xin = xn >= 0 && xn < W;
                const int by_lo = max(0, r0 - ly), by_hi = min(B, r0 + rbn - ly);
                for (int by = by_lo; by < by_hi; ++by) {
                    const int yn = py - bd + by;
                    const bool inb = xin && yn >= 0 && yn < H;
                    const bool centre = xn == px && yn == py;
                    const float4 *cp = reinterpret_cast<const float4 *>(col + (ly + by - r0) * RS);
                    // a neighbour pixel's P acceptance bits are collected in a register and appended to the 64-bit
                    // bit buffer at once (chunks of 32 samples when spp is larger)
                    for (int s0 = 0; s0 < P; s0 += 32) {
                        const int ns = min(32, P - s0);
                        uint32_t bits = 0u;
                        if (!inb) {
                            cp += ns * QS;  // outside the image: no candidate
                        } else if (centre) {
                            bits = ns == 32 ? 0xffffffffu : (1u << ns) - 1u;  // the pixel's own samples always belong (rpf.cpp:558)
                            cp += ns * QS;
                        } else {
                            // allLessThan(|f - mu|, 3 sigma) (ops.h:100-107) as the exact float interval test. Four samples at a
                            // time: the first float4 of each (most candidates fail there), the other features only if one of
                            // the four is still in for this pixel.
                            int s = 0;
                            for (; s + 4 <= ns; s += 4, cp += 4 * QS) {
                                bool out[4];
#pragma unroll
                                for (int j = 0; j < 4; ++j) out[j] = outside_head(cp[j * QS], lo, hi);
                                if (NF > 4 && !(out[0] && out[1] && out[2] && out[3])) {
#pragma unroll
                                    for (int j = 0; j < 4; ++j) out[j] = outside_tail(cp + j * QS, out[j], lo, hi);
                                }
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    bool acc = !out[j];
                                    if (SUBSAMPLED && acc) acc = rpf_subsample_keep(ckey, p.sub.keep_q24, xn, yn + p.sub.row_g0, s0 + s + j);
                                    if (acc) bits |= 1u << (s + j);
                                }
                            }
                            for (; s < ns; ++s, cp += QS) {
                                bool out = outside_head(cp[0], lo, hi);
                                if (NF > 4 && !out) out = outside_tail(cp, out, lo, hi);
                                bool acc = !out;
                                if (SUBSAMPLED && acc) acc = rpf_subsample_keep(ckey, p.sub.keep_q24, xn, yn + p.sub.row_g0, s0 + s);
                                if (acc) bits |= 1u << s;
                            }
                        }
                        buf |= static_cast<unsigned long long>(bits) << nbits;
                        nbits += ns;
                        if (nbits >= 32) {
                            mrow[wi++] = static_cast<uint32_t>(buf);
                            buf >>= 32;
                            nbits -= 32;
                        }
                    }
                }
            }
        }
    }
    if (live) {
        if (nbits) mrow[wi++] = static_cast<uint32_t>(buf);
        for (; wi < p.mw; ++wi) mrow[wi] = 0u;
    }
}

// =====================================================================================================
constexpr int KO_WARPS = 4;  // pixels per CTA of the ordered statistics kernel
#ifndef RPF_KO_LIST
#define RPF_KO_LIST 1536
#endif
// capacity of a warp's sample list: mask words are decoded 32 at a time (1,024 candidates) until another 1,024 might not fit,
// then the list is walked -- a few hundred samples per walk at the reference's sigma instead of the few dozen that the 512
// candidates of a 16-word round used to yield (every walk starts with a batch of loads that nothing overlaps)
constexpr int KO_LIST = RPF_KO_LIST;
static_assert(KO_LIST >= 1024 + 32, "the list must hold one group of mask words");

__global__ void __launch_bounds__(KO_WARPS * 32) ordered_stats_kernel(const GeneralParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int W = p.f.W, P = p.f.P, B = p.box, bd = (B - 1) / 2;
    if (blockIdx.x * KO_WARPS + warp >= p.nrows * W) return;
    const int pl = strip_order(blockIdx.x * KO_WARPS + warp, W, p.nrows);
    const int py = chunk_row(pl / W, p.row0, p.gap_at, p.gap), px = pl - (pl / W) * W;
    const size_t pix = static_cast<size_t>(py) * W + px;
    const uint32_t base = static_cast<uint32_t>(pix * P);
    const uint32_t *mask = p.mask + static_cast<size_t>(pl) * p.pl.mw;

    // lane c < 19 owns channel c: its plane pointer, the running sums (ops.h:111-144 order) and extremes
    const int c = lane < NCH ? lane : 0;
    const double *pr = p.rec19 + c;  // all-double record: one 8-byte load per lane and sample
    double acc = 0.0, sq = 0.0;
    double lo = __longlong_as_double(0x7ff0000000000000ll), hi = __longlong_as_double(0xfff0000000000000ll);
    auto take = [&](double v) {
        acc = __dadd_rn(acc, v);
        sq = __dadd_rn(sq, __dmul_rn(v, v));
        lo = v < lo ? v : lo;
        hi = hi < v ? v : hi;
    };
    auto fetch = [&](uint32_t g) { return __ldg(pr + static_cast<size_t>(g) * REC19); };

    // The accepted samples are collected, in the reference's order, in a per-warp list of sample indices; every lane then
    // walks the list for its own channel, loads batched eight ahead of the dependent adds.
    __shared__ uint32_t glist[KO_WARPS][KO_LIST];
    uint32_t *gl = glist[warp];
    auto walk = [&](int cnt) {  // cnt > 0 entries, written by this warp
        __syncwarp();
        double v[8], u[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = fetch(gl[min(i, cnt - 1)]);
        for (int i0 = 0; i0 < cnt; i0 += 16) {
#pragma unroll
            for (int i = 0; i < 8; ++i) u[i] = fetch(gl[min(i0 + 8 + i, cnt - 1)]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i0 + i < cnt) take(v[i]);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = fetch(gl[min(i0 + 16 + i, cnt - 1)]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i0 + 8 + i < cnt) take(u[i]);
        }
        __syncwarp();  // the list is rewritten next
    };

    // own samples first, unconditionally and in order (rpf.cpp:558): they open the list
    int N = 0, cnt = 0;
    for (int s0 = 0; s0 < P; s0 += KO_LIST - 1024) {
        if (cnt) {
            N += cnt;
            walk(cnt);
        }
        cnt = min(KO_LIST - 1024, P - s0);
        for (int i = lane; i < cnt; i += 32) gl[i] = base + s0 + i;
    }

    // then the mask in candidate order, skipping the centre pixel (already taken): per mask word the 32 lanes decode one
    // candidate each and append the accepted sample indices (in order) to the list. The next 32 mask words are already
    // on their way while a group is decoded.
    const int BP = B * P;
    const int tc0 = (bd * B + bd) * P;  // candidate numbers of the centre pixel: [tc0, tc0 + P)
    const int nwords = (B * BP + 31) >> 5;
    unsigned myw_next = lane < nwords ? mask[lane] : 0u;
    for (int w0 = 0; w0 < nwords; w0 += 32) {
        const unsigned myw = myw_next;
        if (w0 + 32 < nwords) myw_next = w0 + 32 + lane < nwords ? mask[w0 + 32 + lane] : 0u;
        unsigned live = __ballot_sync(0xffffffffu, myw != 0u);
        while (live) {  // the non-empty words of the group, in order
            const int j = __ffs(live) - 1;
            live &= live - 1u;
            const unsigned bits = __shfl_sync(0xffffffffu, myw, j);
            const int t = (w0 + j) * 32 + lane;
            const bool on = ((bits >> lane) & 1u) && !(t >= tc0 && t < tc0 + P);
            const unsigned eff = __ballot_sync(0xffffffffu, on);
            if (on) gl[cnt + __popc(eff & ((1u << lane) - 1u))] = candidate_index(p.cd, t, px, py, B, P, W);
            cnt += __popc(eff);
        }
        if (cnt > KO_LIST - 1024) {  // another group might not fit
            N += cnt;
            walk(cnt);
            cnt = 0;
        }
    }
    if (cnt) {
        N += cnt;
        walk(cnt);
    }

    double *rec = p.rec + static_cast<size_t>(pl) * REC_DOUBLES;
    if (lane < NCH) {
        const double n = static_cast<double>(N);
        const double m = __ddiv_rn(acc, n);
        double var = __dsub_rn(__ddiv_rn(sq, n), __dmul_rn(m, m));
        if ((p.flags & FLAG_P3) && !(var > 0.0)) var = 0.0;  // P3: sqrt(max(0,var)); std::max(0.0, NaN) is 0 too (ops.h:141)
        rec[REC_M + lane] = m;
        rec[REC_SD + lane] = __dsqrt_rn(var);
        rec[REC_RLO + lane] = lo;
        rec[REC_RHI + lane] = hi;
    }
    if (lane == 0) {
        rec[REC_N] = static_cast<double>(N);
        atomicAdd(&p.counters[0], static_cast<unsigned long long>(N));
        atomicMin(&p.counters[2], static_cast<unsigned long long>(N));
        atomicMax(&p.counters[3], static_cast<unsigned long long>(N));
    }
}

}  // namespace RPF_NS
