// rpf_large.cuh -- dependency estimation for neighbourhoods beyond 65,535 candidates (64 spp with boxes 35 / 55:
// 78,400 / 193,600 candidates, BASELINE.json configs[4]).
//
// Same arithmetic as dependency_kernel (rpf_general.cuh), other containers: 32-bit candidate numbers, u16 bins
// (bins = floor(sqrt(N)) passes 255 at N = 65,536, mi.cpp:53-58) and u32 histogram cells (a cell can hold all N samples).
// Candidate list and bins do not fit in shared memory any more and live in a per-CTA scratch in HBM (L2-resident: 42 B
// per candidate); shared memory holds the 19 marginals and one histogram region. A joint table of nb x nb u32 cells
// may exceed the region (nb = 440 -> 774 KB): it is then built in slabs of rows of the row channel, each slab a pass
// over the bins that only counts the samples whose row bin falls into the slab. When tables are small, up to four
// (one row channel against four column channels, the pair groups of rpf.cpp:416-442) are built per pass.
//
// Fixed-point entropy sums use the unit 2^-38 here (n ln n of 2^20 samples must stay below 2^64).
// Reference layout only (RPF_FULL in rpf_layouts.cuh): wider channel layouts take the general path.

namespace RPF_NS {

constexpr int KL_THREADS = 1024;
constexpr double KL_UNIT = 1.0 / 274877906944.0;  // 2^-38

struct LargeScratch {
    uint32_t *idx;      // [ncap] accepted candidate numbers
    uint16_t *bins;     // [19][cstride]
};

__device__ __forceinline__ LargeScratch large_scratch(const GeneralParams &p) {
    uint8_t *base = p.big_scratch + static_cast<size_t>(blockIdx.x) * p.big_stride;
    LargeScratch s;
    s.idx = reinterpret_cast<uint32_t *>(base);
    s.bins = reinterpret_cast<uint16_t *>(base + static_cast<size_t>(p.pl.nq) * 16);
    return s;
}

// Sum over the cells of `cells` u32 counters of n ln n (fixed point), zeroing nothing. Whole CTA; result added to *dst.
__device__ __forceinline__ void large_sweep(const GeneralParams &p, const unsigned *tab, int cells, const unsigned long long *gsm,
                                            unsigned long long *dst) {
    unsigned long long acc = 0ull;
    for (int k = threadIdx.x; k < cells; k += KL_THREADS) {
        const unsigned c = tab[k];
        if (c > 1u) acc += c < KB_GSM ? gsm[c] : p.hq[c];
    }
    acc = warp_sum_u64(acc);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(dst, acc);
}

// Rare: one pair recomputed cell by cell with the reference's expression and rounding (mi.cpp:66-87), slab by slab.
__device__ void large_exact_pair(const GeneralParams &p, KBSmall &ks, const LargeScratch &sc, unsigned *region, const unsigned *marg,
                                 int t, int N, int nb, int slab_rows) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nbs = p.pl.nbs;
    const size_t cs = static_cast<size_t>(p.pl.nq) * 4;
    int a, b;
    task_channels(t, a, b);
    const uint16_t *ba = sc.bins + a * cs, *bb = sc.bins + b * cs;
    const double dn = static_cast<double>(N);
    double sr = 0.0;
    for (int r0 = 0; r0 < nb; r0 += slab_rows) {
        const int rows = min(slab_rows, nb - r0);
        __syncthreads();
        for (int k = tid; k < rows * nb; k += KL_THREADS) region[k] = 0u;
        __syncthreads();
        for (int n = tid; n < N; n += KL_THREADS) {
            const unsigned ra = static_cast<unsigned>(ba[n]) - static_cast<unsigned>(r0);
            if (ra < static_cast<unsigned>(rows)) atomicAdd(&region[ra * nb + bb[n]], 1u);
        }
        __syncthreads();
        for (int cell = tid; cell < rows * nb; cell += KL_THREADS) {
            const unsigned n = region[cell];
            if (n == 0) continue;
            const int i = cell / nb, jj = cell - i * nb;
            const double pxv = __ddiv_rn(static_cast<double>(marg[a * nbs + r0 + i]), dn);
            const double pyv = __ddiv_rn(static_cast<double>(marg[b * nbs + jj]), dn);
            const double pxy = __ddiv_rn(static_cast<double>(n), dn);
            const double pp = __dmul_rn(pxv, pyv);
            if (pp != 0.0) sr = __dadd_rn(sr, __dmul_rn(pxy, log(__ddiv_rn(pxy, pp))));
        }
    }
    sr = warp_sum(sr);
    if (lane == 0) ks.redd[warp] = sr;
    __syncthreads();
    if (tid == 0) {
        double r = 0.0;
        for (int w = 0; w < KL_THREADS / 32; ++w) r += ks.redd[w];
        ks.mi[t] = r;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(KL_THREADS, 1) dependency_large_kernel(const GeneralParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ KBSmall ks;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = p.f.W, P = p.f.P, B = p.box;
    const GeneralPlan &L = p.pl;
    unsigned long long *gsm = reinterpret_cast<unsigned long long *>(smem + L.off_gsm);
    unsigned *marg = reinterpret_cast<unsigned *>(smem + L.off_marg);  // [19][nbs]
    unsigned *region = reinterpret_cast<unsigned *>(smem + L.off_region);
    const LargeScratch sc = large_scratch(p);
    const size_t cs = static_cast<size_t>(L.nq) * 4;  // bins per channel row of the scratch
    const int npix = p.nrows * W;

    for (int k = tid; k < KB_GSM; k += KL_THREADS) gsm[k] = p.hq[k];

    for (;;) {
        __syncthreads();
        if (tid == 0) ks.pixel = atomicAdd(p.work_counter, 1u);
        __syncthreads();
        if (static_cast<int>(ks.pixel) >= npix) break;
        const int pl = strip_order(static_cast<int>(ks.pixel), W, p.nrows);
        const int py = chunk_row(pl / W, p.row0, p.gap_at, p.gap), px = pl - (pl / W) * W;
        const size_t pix = static_cast<size_t>(py) * W + px;
        const uint32_t *mask = p.mask + static_cast<size_t>(pl) * L.mw;
        double *rec = p.rec + static_cast<size_t>(pl) * REC_DOUBLES;

        const int N = static_cast<int>(rec[REC_N]);
        if (N < p.large_min_n) continue;  // block-uniform: dependency_kernel has done this pixel
        int nb = static_cast<int>(__dsqrt_rn(static_cast<double>(N)));  // mi.cpp:53-58
        if (nb < 1) nb = 1;
        const double dnb = static_cast<double>(nb);
        kb_prepare_pixel(ks, rec, N, dnb, tid);
        for (int k = tid; k < N_JOINT; k += KL_THREADS) { ks.susp[k] = 0; ks.Sab[k] = 0ull; }
        for (int k = tid; k < NCH * L.nbs; k += KL_THREADS) marg[k] = 0u;

        // ---- accepted candidates -> idx (candidate numbers, any order) ---------------------------------------------
        {
            const int wpt = (L.mw + KL_THREADS - 1) / KL_THREADS;
            const int w0 = tid * wpt, w1 = min(w0 + wpt, L.mw);
            int cnt = 0;
            for (int w = w0; w < w1; ++w) cnt += __popc(mask[w]);
            int total;
            int off = block_exclusive_scan<KL_THREADS>(cnt, ks.wsum, &total);
            for (int w = w0; w < w1; ++w) {
                unsigned bits = mask[w];
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    sc.idx[off++] = static_cast<uint32_t>(w) * 32u + b;
                }
            }
        }
        __syncthreads();  // global writes of this CTA are visible to it after the barrier

        // ---- normalise + bin (u16 bins, channel-major), marginal counts ---------------------------------------------
        {
            const unsigned long long Q = static_cast<unsigned long long>(ks.Q), uN = static_cast<unsigned long long>(N);
            for (unsigned m = tid; m < static_cast<unsigned>(N); m += KL_THREADS) {
                const int t = static_cast<int>(sc.idx[(m * Q) % uN]);
                const uint32_t g = candidate_index(p.cd, t, px, py, B, P, W);
                const Rec16 r16 = load_rec16(p.rec16, g);
                const double4 cc = load_col4(p.col4, g);
                auto put = [&](int c, double x) {
                    const int b = bin_value(x, c, ks, nb, dnb, ks.tol[c]);
                    sc.bins[c * cs + m] = static_cast<uint16_t>(b);
                    atomicAdd(&marg[c * L.nbs + b], 1u);
                };
                put(CH_COL + 0, cc.x);
                put(CH_COL + 1, cc.y);
                put(CH_COL + 2, cc.z);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 f = r16.q[q];
                    put(slot_channel(4 * q + 0), f.x);
                    put(slot_channel(4 * q + 1), f.y);
                    put(slot_channel(4 * q + 2), f.z);
                    put(slot_channel(4 * q + 3), f.w);
                }
            }
        }
        __syncthreads();

        // ---- marginals: S_c = sum_i n_i ln n_i ----------------------------------------------------------------------
        for (int c = warp; c < NCH; c += KL_THREADS / 32) {
            unsigned long long s = 0ull;
            for (int i = lane; i < nb; i += 32) s += p.hq[marg[c * L.nbs + i]];
            s = warp_sum_u64(s);
            if (lane == 0) ks.Sm[c] = s;
        }

        // ---- 96 joint histograms, G tables (or one slab of one table) at a time ----------------------------------------
        const int cells = nb * nb;
        int G = L.region_words / cells;
        G = G >= 4 ? 4 : (G >= 2 ? 2 : 1);
        const int slab_rows = cells <= L.region_words ? nb : L.region_words / nb;
        for (int grp = 0; grp < 24; ++grp) {
            int a, t0, g4;
            unit_lookup(grp, 4, a, t0, g4);  // reference layout: 24 units of exactly four tables
            if (ks.flat[a]) continue;
            const uint16_t *ba = sc.bins + a * cs;
            for (int sub = 0; sub < 4; sub += G) {
                const uint16_t *bb[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    int aa, b;
                    task_channels(t0 + min(sub + j, 3), aa, b);
                    bb[j] = sc.bins + b * cs;
                }
                for (int r0 = 0; r0 < nb; r0 += slab_rows) {
                    const int rows = min(slab_rows, nb - r0);
                    const int tcells = rows * nb;
                    __syncthreads();
                    for (int k = tid; k < G * tcells; k += KL_THREADS) region[k] = 0u;
                    __syncthreads();
                    for (int n = tid; n < N; n += KL_THREADS) {
                        const unsigned ra = static_cast<unsigned>(ba[n]) - static_cast<unsigned>(r0);
                        if (ra < static_cast<unsigned>(rows)) {
                            const unsigned row = ra * nb;
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (j < G) atomicAdd(&region[j * tcells + row + bb[j][n]], 1u);
                        }
                    }
                    __syncthreads();
                    for (int j = 0; j < G; ++j) large_sweep(p, region + j * tcells, tcells, gsm, &ks.Sab[t0 + sub + j]);
                }
            }
        }
        __syncthreads();
        if (tid < N_JOINT) {
            int a, b;
            task_channels(tid, a, b);
            double mi = 0.0;
            if (!ks.flat[a] && !ks.flat[b]) {
                const double dn = static_cast<double>(N);
                const long long num = static_cast<long long>((ks.Sab[tid] + p.hq[N]) - (ks.Sm[a] + ks.Sm[b]));
                mi = static_cast<double>(num) * KL_UNIT / dn;
                if (mi != 0.0 && fabs(mi) < 1e-4 / (dn * dn) + 4.0 * static_cast<double>(nb) * nb * KL_UNIT / dn) {
                    ks.susp[tid] = 1;
                    ks.any_susp = 1;
                }
            }
            ks.mi[tid] = mi;
        }
        __syncthreads();
        if (ks.any_susp) {  // block-uniform, rare
            for (int t = 0; t < N_JOINT; ++t)
                if (ks.susp[t]) large_exact_pair(p, ks, sc, region, marg, t, N, nb, slab_rows);
        }

        cf_weights_from_mi(ks, p.flags, p.w_eps, tid);
        if (tid < NC) rec[REC_ALPHA + tid] = ks.alpha[tid];
        if (tid < NF) rec[REC_BETA + tid] = ks.beta[tid];
        if (tid == 0) rec[REC_WRC] = ks.wrc;
        if (p.probe.mi && static_cast<long long>(pix) == p.probe.pixel) {
            if (tid < N_JOINT) p.probe.mi[tid] = ks.mi[tid];
            if (tid == 0) { p.probe.mi[N_JOINT] = static_cast<double>(N); p.probe.mi[N_JOINT + 1] = static_cast<double>(nb); }
            for (int k = tid; k < NCH * PROBE_BINS; k += KL_THREADS) {
                const int c = k / PROBE_BINS, i = k - c * PROBE_BINS;
                p.probe.marg[k] = i < nb ? marg[c * L.nbs + i] : 0u;
            }
        }
        if (p.dbg_n) {
            if (tid == 0) { p.dbg_n[pix] = N; p.dbg_wrc[pix] = ks.wrc; }
            if (tid < NC) p.dbg_alpha[pix * NC + tid] = ks.alpha[tid];
            if (tid < NF) p.dbg_beta[pix * NF + tid] = ks.beta[tid];
        }
    }
}

}  // namespace RPF_NS
