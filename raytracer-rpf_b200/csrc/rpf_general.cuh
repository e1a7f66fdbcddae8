// rpf_general.cuh -- the general (any neighbourhood size) RPF path: four kernels per ApplyRPFFilter pass.
//
//   KA  membership_kernel + ordered_stats_kernel (rpf_membership.cuh): 3-sigma membership over the box -> per-pixel
//       acceptance bit mask, then N and the neighbourhood mean / sigma / min / max of all 19 channels  [rpf.cpp:556-601]
//   KB  dependency_kernel  normalise + bin every accepted sample, 96 joint histograms in shared memory
//                          (shared atomics), mutual information, alpha / beta / W_r_c        [rpf.cpp:603-618,356-488; mi.cpp]
//       (beyond 65,535 accepted samples: dependency_large_kernel, rpf_large.cuh)
//   KC  blend_kernel       cross-bilateral weights of the P own samples against the N accepted samples, blend
//       (16 spp and more: blend_wide_kernel, rpf_blend_wide.cuh)                               [rpf.cpp:627-717]
//
// The fused one-CTA-per-pixel kernel (rpf_pixel_kernel.cuh) keeps the whole neighbourhood in shared memory and
// lane-private histograms, which stops at N = 1279. Here nothing per-sample is resident: KA leaves one bit per
// candidate in HBM (candidate order = reference enumeration order xn, yn, sample: rpf.cpp:562-573), KB and KC
// regenerate the sample list from it. What must follow the reference's order -- the sequential FP64 sums behind
// mean and sigma (ops.h:111-144), whose rounding decides bin boundaries -- is done in KA by one lane per channel
// walking the accepted samples in that order (ordered_stats_kernel); everything else is order independent (integer counts) or differs
// from the CPU at the 1e-15 level (sums of logs, sums of weights).
// RE-INCLUDABLE (once per channel layout, rpf_layouts.cuh).

namespace RPF_NS {

// per-pixel record in HBM scratch: M, SD, raw min, raw max of every channel, N, alpha[3], beta[NF], W_r_c
constexpr int REC_M = 0, REC_SD = NCH, REC_RLO = 2 * NCH, REC_RHI = 3 * NCH, REC_N = 4 * NCH, REC_ALPHA = REC_N + 1,
              REC_BETA = REC_ALPHA + NC, REC_WRC = REC_BETA + NF;
constexpr int REC_DOUBLES = (REC_WRC + 1 + 7) & ~7;  // 96 for the reference layout
constexpr int KB_MAX_WARPS = 32, KB_GSM = 512;
constexpr int KC_THREADS = 128, KC_WPT = 2, KC_SEG_WORDS = KC_WPT * KC_THREADS, KC_IC = 8;  // KC_WPT mask words per thread and segment

struct GeneralPlan {
    int ncap;        // box*box*spp candidates
    int mw;          // mask words per pixel (multiple of 4)
    int nq;          // ceil(ncap/4): words per channel of packed u8 bins
    int nt;          // threads per CTA of the dependency kernel: 256 (two CTAs per SM) or 512
    int hw;          // words per joint histogram (two u16 cells per word), multiple of 4
    int nbs;         // words per marginal histogram
    int bins_in_smem;
    int bins_dyn;    // bins_in_smem == 0 only: decide PER PIXEL -- a pixel whose N accepted samples leave room for eight tables of
                     // its own size keeps its bins in shared memory (end of the histogram region), the others use bins_scratch
    int off_gsm, off_marg, off_region, off_bins, kb_smem;  // KB dynamic shared memory layout (bytes)
    int region_words;  // capacity of the histogram region in 32-bit words
    int kb_grid;
    int kc_smem;
    int idx_global;  // 1: box*box*spp > 65,535: candidate numbers are 32-bit and live in big_scratch; pixels with N > 65,535 are left
                     //    to dependency_large_kernel
    int large;       // 1: dependency_large_kernel (rpf_large.cuh): u32 candidates, u16 bins, u32 cells, slabbed tables
};

// Candidate number t -> (box column bx, box row by, sample s) without integer division: t / d = (t * ceil(2^40 / d)) >> 40,
// exact while t * d < 2^40 (t < 2^20 candidates, d = box*spp or spp < 2^20).
struct CandDecode {
    unsigned long long mBP, mP;
    int BP, P;
};
inline CandDecode make_cand_decode(int box, int spp) {
    CandDecode d;
    d.BP = box * spp;
    d.P = spp;
    d.mBP = ((1ull << 40) + d.BP - 1) / d.BP;
    d.mP = ((1ull << 40) + d.P - 1) / d.P;
    return d;
}
__device__ __forceinline__ void cand_split(const CandDecode &d, int t, int &bx, int &by, int &s) {
    bx = static_cast<int>((static_cast<unsigned long long>(t) * d.mBP) >> 40);
    const int r = t - bx * d.BP;
    by = static_cast<int>((static_cast<unsigned long long>(r) * d.mP) >> 40);
    s = r - by * d.P;
}

struct GeneralParams {
    FrameView f;
    CandDecode cd;
    const double *col_in;
    double *col_out;
    const double *pstat;
    const float4 *rec16;   // [S][NQ4]: the non-colour channels of each sample, slot order (rpf_kernels.cuh)
    const double4 *col4;   // [S]: current colours {r, g, b, 0}
    const double *rec19;   // [S][REC19]: all 19 channels as doubles, channel order (ordered_stats_kernel)
    // Fixed-point entropy tables (unit 2^-43): gq[k] = round(((k+1)ln(k+1) - k ln k) * 2^43), hq[n] = gq[0]+...+gq[n-1]
    // (~ n ln n). Integer sums are exact and order independent, so the atomics-ordered accumulation is deterministic.
    const unsigned long long *gq;
    const unsigned long long *hq;
    int box, flags;
    double seed;
    double w_eps;         // paper mode: epsilon of the W ratios (0 = reference)
    int row0, nrows;      // pixel rows of this chunk
    int gap_at, gap;      // chunk_row(): `gap` context rows are skipped after the chunk's first gap_at rows (none: gap = 0)
    uint32_t *mask;       // [chunk pixels][mw]
    double *rec;          // [chunk pixels][REC_DOUBLES]
    uint8_t *bins_scratch;  // [kb_grid][19][nq*4] when bins do not fit in shared memory
    uint8_t *big_scratch;   // rpf_large.cuh: [kb_grid][big_stride] candidate list + u16 bins of one pixel
    size_t big_stride;
    int large_min_n;        // dependency_large_kernel skips pixels with fewer accepted samples (0: takes all)
    unsigned int *work_counter;
    GeneralPlan pl;
    int *dbg_n;
    double *dbg_alpha, *dbg_beta, *dbg_wrc;
    unsigned long long *counters;
    DebugProbe probe;
};

// 256-bit gathers (LDG.E.256 on sm_100): a sample's 64-byte feature record in two requests instead of four, its
// {r, g, b, 0} double4 in one instead of two. Addresses are 32-byte aligned (cudaMalloc base + multiples of 32).
struct Rec16 { float4 q[NQ4]; };
__device__ __forceinline__ void ldg256(const float4 *p, float4 &a, float4 &b) {
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
__device__ __forceinline__ Rec16 load_rec16(const float4 *rec16, uint32_t g) {
    const float4 *p = rec16 + static_cast<size_t>(g) * NQ4;
    Rec16 r;
#pragma unroll
    for (int q = 0; q < NQ4; q += 2) ldg256(p + q, r.q[q], r.q[q + 1]);
    return r;
}
// the same record as plain 128-bit loads into a flat array (blend kernels: fewer live registers than the 256-bit form)
__device__ __forceinline__ void load_slots(const float4 *rec16, uint32_t g, float (&fs)[NQ4 * 4]) {
    const float4 *p = rec16 + static_cast<size_t>(g) * NQ4;
#pragma unroll
    for (int q = 0; q < (NSLOT + 3) / 4; ++q) {
        const float4 f = __ldg(p + q);
        fs[4 * q] = f.x; fs[4 * q + 1] = f.y; fs[4 * q + 2] = f.z; fs[4 * q + 3] = f.w;
    }
}
// Weight-space channel order is colour (3), features (NF), position (2) -- most selective first. Slot (rpf_common.cuh) of
// weight channel k >= 3: feature k - 3 sits behind the 2 positions and the NR random parameters, position k - 3 - NF in slot 0 / 1.
__host__ __device__ constexpr int weight_slot(int k) { return k < 3 + NF ? k + NR - 1 : k - (3 + NF); }
__device__ __forceinline__ double4 load_col4(const double4 *col4, uint32_t g) {
    double4 c;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(c.x), "=d"(c.y), "=d"(c.z), "=d"(c.w) : "l"(col4 + g));  // constant per launch
    return c;
}

__device__ __forceinline__ float warp_min_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// 64-bit warp sum from three hardware integer reductions (REDUX) over 22-bit slices: a slice sum of 32 lanes stays
// below 2^27, and the slices add back up exactly.
__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
    const unsigned s0 = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(v) & 0x3fffffu);
    const unsigned s1 = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(v >> 22) & 0x3fffffu);
    const unsigned s2 = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(v >> 44));
    return static_cast<unsigned long long>(s0) + (static_cast<unsigned long long>(s1) << 22) + (static_cast<unsigned long long>(s2) << 44);
}
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Work order of the general path. Work item w of a chunk of nrows rows -> chunk-local pixel number (row-major, the index of
// the pixel's mask and record in the scratch). The items run through COLUMN STRIPS of STRIP_W pixels, row-major inside a
// strip: the pixels in flight at any time then gather from a window of box rows x (STRIP_W + box) columns, about 13 MB of
// gather records at 8 spp, which stays L2-resident. Row-major over a full-HD row the window is box rows x 1920 columns =
// 81 MB: it falls out of L2 before the next row comes round, and every row of pixels re-reads its whole window from HBM
// (measured: 26 GB of DRAM traffic per dependency_kernel launch, profiles/r2/traffic.json v25).
constexpr int STRIP_W = 256;
__device__ __forceinline__ int strip_order_n(int w, int W, int nrows, int strip) {
    const int per = strip * nrows, s = w / per, r = w - s * per;
    const int x0 = s * strip, sw = min(strip, W - x0);
    const int y = r / sw;
    return y * W + x0 + (r - y * sw);
}
__device__ __forceinline__ int strip_order(int w, int W, int nrows) { return strip_order_n(w, W, nrows, STRIP_W); }
// Context row of chunk-local row r. A chunk may consist of two row ranges with `gap` rows left out after its first gap_at
// rows: the band schedule filters the rows next to BOTH band edges in one set of launches (rpf_bands.inl), which keeps the
// grids large enough to fill the GPU.
__device__ __forceinline__ int chunk_row(int r, int row0, int gap_at, int gap) { return row0 + r + (r >= gap_at ? gap : 0); }

// candidate t -> global sample index, or false when the neighbour pixel is out of bounds (rpf.cpp:569)
// Sample index of ACCEPTED candidate t of pixel (px, py). Mask bits are only ever set for candidates inside the image
// (membership_kernel), so there is no bound test; sample indices fit 32 bits (rpf_create checks S).
__device__ __forceinline__ uint32_t candidate_index(const CandDecode &cd, int t, int px, int py, int B, int P, int W) {
    const int bd = (B - 1) / 2;
    int bx, by, s;
    cand_split(cd, t, bx, by, s);
    return static_cast<uint32_t>((py - bd + by) * W + (px - bd + bx)) * static_cast<uint32_t>(P) + static_cast<uint32_t>(s);
}

// =====================================================================================================
// shared helpers for KB / KC

// Exclusive prefix of one int per thread over a CTA of NT threads; returns the prefix, total in *total.
template <int NT>
__device__ __forceinline__ int block_exclusive_scan(int v, int *wsum /* [NT/32 + 1] shared */, int *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < NT / 32 ? wsum[lane] : 0, winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += u;
        }
        if (lane < NT / 32) wsum[lane] = winc - w;
        if (lane == 31) wsum[NT / 32] = winc;
    }
    __syncthreads();
    *total = wsum[NT / 32];
    const int r = wsum[warp] + inc - v;
    __syncthreads();  // wsum may be reused right away
    return r;
}

#ifdef RPF_PHASE_TIMERS
// Development build only (lib/librpf_b200_timers.so): SM cycles per phase of dependency_kernel, summed over CTAs by thread 0.
__device__ unsigned long long g_phase_cycles[8];
#define PHASE_MARK(i) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_phase_cycles[i], static_cast<unsigned long long>(t_ - t_phase)); t_phase = t_; } } while (0)
#else
#define PHASE_MARK(i) do { } while (0)
#endif

struct KBSmall {
    double M[NCH], SD[NCH], lo[NCH], hi[NCH];
    double2 AB[NCH];                   // bin estimate te = x*A + Bc ({A, Bc}: one 16-byte load), trusted unless within
    double tol[NCH];                   // tol[c] of an integer; the binning loop uses the largest of the 19 for every channel
    int flat[NCH];
    unsigned long long Sm[NCH];        // sum n ln n of each marginal (fixed point)
    unsigned long long Sab[N_JOINT];   // sum n ln n of each joint table (fixed point)
    double redd[KB_MAX_WARPS];
    double mi[N_JOINT];
    int susp[N_JOINT];                 // pairs whose MI must be recomputed with the reference's own expression
    double alpha[NC], beta[NF], wrc;
    double wrf[NF], dcf[NF], dsum[9], wrck[NC];
    int wsum[KB_MAX_WARPS + 1];
    int N, nb, Q, any_susp;
    unsigned int pixel;
};

// mi.cpp:14-16 on top of sd.h:224-235: bin of raw value x of channel c. The reference value is
// trunc((((x-M)/SD) - lo)/(hi-lo) * nb) with every operation rounded; one FMA estimates it, and only when the estimate is
// too close to an integer for its error bound (or is NaN) do the reference's exact divisions run. A flat channel has
// A = 0, Bc = 0.5 (always bin 0); a NaN sigma makes A NaN (always the exact path).
__device__ __forceinline__ int bin_value(double x, int c, const KBSmall &ks, int nb, double dnb, double tol) {
    const double2 ab = ks.AB[c];
    const double te = fma(x, ab.x, ab.y);
    const double r = rint(te);
    int b;
    if (fabs(te - r) > tol) {
        b = __double2int_rz(te);
    } else {
        if (ks.flat[c]) return 0;
        const double z = normalise(x, ks.M[c], ks.SD[c]);
        b = __double2int_rz(__dmul_rn(__ddiv_rn(__dsub_rn(z, ks.lo[c]), __dsub_rn(ks.hi[c], ks.lo[c])), dnb));
    }
    return max(min(b, nb - 1), 0);
}

// Per-pixel constants of the dependency kernels: the slot permutation step, and per channel the normalised extremes and
// the FMA bin estimate te = x*A + Bc with its (absolute) trust tolerance. Threads 0..NCH-1 own a channel, thread 32 the permutation.
__device__ __forceinline__ void kb_prepare_pixel(KBSmall &ks, const double *rec, int N, double dnb, int tid) {
    if (tid == 32) {
        // slot -> sample permutation step: adjacent slots hold samples far apart in enumeration order, so the lanes
        // of a warp rarely hit one histogram cell together even for smooth channels
        int q = static_cast<int>(0.6180339887 * N) | 1;
        if (q >= N) q = 1;
        for (;; q += 2) {
            int a = q, b = N;
            while (b) { const int r = a % b; a = b; b = r; }
            if (a == 1) break;
        }
        ks.Q = N > 1 ? q % N : 0;
        if (N > 1 && ks.Q == 0) ks.Q = 1;
        ks.any_susp = 0;
    }
    if (tid < NCH) {
        const int c = tid;
        const double m = rec[REC_M + c], sd = rec[REC_SD + c], rlo = rec[REC_RLO + c], rhi = rec[REC_RHI + c];
        // (x-M)/SD is monotone in x, so the extremes of the normalised channel are the normalised extremes
        const double lo = normalise(rlo, m, sd), hi = normalise(rhi, m, sd);
        ks.M[c] = m; ks.SD[c] = sd; ks.lo[c] = lo; ks.hi[c] = hi;
        const bool flat = hi == lo;
        ks.flat[c] = flat;
        const double rr = dnb / (hi - lo), rsd = 1.0 / sd;
        double A = rsd * rr, Bc = (-m * rsd - lo) * rr;
        if (flat) { A = 0.0; Bc = 0.5; }
        if (!flat && !(hi > lo)) A = __longlong_as_double(0x7ff8000000000000ll);  // sigma is NaN: always the exact path
        ks.AB[c] = make_double2(A, Bc);
        // te lies in [0, nb] (lo / hi are the true extremes), so tol * (nb + 2) bounds the relative form tol * (|te| + 1)
        ks.tol[c] = (1e-9 + 1e-14 * (fabs(m * A) + fabs(Bc) + fabs(rhi * A) + fabs(rlo * A))) * (dnb + 2.0);
    }
}

// rpf.cpp:416-487 from the 96 mutual informations (same code path as phase 6 of the fused kernel)
__device__ __forceinline__ void cf_weights_from_mi(KBSmall &ks, int flags, double eps, int tid) {
    const bool p2 = flags & FLAG_P2;
    auto ratio = [&](double num, double den) {
        if (eps != 0.0) den = den + eps;  // paper mode (rpf_set_paper_mode); 0 = the reference's ratios
        return (p2 && den == 0.0) ? 0.0 : num / den;
    };
    if (tid < NF) {
        const int i = tid;
        double drf = 0.0, dpf = 0.0, dcf = 0.0;  // every sum starts at 0 and adds in index order, as the reference's loops do
        for (int q = 0; q < NR; ++q) drf += ks.mi[NPCOL * i + q];            // rpf.cpp:421
        for (int q = 0; q < NP; ++q) dpf += ks.mi[NPCOL * i + NR + q];       // rpf.cpp:425
        ks.wrf[i] = ratio(drf, drf + dpf);                                   // rpf.cpp:465
        for (int c = 0; c < NC; ++c) dcf += ks.mi[T0_COL + NCCOL * c + NPCOL + i];  // P1: sum over colours of MI(c_i, f_j)
        ks.dcf[i] = dcf;
    } else if (tid >= 32 && tid < 32 + NC) {
        const int i = tid - 32, t = T0_COL + NCCOL * i;
        double drc = 0.0, dpc = 0.0;
        for (int q = 0; q < NR; ++q) drc += ks.mi[t + q];                    // rpf.cpp:432
        for (int q = 0; q < NP; ++q) dpc += ks.mi[t + NR + q];               // rpf.cpp:436
        double dfc = 0.0;
        for (int j = 0; j < NF; ++j) dfc += ks.mi[t + NPCOL + j];            // rpf.cpp:440
        ks.dsum[i] = dfc;
        ks.dsum[3 + i] = drc;
        ks.dsum[6 + i] = dpc;
        const double w = ratio(drc, drc + dpc);                              // rpf.cpp:470
        ks.wrck[i] = w;
        ks.alpha[i] = 1.0 - w;                                               // rpf.cpp:475
    }
    __syncthreads();
    if (tid < NF) {
        const int i = tid;
        const double Dfc = ((0.0 + ks.dsum[0]) + ks.dsum[1]) + ks.dsum[2];   // rpf.cpp:449-456
        const double Drc = ((0.0 + ks.dsum[3]) + ks.dsum[4]) + ks.dsum[5];
        const double Dpc = ((0.0 + ks.dsum[6]) + ks.dsum[7]) + ks.dsum[8];
        const double num = (flags & FLAG_P1) ? ks.dcf[i] : (i < NC ? ks.dsum[i] : __longlong_as_double(0x7ff8000000000000ll));
        ks.beta[i] = (1.0 - ks.wrf[i]) * ratio(num, Dfc + Drc + Dpc);        // rpf.cpp:464,479
    } else if (tid == 32) {
        ks.wrc = (((0.0 + ks.wrck[0]) + ks.wrck[1]) + ks.wrck[2]) / NC;      // rpf.cpp:483-487
    }
    __syncthreads();
}

// Units of work of the joint phase: one row channel a against up to G column channels (G tables that share the row index).
// A feature has NPCOL partners, a colour NCCOL; each list is cut into chunks of G, the last one possibly shorter. With the
// reference layout and G = 4 these are the 24 groups {feature g: r0, r1, p0, p1} and {colour c, quarter q of its 16 partners}.
// Task index (rpf_common.cuh task_channels) of member j of unit u: t0 + j, j < g.
__host__ __device__ inline int unit_count(int G) { return NF * ((NPCOL + G - 1) / G) + NC * ((NCCOL + G - 1) / G); }
__device__ __forceinline__ void unit_lookup(int u, int G, int &a, int &t0, int &g) {
    const int cf = (NPCOL + G - 1) / G;
    if (u < NF * cf) {
        const int i = u / cf, c = u - i * cf;
        a = CH_FEAT + i;
        t0 = NPCOL * i + c * G;
        g = min(G, NPCOL - c * G);
    } else {
        u -= NF * cf;
        const int cc = (NCCOL + G - 1) / G, i = u / cc, c = u - i * cc;
        a = CH_COL + i;
        t0 = T0_COL + NCCOL * i + c * G;
        g = min(G, NCCOL - c * G);
    }
}


// fire-and-forget add on a 32-bit shared-memory address
__device__ __forceinline__ void smem_red_add(unsigned addr, unsigned v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

__device__ __forceinline__ void team_sync(int team, int nthreads) {
    if (nthreads == 32) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(nthreads) : "memory");
}

// Joint histograms of one pixel: fire-and-forget shared atomics build the tables (u16 cells, two per word), then the
// team sweeps each table once, adding n ln n per non-empty cell from a fixed-point table (unit 2^-43, exact integer
// sums, so the result does not depend on the order of the atomics) and zeroing it for its next use.
// MI(a,b) = (S_ab - S_a - S_b + N ln N)/N equals mi.cpp:79-87 up to rounding.
// The NT/32 warps form `teams` teams; a team owns `sets` x G histogram slots and takes every teams-th unit of work
// (one row channel against G column channels). With two slot sets a unit costs one team barrier: unit i+1 fills the
// other set while slower threads may still be sweeping the set of unit i.
template <int G, int NT>
__device__ __forceinline__ void joint_phase(const GeneralParams &p, KBSmall &ks, const uint32_t *bins, unsigned *region,
                                            const unsigned long long *hsm, int N, int nb, int hwn, int teams, int sets, int nq) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tw = (NT / 32) / teams, team = warp / tw;
    const int tn = tw * 32, tt = tid - team * tn;
    const int nfull = N >> 2, ntail = N & 3;
    // every unit holds exactly G tables when G divides both partner counts (always so for the reference layout): the
    // `j < ge` tests below then fold away at compile time
    constexpr bool FULL_UNITS = (NPCOL % G == 0) && (NCCOL % G == 0);
    // Table layout. With two or more tables per unit, tables 2i and 2i+1 share one table of 32-bit WORDS: cell c of the even
    // table is the low half of word c, cell c of the odd one the high half. Which half an update goes to is then a compile-time
    // constant (the increment is the immediate 1 or 0x10000 and the address one LEA from the cell number), where one table of
    // packed cells needs the cell's parity for a variable shift and a halved index: five instructions per update instead of ten.
    // Same bytes, same sweep volume. A single table (G = 1) keeps two cells per word.
    constexpr bool PAIRED = G >= 2;
    const int pw = 2 * hwn;  // words of a pair of tables (>= nb * nb)
    int set = 0;
    for (int u = team; u < unit_count(G); u += teams) {
        int a, t0, g;
        unit_lookup(u, G, a, t0, g);
        const int ge = FULL_UNITS ? G : g;
        if (ks.flat[a]) continue;  // a constant channel has MI exactly 0 with everything; Sab stays unused
        unsigned *th = region + (team * sets + set) * G * hwn;
        const unsigned th_s = static_cast<unsigned>(__cvta_generic_to_shared(th)), nb4 = 4u * nb;
        const uint32_t *ba = bins + a * nq;
        const uint32_t *bb[G];
#pragma unroll
        for (int j = 0; j < G; ++j) { int aa, b; task_channels(t0 + min(j, ge - 1), aa, b); bb[j] = bins + b * nq; }
        // the packed bins of the next round are requested before this round's atomics go out: when the bins live in the
        // L2-resident scratch (large neighbourhoods) their latency is what the loop would otherwise wait for
        uint32_t wa_n = 0u, wb_n[G];
#pragma unroll
        for (int j = 0; j < G; ++j) wb_n[j] = 0u;
        if (tt < nfull) {
            wa_n = ba[tt];
#pragma unroll
            for (int j = 0; j < G; ++j) wb_n[j] = bb[j][tt];
        }
        for (int q = tt; q < nfull; q += tn) {
            const uint32_t wa = wa_n;
            uint32_t wb[G];
#pragma unroll
            for (int j = 0; j < G; ++j) wb[j] = wb_n[j];
            if (q + tn < nfull) {
                wa_n = ba[q + tn];
#pragma unroll
                for (int j = 0; j < G; ++j) wb_n[j] = bb[j][q + tn];
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                // byte k of a packed word in one PRMT; an update is then PRMT + LEA (row + 4 * bin) + ATOMS
                const unsigned ca = __byte_perm(wa, 0u, 0x4440u + k) * nb;
                if (PAIRED) {
#pragma unroll
                    for (int jp = 0; jp < G / 2; ++jp) {
                        // 32-bit shared address of the row of the pair's table, one IMAD per (sample, pair)
                        unsigned row = __byte_perm(wa, 0u, 0x4440u + k) * nb4 + (th_s + jp * pw * 4);
                        asm volatile("" : "+r"(row));  // keep it: the compiler would otherwise refold (nb * a + b) * 4 + base per update
                        if (2 * jp < ge) smem_red_add(row + 4u * __byte_perm(wb[2 * jp], 0u, 0x4440u + k), 1u);
                        if (2 * jp + 1 < ge) smem_red_add(row + 4u * __byte_perm(wb[2 * jp + 1], 0u, 0x4440u + k), 0x10000u);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < G; ++j) {
                        if (j < ge) {
                            const unsigned cell = ca + ((wb[j] >> (8 * k)) & 255u);
                            atomicAdd(&th[j * hwn + (cell >> 1)], 1u << ((cell & 1u) << 4));
                        }
                    }
                }
            }
        }
        if (ntail && tt == 0) {
            const uint32_t wa = ba[nfull];
            for (int k = 0; k < ntail; ++k) {
                const unsigned ca = ((wa >> (8 * k)) & 255u) * nb;
#pragma unroll
                for (int j = 0; j < G; ++j) {
                    if (j < ge) {
                        const unsigned cell = ca + ((bb[j][nfull] >> (8 * k)) & 255u);
                        if (PAIRED) atomicAdd(&th[(j >> 1) * pw + cell], (j & 1) ? 0x10000u : 1u);
                        else atomicAdd(&th[j * hwn + (cell >> 1)], 1u << ((cell & 1u) << 4));
                    }
                }
            }
        }
#ifdef RPF_PHASE_TIMERS
        long long t_fill = 0;
        if (tid == 0) t_fill = clock64();
#endif
        team_sync(team, tn);
#ifdef RPF_PHASE_TIMERS
        if (tid == 0) { const long long t_ = clock64(); atomicAdd(&g_phase_cycles[6], static_cast<unsigned long long>(t_ - t_fill)); t_fill = t_; }
#endif
        // sweep: sum n ln n over the cells, zero them. Counts 0 and 1 contribute nothing (n ln n = 0) and skip the lookup.
        if (PAIRED) {
#pragma unroll
            for (int jp = 0; jp < G / 2; ++jp) {
                if (2 * jp >= ge) break;
                uint4 *r4 = reinterpret_cast<uint4 *>(th + jp * pw);
                unsigned long long acc0 = 0ull, acc1 = 0ull;  // even table: low halves, odd table: high halves
                const int nk = pw >> 2;
                uint4 vnext = tt < nk ? r4[tt] : make_uint4(0u, 0u, 0u, 0u);
                for (int k = tt; k < nk; k += tn) {
                    const uint4 v = vnext;  // the next cells are already on their way while these are looked up
                    if (k + tn < nk) vnext = r4[k + tn];
                    const unsigned any = v.x | v.y | v.z | v.w;
                    if (any == 0u) continue;
                    r4[k] = make_uint4(0u, 0u, 0u, 0u);
                    const unsigned wv[4] = {v.x, v.y, v.z, v.w};
                    if ((any & 0xfe00fe00u) == 0u) {  // every count below KB_GSM: shared-memory table only
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const unsigned c0 = wv[e] & 0xffffu, c1 = wv[e] >> 16;
                            // unconditional: hsm[0] = hsm[1] = 0, and the lanes' addresses are a handful of small ones either
                            // way (one wavefront); the test per cell cost more than it saved (profiles/r2/ab_v44_sweep_uncond.txt)
                            acc0 += hsm[c0];
                            acc1 += hsm[c1];
                        }
                    } else {
                        for (int e = 0; e < 4; ++e) {
                            const unsigned c0 = wv[e] & 0xffffu, c1 = wv[e] >> 16;
                            acc0 += p.hq[c0];
                            acc1 += p.hq[c1];
                        }
                    }
                }
                acc0 = warp_sum_u64(acc0);
                if (lane == 0 && acc0) atomicAdd(&ks.Sab[t0 + 2 * jp], acc0);
                if (2 * jp + 1 < ge) {
                    acc1 = warp_sum_u64(acc1);
                    if (lane == 0 && acc1) atomicAdd(&ks.Sab[t0 + 2 * jp + 1], acc1);
                }
            }
        } else
#pragma unroll
        for (int j = 0; j < G; ++j) {
            if (j >= ge) break;
            uint4 *r4 = reinterpret_cast<uint4 *>(th + j * hwn);
            unsigned long long acc = 0ull;
            const int nk = hwn >> 2;
            uint4 vnext = tt < nk ? r4[tt] : make_uint4(0u, 0u, 0u, 0u);
            for (int k = tt; k < nk; k += tn) {
                const uint4 v = vnext;  // the next cells are already on their way while these are looked up
                if (k + tn < nk) vnext = r4[k + tn];
                const unsigned any = v.x | v.y | v.z | v.w;
                if (any == 0u) continue;
                r4[k] = make_uint4(0u, 0u, 0u, 0u);
                const unsigned wv[4] = {v.x, v.y, v.z, v.w};
                // Predicated lookups on purpose: a quarter of the lanes take part, all at a handful of small addresses, so each costs
                // one shared-memory wavefront. Measured alternatives (profiles/r2/ab_v24.txt): one unconditional lookup per
                // PAIR of cells from a 256-entry table (fewer instructions, but 32 random 8-byte addresses per load) +2.5 %; counting
                // the cells equal to 2 and 3 in packed form without any table +6 %.
                if ((any & 0xfe00fe00u) == 0u) {  // every count below KB_GSM: shared-memory table only
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const unsigned c0 = wv[e] & 0xffffu, c1 = wv[e] >> 16;
                        if (c0 > 1u) acc += hsm[c0];
                        if (c1 > 1u) acc += hsm[c1];
                    }
                } else {
                    for (int e = 0; e < 4; ++e) {
                        const unsigned c0 = wv[e] & 0xffffu, c1 = wv[e] >> 16;
                        acc += p.hq[c0] + p.hq[c1];
                    }
                }
            }
            acc = warp_sum_u64(acc);
            if (lane == 0 && acc) atomicAdd(&ks.Sab[t0 + j], acc);
        }
        if (sets == 1) team_sync(team, tn);
        else set ^= 1;
#ifdef RPF_PHASE_TIMERS
        if (tid == 0) atomicAdd(&g_phase_cycles[7], static_cast<unsigned long long>(clock64() - t_fill));
#endif
    }
}

// Rare: recompute one pair cell by cell with the reference's expression and rounding (mi.cpp:66-87). Whole CTA.
template <int NT>
__device__ void exact_pair(const GeneralParams &p, KBSmall &ks, const uint32_t *bins, unsigned *region, const unsigned *marg,
                           int t, int N, int nb, int nq) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nbs = p.pl.nbs;
    int a, b;
    task_channels(t, a, b);
    const double dn = static_cast<double>(N);
    for (int k = tid; k < (nb * nb + 1) / 2; k += NT) region[k] = 0u;
    __syncthreads();
    const uint8_t *ba = reinterpret_cast<const uint8_t *>(bins + a * nq), *bb = reinterpret_cast<const uint8_t *>(bins + b * nq);
    for (int n = tid; n < N; n += NT) {
        const unsigned cell = ba[n] * nb + bb[n];
        atomicAdd(&region[cell >> 1], (cell & 1u) ? 0x10000u : 1u);
    }
    __syncthreads();
    double sr = 0.0;
    for (int cell = tid; cell < nb * nb; cell += NT) {
        const unsigned wv = region[cell >> 1];
        const unsigned n = (cell & 1) ? (wv >> 16) : (wv & 0xffffu);
        if (n == 0) continue;
        const int i = cell / nb, jj = cell - i * nb;
        const double pxv = __ddiv_rn(static_cast<double>(marg[a * nbs + i]), dn);
        const double pyv = __ddiv_rn(static_cast<double>(marg[b * nbs + jj]), dn);
        const double pxy = __ddiv_rn(static_cast<double>(n), dn);
        const double pp = __dmul_rn(pxv, pyv);
        if (pp != 0.0) sr = __dadd_rn(sr, __dmul_rn(pxy, log(__ddiv_rn(pxy, pp))));
    }
    sr = warp_sum(sr);
    if (lane == 0) ks.redd[warp] = sr;
    __syncthreads();
    if (tid == 0) {
        double r = 0.0;
        for (int w = 0; w < NT / 32; ++w) r += ks.redd[w];
        ks.mi[t] = r;
    }
    __syncthreads();
}

// =====================================================================================================
// KB: bins + joint histograms + MI + alpha/beta. Persistent CTAs of NT threads (512: two per SM, 1024: one per SM),
// 64 registers per thread so that 32 warps per SM hide the shared-memory latencies.
template <int NT>
__global__ void __launch_bounds__(NT, NT == 1024 ? 1 : 2) dependency_kernel(const GeneralParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ KBSmall ks;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = p.f.W, P = p.f.P, B = p.box;
    const GeneralPlan &L = p.pl;
    unsigned long long *gsm = reinterpret_cast<unsigned long long *>(smem + L.off_gsm);
    unsigned *marg = reinterpret_cast<unsigned *>(smem + L.off_marg);  // [19][nbs]
    unsigned *region = reinterpret_cast<unsigned *>(smem + L.off_region);
    unsigned short *idx16 = reinterpret_cast<unsigned short *>(region);
    uint32_t *idx32 = reinterpret_cast<uint32_t *>(p.big_scratch + static_cast<size_t>(blockIdx.x) * p.big_stride);
    uint32_t *const bins_fixed = L.bins_in_smem ? reinterpret_cast<uint32_t *>(smem + L.off_bins)
                                                : reinterpret_cast<uint32_t *>(p.bins_scratch + static_cast<size_t>(blockIdx.x) * NCH * L.nq * 4);
    const int npix = p.nrows * W;

    for (int k = tid; k < KB_GSM; k += NT) gsm[k] = p.hq[k];  // n ln n for small counts

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

        // ---- per-channel constants --------------------------------------------------------------------
#ifdef RPF_PHASE_TIMERS
        long long t_phase = clock64();
#endif
        const int N = static_cast<int>(rec[REC_N]);
        if (N > 65535) continue;  // block-uniform: u16 cells / u8 bins stop here; dependency_large_kernel takes the pixel
        int nb = static_cast<int>(__dsqrt_rn(static_cast<double>(N)));  // mi.cpp:53-58
        if (nb < 1) nb = 1;
        const double dnb = static_cast<double>(nb);
        // where this pixel's packed bins live (block-uniform): see GeneralPlan::bins_dyn
        uint32_t *bins = bins_fixed;
        int nq = L.nq, region_words = L.region_words;
        if (L.bins_dyn) {
            const int nqp = (N + 3) >> 2, hwp = (((nb * nb + 1) >> 1) + 3) & ~3;
            const int need = (NCH * nqp + 3) & ~3;  // words, 16-byte granular
            if (need + 8 * hwp <= L.region_words) {
                nq = nqp;
                region_words = L.region_words - need;
                bins = region + region_words;
            }
        }
        kb_prepare_pixel(ks, rec, N, dnb, tid);
        for (int k = tid; k < N_JOINT; k += NT) { ks.susp[k] = 0; ks.Sab[k] = 0ull; }
        for (int k = tid; k < NCH * L.nbs; k += NT) marg[k] = 0u;

        // ---- accepted candidates -> idx16 (candidate numbers, any order) ---------------------------------
        {
            const int wpt = (L.mw + NT - 1) / NT;
            const int w0 = tid * wpt;
            int cnt = 0;
            for (int w = w0; w < min(w0 + wpt, L.mw); ++w) cnt += __popc(mask[w]);
            int total;
            int off = block_exclusive_scan<NT>(cnt, ks.wsum, &total);
            for (int w = w0; w < min(w0 + wpt, L.mw); ++w) {
                unsigned bits = mask[w];
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    if (L.idx_global) idx32[off++] = static_cast<uint32_t>(w * 32 + b);
                    else idx16[off++] = static_cast<unsigned short>(w * 32 + b);
                }
            }
        }
        __syncthreads();
        PHASE_MARK(0);  // prepare + mask decode

        // ---- normalise + bin: one slot per thread, u8 bins (channel-major byte arrays), marginal counts ------------
        {
            const unsigned Q = static_cast<unsigned>(ks.Q), uN = static_cast<unsigned>(N);
            double tolmax = ks.tol[0];  // one trust tolerance for all channels: the largest (a NaN one is skipped; its channel's
#pragma unroll                          // estimate is NaN as well and takes the exact path anyway)
            for (int c = 1; c < NCH; ++c) tolmax = fmax(tolmax, ks.tol[c]);
            uint8_t *bin8 = reinterpret_cast<uint8_t *>(bins);
            const int cstride = nq * 4;
            for (unsigned m = tid; m < uN; m += NT) {
                const unsigned slot = (m * Q) % uN;
                const int t = L.idx_global ? static_cast<int>(idx32[slot]) : idx16[slot];
                const uint32_t g = candidate_index(p.cd, t, px, py, B, P, W);
                const Rec16 r16 = load_rec16(p.rec16, g);
                const double4 cc = load_col4(p.col4, g);
                auto put = [&](int c, double x) {
                    const int b = bin_value(x, c, ks, nb, dnb, tolmax);
                    bin8[c * cstride + m] = static_cast<uint8_t>(b);
                    atomicAdd(&marg[c * L.nbs + b], 1u);
                };
                put(CH_COL + 0, cc.x);
                put(CH_COL + 1, cc.y);
                put(CH_COL + 2, cc.z);
#pragma unroll
                for (int q = 0; q < NQ4; ++q) {
                    const float4 f = r16.q[q];
                    if (4 * q + 0 < NSLOT) put(slot_channel(4 * q + 0), f.x);
                    if (4 * q + 1 < NSLOT) put(slot_channel(4 * q + 1), f.y);
                    if (4 * q + 2 < NSLOT) put(slot_channel(4 * q + 2), f.z);
                    if (4 * q + 3 < NSLOT) put(slot_channel(4 * q + 3), f.w);
                }
            }
        }
        __syncthreads();
        PHASE_MARK(1);  // gather + binning

        // ---- marginals: S_c = sum_i n_i ln n_i ---------------------------------------------------------------------
        for (int c = warp; c < NCH; c += NT / 32) {
            unsigned long long s = 0ull;
            for (int i = lane; i < nb; i += 32) s += p.hq[marg[c * L.nbs + i]];
            s = warp_sum_u64(s);
            if (lane == 0) ks.Sm[c] = s;
        }
        // ---- the N_JOINT joint histograms (96 for the reference layout) ----------------------------------------------------------------------------------
        // The histogram region (idx16 is dead now) is cut into slots of the ACTUAL table size (nb*nb u16 cells).
        {
            const int hwn = (((nb * nb + 1) >> 1) + 3) & ~3;
            const int slots = region_words / hwn;
            int Gp, sets = 1, teams = 1;
            if (slots >= 4) {
                Gp = 4;
                // One set of tables per team; as many teams as fit (more independent teams hide more latency: 8 teams
                // of 2 warps beat 4 of 4 by 10-20 %), then as many tables per team as still fit. Multi-warp teams
                // synchronise on named barriers 1..15, so at most 8 of them.
                sets = 1;
                auto teams_for = [&](int g) {
                    int t = 1;
                    while (t * 2 <= NT / 32 && t * 2 * g <= slots) t *= 2;
                    if (t > 8 && t < NT / 32) t = 8;
                    return t;
                };
                const int t2 = teams_for(2), t4 = teams_for(4);
                if (t4 == t2) { Gp = 4; teams = t4; } else { Gp = 2; teams = t2; }
            } else Gp = slots >= 4 ? 4 : (slots >= 2 ? 2 : 1);  // few large tables: one team (splitting it does not pay, L2-bound)
            {
                uint4 *r4 = reinterpret_cast<uint4 *>(region);
                for (int k = tid; k < (teams * sets * Gp * hwn) >> 2; k += NT) r4[k] = make_uint4(0u, 0u, 0u, 0u);
            }
            __syncthreads();
            PHASE_MARK(2);  // marginal sums + table clear
            if (Gp == 4) joint_phase<4, NT>(p, ks, bins, region, gsm, N, nb, hwn, teams, sets, nq);
            else if (Gp == 2) joint_phase<2, NT>(p, ks, bins, region, gsm, N, nb, hwn, teams, sets, nq);
            else joint_phase<1, NT>(p, ks, bins, region, gsm, N, nb, hwn, teams, sets, nq);
        }
        __syncthreads();
        PHASE_MARK(3);  // joint tables: atomics + sweeps
        if (tid < N_JOINT) {
            int a, b;
            task_channels(tid, a, b);
            double mi = 0.0;
            if (!ks.flat[a] && !ks.flat[b]) {  // a constant channel has MI exactly 0 with everything (mi.cpp: every ratio is 1)
                const double dn = static_cast<double>(N);
                const long long num = static_cast<long long>((ks.Sab[tid] + p.hq[N]) - (ks.Sm[a] + ks.Sm[b]));
                mi = static_cast<double>(num) * (1.0 / 8796093022208.0) / dn;  // 2^-43
                // Exactly independent tables give exactly 0 in mi.cpp (every ratio is 1.0) but rounding noise here; the
                // P2 guards branch on such zeros, so suspiciously small values are recomputed with mi.cpp's own expression.
                if (mi != 0.0 && fabs(mi) < 1e-4 / (dn * dn)) { ks.susp[tid] = 1; ks.any_susp = 1; }
            }
            ks.mi[tid] = mi;
        }
        __syncthreads();
        if (ks.any_susp) {  // block-uniform, rare
            for (int t = 0; t < N_JOINT; ++t)
                if (ks.susp[t]) exact_pair<NT>(p, ks, bins, region, marg, t, N, nb, nq);
        }

        // ---- dependencies -> alpha, beta, W_r_c (rpf.cpp:416-487) ------------------------------------------
        cf_weights_from_mi(ks, p.flags, p.w_eps, tid);
        if (tid < NC) rec[REC_ALPHA + tid] = ks.alpha[tid];
        if (tid < NF) rec[REC_BETA + tid] = ks.beta[tid];
        if (tid == 0) rec[REC_WRC] = ks.wrc;
        if (p.dbg_n) {
            if (tid == 0) { p.dbg_n[pix] = N; p.dbg_wrc[pix] = ks.wrc; }
            if (tid < NC) p.dbg_alpha[pix * NC + tid] = ks.alpha[tid];
            if (tid < NF) p.dbg_beta[pix * NF + tid] = ks.beta[tid];
        }
        PHASE_MARK(4);  // MI, exact recomputes, weights
        if (p.probe.mi && static_cast<long long>(pix) == p.probe.pixel) {
            if (tid < N_JOINT) p.probe.mi[tid] = ks.mi[tid];
            if (tid == 0) { p.probe.mi[N_JOINT] = static_cast<double>(N); p.probe.mi[N_JOINT + 1] = static_cast<double>(nb); }
            for (int k = tid; k < NCH * PROBE_BINS; k += NT) {
                const int c = k / PROBE_BINS, i = k - c * PROBE_BINS;
                p.probe.marg[k] = i < nb ? marg[c * L.nbs + i] : 0u;
            }
        }
    }
}

// =====================================================================================================
// KC: cross-bilateral weights and blend (rpf.cpp:627-717). One CTA of 128 threads per pixel.
// w_ij = exp(-sum_p d^2/(2 sp2)) * exp(-sum_c alpha d^2/(2 s2)) * exp(-sum_f beta d^2/(2 s2)) on NORMALISED values;
// the three exponents are summed and each channel is pre-scaled by sqrt(coefficient)/sigma. Every term is >= 0, so
// once the running exponent passes 745.2 the weight is exactly 0 (exp underflows) and the rest is skipped.
struct KCSmall {
    double scl[NWZ], mz[NWZ];
    double red[4][KC_IC * 4];
    int wsum[KC_THREADS / 32 + 1];
};

#ifndef RPF_KC_MINB
#define RPF_KC_MINB 4  // CTAs per SM the register allocation aims at (128 registers)
#endif
__global__ void __launch_bounds__(KC_THREADS, RPF_KC_MINB) blend_kernel(const GeneralParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ KCSmall ks;
    __shared__ unsigned short segidx[KC_SEG_WORDS * 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = p.f.W, P = p.f.P, B = p.box;
    const size_t S = p.f.S;
    const int pl = strip_order(blockIdx.x, W, p.nrows);
    const int py = chunk_row(pl / W, p.row0, p.gap_at, p.gap), px = pl - (pl / W) * W;
    const size_t pix = static_cast<size_t>(py) * W + px;
    const uint32_t base = static_cast<uint32_t>(pix * P);
    const uint32_t *mask = p.mask + static_cast<size_t>(pl) * p.pl.mw;
    const double *rec = p.rec + static_cast<size_t>(pl) * REC_DOUBLES;
    double *zi = reinterpret_cast<double *>(smem);  // [P][ZS]: order colour(3), features(NF), position(2), pad

    // channel order inside z: k = 0..2 colour, 3..14 features, 15..16 position (most selective first)
    if (tid < NWZ) {
        const int k = tid;
        const int c = k < 3 ? CH_COL + k : (k < 3 + NF ? CH_FEAT + (k - 3) : CH_POS + (k - (3 + NF)));
        const double wrc = rec[REC_WRC];
        const double sigma_p = static_cast<double>(B / 4);                       // rpf.cpp:531 integer division
        const double s2 = p.seed * p.seed / (1.0 - wrc) / (1.0 - wrc);           // rpf.cpp:662
        double coef;
        if (k < 3) coef = rec[REC_ALPHA + k] / (2.0 * s2);
        else if (k < 3 + NF) coef = rec[REC_BETA + (k - 3)] / (2.0 * s2);
        else coef = 1.0 / (2.0 * (sigma_p * sigma_p));
        const double sd = rec[REC_SD + c];
        ks.scl[k] = sd == 0.0 ? 0.0 : sqrt(coef) / sd;
        ks.mz[k] = rec[REC_M + c];
    }
    __syncthreads();
    auto load_z = [&](uint32_t g, double *z, double *col) {
        float fs[NQ4 * 4];
        load_slots(p.rec16, g, fs);
        const double4 cc = load_col4(p.col4, g);
        col[0] = cc.x; col[1] = cc.y; col[2] = cc.z;
#pragma unroll
        for (int k = 0; k < NWZ; ++k) {
            // z order: colour, features (slots 4..15), position (slots 0, 1)
            const double x = k < 3 ? col[k] : static_cast<double>(fs[weight_slot(k)]);
            z[k] = scaled(x, ks.mz[k], ks.scl[k]);
        }
    };
    for (int n = tid; n < P; n += KC_THREADS) {
        double z[NWZ], col[NC];
        load_z(base + n, z, col);
#pragma unroll
        for (int k = 0; k < NWZ; ++k) zi[n * ZS + k] = z[k];
#pragma unroll
        for (int k = NWZ; k < ZS; ++k) zi[n * ZS + k] = 0.0;
    }
    __syncthreads();

    const int nseg = (p.pl.mw + KC_SEG_WORDS - 1) / KC_SEG_WORDS;
    for (int i0 = 0; i0 < P; i0 += KC_IC) {
        double sw[KC_IC], swc[KC_IC][NC];
#pragma unroll
        for (int ii = 0; ii < KC_IC; ++ii) { sw[ii] = 0.0; swc[ii][0] = swc[ii][1] = swc[ii][2] = 0.0; }
        const bool full_round = P - i0 >= KC_IC;
        for (int seg = 0; seg < nseg; ++seg) {
            unsigned wbits[KC_WPT];
            int cnt = 0;
#pragma unroll
            for (int k = 0; k < KC_WPT; ++k) {
                const int w = seg * KC_SEG_WORDS + tid * KC_WPT + k;
                wbits[k] = w < p.pl.mw ? mask[w] : 0u;
                cnt += __popc(wbits[k]);
            }
            int total;
            int off = block_exclusive_scan<KC_THREADS>(cnt, ks.wsum, &total);
#pragma unroll
            for (int k = 0; k < KC_WPT; ++k) {
                unsigned bits = wbits[k];
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    segidx[off++] = static_cast<unsigned short>((tid * KC_WPT + k) * 32 + b);
                }
            }
            __syncthreads();
            for (int n = tid; n < total; n += KC_THREADS) {
                const int t = seg * (KC_SEG_WORDS * 32) + segidx[n];
                const uint32_t g = candidate_index(p.cd, t, px, py, B, P, W);
                // colour terms first, for all own samples of this round: at the reference's sigma most neighbours are cut
                // by every own sample, and then the other 14 channels are never scaled (they are two thirds of the work)
                const double4 cc = load_col4(p.col4, g);
                const double cj[NC] = {cc.x, cc.y, cc.z};
                double zj[NWZ];
#pragma unroll
                for (int k = 0; k < 3; ++k) zj[k] = scaled(cj[k], ks.mz[k], ks.scl[k]);
                unsigned pass = 0u;
                if (full_round) {  // the usual case (spp a multiple of KC_IC): no bound test per own sample
#pragma unroll
                    for (int ii = 0; ii < KC_IC; ++ii) {
                        const double *zr = zi + (i0 + ii) * ZS;
                        double arg = 0.0;
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            const double d = zr[k] - zj[k];
                            arg = fma(d, d, arg);
                        }
                        if (!(arg > 745.2)) pass |= 1u << ii;  // NaN passes on to exp
                    }
                } else {
#pragma unroll
                    for (int ii = 0; ii < KC_IC; ++ii) {
                        if (i0 + ii < P) {
                            const double *zr = zi + (i0 + ii) * ZS;
                            double arg = 0.0;
#pragma unroll
                            for (int k = 0; k < 3; ++k) {
                                const double d = zr[k] - zj[k];
                                arg = fma(d, d, arg);
                            }
                            if (!(arg > 745.2)) pass |= 1u << ii;
                        }
                    }
                }
                // A neighbour whose raw colour is Inf or NaN poisons every own sample it is blended into even with weight 0
                // (rpf.cpp:692: 0 * inf = NaN), so it is never skipped: the cut is only a shortcut for finite colours.
                const bool finite_c = (cj[0] - cj[0] == 0.0) && (cj[1] - cj[1] == 0.0) && (cj[2] - cj[2] == 0.0);
                if (!finite_c) pass = (P - i0 >= KC_IC) ? ((1u << KC_IC) - 1u) : ((1u << (P - i0)) - 1u);
                if (pass == 0u) continue;
                // the 64-byte feature record is fetched only now: at the reference's sigma almost every neighbour is cut by its
                // colour terms alone, and its 32-byte colour record is then all that ever leaves L2
                float fs[NQ4 * 4];
                load_slots(p.rec16, g, fs);
#pragma unroll
                for (int k = 3; k < NWZ; ++k) {
                    const double x = static_cast<double>(fs[weight_slot(k)]);
                    zj[k] = scaled(x, ks.mz[k], ks.scl[k]);
                }
#pragma unroll
                for (int ii = 0; ii < KC_IC; ++ii) {
                    if ((pass >> ii) & 1u) {
                        const double *zr = zi + (i0 + ii) * ZS;
                        double arg = 0.0;
#pragma unroll
                        for (int k = 0; k < NWZ; ++k) {
                            const double d = zr[k] - zj[k];
                            arg = fma(d, d, arg);
                        }
                        const double wgt = arg > 745.2 ? 0.0 : exp(-arg);
                        sw[ii] += wgt;
                        swc[ii][0] = fma(wgt, cj[0], swc[ii][0]);
                        swc[ii][1] = fma(wgt, cj[1], swc[ii][1]);
                        swc[ii][2] = fma(wgt, cj[2], swc[ii][2]);
                    }
                }
            }
            __syncthreads();
        }
#pragma unroll
        for (int ii = 0; ii < KC_IC; ++ii) {
            const double a = warp_sum(sw[ii]), b0 = warp_sum(swc[ii][0]), b1 = warp_sum(swc[ii][1]), b2 = warp_sum(swc[ii][2]);
            if (lane == 0) {
                ks.red[warp][ii * 4 + 0] = a;
                ks.red[warp][ii * 4 + 1] = b0;
                ks.red[warp][ii * 4 + 2] = b1;
                ks.red[warp][ii * 4 + 3] = b2;
            }
        }
        __syncthreads();
        if (tid < KC_IC * NC) {
            const int ii = tid / NC, k = tid - ii * NC;
            if (i0 + ii < P) {
                const double a = ((ks.red[0][ii * 4] + ks.red[1][ii * 4]) + ks.red[2][ii * 4]) + ks.red[3][ii * 4];
                const double b = ((ks.red[0][ii * 4 + 1 + k] + ks.red[1][ii * 4 + 1 + k]) + ks.red[2][ii * 4 + 1 + k]) +
                                 ks.red[3][ii * 4 + 1 + k];
                const double c = b / a;  // rpf.cpp:700
                if (c != c) atomicAdd(&p.counters[1], 1ull);  // P4: count, never exit (rpf.cpp:702-705)
                p.col_out[static_cast<size_t>(k) * S + base + i0 + ii] = c;
            }
        }
        __syncthreads();
    }
}

}  // namespace RPF_NS
