// rpf_pixel_kernel.cuh -- K2: the per-pixel RPF kernel (one CTA of 128 threads per pixel).
// Phases (reference lines in brackets):
//   1 gather      3-sigma membership over the box, ordered compaction            [rpf.cpp:556-586]
//   2 stage       accepted samples' 19 channels -> shared memory (channel-major)
//   3 statistics  sequential sum / sum of squares, min / max per channel         [rpf.cpp:596-601, ops.h:111-144]
//   4 binning     normalise and bin every (sample, channel)                      [sd.h:224-235, mi.cpp:14-16]
//   5 MI          96 joint + 19 marginal histograms, thread-private, no atomics  [mi.cpp:23-42,65-87]
//   6 alpha/beta  dependency sums and fractional contributions                   [rpf.cpp:416-487]
//   7 weights     cross-bilateral weights and blend                              [rpf.cpp:637-717]
// Reference layout only (RPF_FULL in rpf_layouts.cuh): wider channel layouts take the general path.

namespace RPF_NS {

// Histogram cell address: lane-private column of a [words][32 lanes] array, four u8 cells per 32-bit word.
__device__ __forceinline__ uint8_t *cell_u8(unsigned char *colbase, int c) { return colbase + (c >> 2) * 128 + (c & 3); }

// Phase-5 accumulation for one lane. Cells are u8; a cell that passes 255 wraps and its index is remembered in one of
// K registers (at most N/256 wraps can happen), so the true count is low + 256 * (#remembered wraps of that cell).
template <int K>
__device__ __forceinline__ double accumulate_task(unsigned char *colbase, const uint8_t *ba, const uint8_t *bb, int sa, int sb,
                                                  int N, const double *gtab) {
    int ov[K > 0 ? K : 1];
#pragma unroll
    for (int k = 0; k < K; ++k) ov[k] = -1;
    int nov = 0;
    auto true_count = [&](int c, int low) {
#pragma unroll
        for (int k = 0; k < K; ++k) low += (ov[k] == c) ? 256 : 0;
        return low;
    };
    auto wrapped = [&](int c) {
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (k == nov) ov[k] = c;
        ++nov;
    };
    double acc = 0.0;
    int j = 0;
    for (; j + 1 < N; j += 2) {
        const int c0 = ba[j * 20] * sa + bb[j * 20] * sb;
        const int c1 = ba[j * 20 + 20] * sa + bb[j * 20 + 20] * sb;
        uint8_t *p0 = cell_u8(colbase, c0), *p1 = cell_u8(colbase, c1);
        const int l0 = *p0;
        int l1 = *p1;
        if (c0 == c1) l1 = (l0 + 1) & 255;  // both samples fall in one cell: the second sees the first's increment
        *p0 = static_cast<uint8_t>(l0 + 1);
        *p1 = static_cast<uint8_t>(l1 + 1);
        int o0 = l0, o1 = l1;
        if (K > 0) {
            o0 = true_count(c0, l0);
            if (l0 == 255) wrapped(c0);
            o1 = true_count(c1, l1);
            if (l1 == 255) wrapped(c1);
        }
        acc += gtab[o0];
        acc += gtab[o1];
    }
    if (j < N) {
        const int c0 = ba[j * 20] * sa + bb[j * 20] * sb;
        uint8_t *p0 = cell_u8(colbase, c0);
        const int l0 = *p0;
        *p0 = static_cast<uint8_t>(l0 + 1);
        acc += gtab[K > 0 ? true_count(c0, l0) : l0];
    }
    return acc;
}

template <bool SUBSAMPLED>  // neighbour subsampling (include/rpf_subsample.h) compiled out of the default path
__global__ void __launch_bounds__(THREADS, 4) rpf_pixel_kernel(const FilterParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ PixelSmall ps;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = p.f.W, H = p.f.H, P = p.f.P;
    const size_t S = p.f.S;
    const int pix_local = blockIdx.x;
    const int py = p.y0 + pix_local / W, px = pix_local - (pix_local / W) * W;
    const size_t pix = static_cast<size_t>(py) * W + px;
    const uint32_t base = static_cast<uint32_t>(pix * P);

    uint32_t *idx = reinterpret_cast<uint32_t *>(smem + p.sm.off_idx);
    uint8_t *bins = smem + p.sm.off_bins;                          // [n][20]
    double *gtab = reinterpret_cast<double *>(smem + p.sm.off_gtab);
    unsigned char *big = smem + p.sm.off_big;
    const int nstride = p.sm.nstride;
    float *rawf = reinterpret_cast<float *>(big + static_cast<size_t>(NC) * nstride * sizeof(double));  // [16][nstride]
    double *rawc = reinterpret_cast<double *>(big);                                                      // [3][nstride]

    // ---- phase 1: membership ---------------------------------------------------------------------
    if (tid < 2 * NF) {
        const double v = p.pstat[pix * 24 + tid];
        if (tid < NF) ps.mu[tid] = v; else ps.sd3[tid - NF] = v;
    }
    for (int s = tid; s < P; s += THREADS) idx[s] = base + s;  // own samples first, unconditionally (rpf.cpp:558)
    __syncthreads();
    const int B = p.box, bd = (B - 1) / 2, BP = B * P, ncand = B * BP;
    int nacc = P;
    for (int c0 = 0, it = 0; c0 < ncand; c0 += THREADS, ++it) {
        const int t = c0 + tid;
        bool acc = false;
        uint32_t g = 0;
        if (t < ncand) {
            // candidate order: xn ascending (outer), yn ascending, sample order (rpf.cpp:562-573)
            const int bx = t / BP, r = t - bx * BP, by = r / P, s = r - by * P;
            const int xn = px - bd + bx, yn = py - bd + by;
            if (xn >= 0 && xn < W && yn >= 0 && yn < H && !(xn == px && yn == py)) {
                g = static_cast<uint32_t>((static_cast<size_t>(yn) * W + xn) * P + s);
                float fv[NF];
#pragma unroll
                for (int k = 0; k < NF; ++k) fv[k] = __ldg(p.f.planes + static_cast<size_t>(CH_FEAT + k) * S + g);
                acc = true;
#pragma unroll
                for (int k = 0; k < NF; ++k) {
                    // allLessThan(|f - mu|, 3 sigma): reject on >= ; NaN sigma accepts (ops.h:100-107)
                    const double d = fabs(__dsub_rn(static_cast<double>(fv[k]), ps.mu[k]));
                    if (d >= ps.sd3[k]) acc = false;
                }
                if (SUBSAMPLED && acc)
                    acc = rpf_subsample_keep(rpf_subsample_centre_key(p.sub.seed, B, px, py + p.sub.row_g0), p.sub.keep_q24, xn,
                                             yn + p.sub.row_g0, s);
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, acc);
        if (lane == 0) ps.wcnt[it & 1][warp] = __popc(m);
        __syncthreads();
        int off = nacc, tot = 0;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int c = ps.wcnt[it & 1][w];
            if (w < warp) off += c;
            tot += c;
        }
        if (acc) idx[off + __popc(m & ((1u << lane) - 1u))] = g;
        nacc += tot;
    }
    const int N = nacc;
    int nb = static_cast<int>(__dsqrt_rn(static_cast<double>(N)));  // mi.cpp:53-58
    if (nb < 1) nb = 1;
    for (int k = tid; k <= N; k += THREADS) gtab[k] = p.gtab[k];
    __syncthreads();

    // ---- phase 2: stage the neighbourhood -------------------------------------------------------------
    for (int n = tid; n < N; n += THREADS) {
        const uint32_t g = idx[n];
#pragma unroll
        for (int s = 0; s < 16; ++s)
            rawf[s * nstride + n] = __ldg(p.f.planes + static_cast<size_t>(slot_channel(s)) * S + g);
#pragma unroll
        for (int k = 0; k < NC; ++k) rawc[k * nstride + n] = p.col_in[static_cast<size_t>(k) * S + g];
    }
    __syncthreads();

    // ---- phase 3: statistics ----------------------------------------------------------------------------
    // warp 0: sequential sum per channel; warp 1: sequential sum of squares (ops.h:111-144 order);
    // warps 2,3: min / max of the raw values (order independent).
    if (warp < 2) {
        if (lane < NCH) {
            const int c = lane;
            const bool isc = c >= CH_COL && c < CH_COL + NC;
            const double *pd = rawc + (isc ? c - CH_COL : 0) * nstride;
            const float *pf = rawf + (isc ? 0 : channel_slot(c)) * nstride;
            double acc = 0.0;
            if (warp == 0) {
                for (int n = 0; n < N; ++n) {
                    const double v = isc ? pd[n] : static_cast<double>(pf[n]);
                    acc = __dadd_rn(acc, v);
                }
                ps.sumx[c] = acc;
            } else {
                for (int n = 0; n < N; ++n) {
                    const double v = isc ? pd[n] : static_cast<double>(pf[n]);
                    acc = __dadd_rn(acc, __dmul_rn(v, v));
                }
                ps.sumsq[c] = acc;
            }
        }
    } else {
        const int cbeg = warp == 2 ? 0 : 10, cend = warp == 2 ? 10 : NCH;
        for (int c = cbeg; c < cend; ++c) {
            const bool isc = c >= CH_COL && c < CH_COL + NC;
            const double *pd = rawc + (isc ? c - CH_COL : 0) * nstride;
            const float *pf = rawf + (isc ? 0 : channel_slot(c)) * nstride;
            double lo = isc ? pd[0] : static_cast<double>(pf[0]), hi = lo;
            for (int n = lane; n < N; n += 32) {
                const double v = isc ? pd[n] : static_cast<double>(pf[n]);
                if (v < lo) lo = v;
                if (hi < v) hi = v;
            }
            lo = warp_min(lo);
            hi = warp_max(hi);
            if (lane == 0) { ps.rawlo[c] = lo; ps.rawhi[c] = hi; }
        }
    }
    __syncthreads();
    if (tid < NCH) {
        const int c = tid;
        const double n = static_cast<double>(N);
        const double m = __ddiv_rn(ps.sumx[c], n);
        double var = __dsub_rn(__ddiv_rn(ps.sumsq[c], n), __dmul_rn(m, m));
        if ((p.flags & FLAG_P3) && !(var > 0.0)) var = 0.0;  // P3: sqrt(max(0,var)); std::max(0.0, NaN) is 0 too (ops.h:141)
        const double sd = __dsqrt_rn(var);
        ps.M[c] = m;
        ps.SD[c] = sd;
        // (x-M)/SD is monotone in x, so the extremes of the normalised channel are the normalised extremes
        ps.lo[c] = normalise(ps.rawlo[c], m, sd);
        ps.hi[c] = normalise(ps.rawhi[c], m, sd);
    }
    if (tid == 0) { ps.N = N; ps.nb = nb; }
    __syncthreads();

    // ---- phase 4: normalise + bin (mi.cpp:14-16) ------------------------------------------------------
    // The reference bin is trunc(((x-M)/SD - lo)/(hi-lo) * nb) with every operation rounded. A reciprocal-multiply
    // estimate of that value is within ~1e-15*nb of it; only when the estimate is closer than 1e-9 to an integer
    // (or is NaN) can truncation differ, and only then are the exact divisions executed. Extremes are handled by
    // identity: the minimum gives exactly 0 and the maximum exactly nb, which clamps to nb-1.
    {
        const double dnb = static_cast<double>(nb);
        for (int c = 0; c < NCH; ++c) {
            const bool isc = c >= CH_COL && c < CH_COL + NC;
            const double lo = ps.lo[c], hi = ps.hi[c], m = ps.M[c], sd = ps.SD[c];
            const double rlo = ps.rawlo[c], rhi = ps.rawhi[c];
            const bool flat = hi == lo;
            const bool ordered = hi > lo;  // false when sigma is NaN: everything takes the exact path
            const double rsd = 1.0 / sd, rr = dnb / (hi - lo);
            const double *pd = rawc + (isc ? c - CH_COL : 0) * nstride;
            const float *pf = rawf + (isc ? 0 : channel_slot(c)) * nstride;
            for (int n = tid; n < N; n += THREADS) {
                const double x = isc ? pd[n] : static_cast<double>(pf[n]);
                int b = 0;
                if (!flat) {
                    if (ordered && x == rlo) b = 0;
                    else if (ordered && x == rhi) b = nb - 1;
                    else {
                        const double te = ((x - m) * rsd - lo) * rr;
                        const double r = rint(te);
                        if (ordered && fabs(te - r) > 1e-9 * (fabs(te) + 1.0)) {
                            b = static_cast<int>(te);
                        } else {
                            const double z = normalise(x, m, sd);
                            b = __double2int_rz(__dmul_rn(__ddiv_rn(__dsub_rn(z, lo), __dsub_rn(hi, lo)), dnb));
                        }
                        b = max(min(b, nb - 1), 0);
                    }
                }
                bins[n * 20 + c] = static_cast<uint8_t>(b);
            }
        }
    }
    __syncthreads();

    // ---- phase 5: joint + marginal histograms -------------------------------------------------------
    // Each lane owns one task (a pair of channels, or one channel's marginal) and a PRIVATE histogram:
    // column `lane` of a [words][32 lanes] array of 32-bit words, two u16 cells per word. Every lane only
    // ever touches bank `lane`, so the read-modify-write needs no atomics and has no bank conflicts.
    // The sum over cells of n*ln(n) is accumulated incrementally: raising a cell from k to k+1 adds
    // g[k] = (k+1)ln(k+1) - k ln k, so no pass over the cells is needed afterwards.
    // MI(a,b) = (S_ab - S_a - S_b + S_N)/N  with S = sum n ln n  (equals mi.cpp:79-87 up to rounding).
    if (warp < p.sm.mi_warps) {
        uint32_t *hw = reinterpret_cast<uint32_t *>(big) + static_cast<size_t>(warp) * p.sm.wpl * 32;
        unsigned char *colbase = reinterpret_cast<unsigned char *>(hw) + lane * 4;
        for (int grp = warp; grp < N_GROUP; grp += p.sm.mi_warps) {
            const int t = grp * 32 + lane;
            const bool active = t < N_TASK;
            int a = 0, b = -1;
            if (active) task_channels(t, a, b);
            int sa = nb, sb = 1;
            if (b < 0) { b = a; sa = 1; sb = 0; }
            const int ncell = grp == N_GROUP - 1 ? nb : nb * nb;
            const int nwords = (ncell + 3) >> 2;
            for (int w = 0; w < nwords; ++w) hw[w * 32 + lane] = 0u;
            __syncwarp();
            if (active) {
                const uint8_t *ba = bins + a, *bb = bins + b;
                double acc;
                switch (N >> 8) {
                    case 0: acc = accumulate_task<0>(colbase, ba, bb, sa, sb, N, gtab); break;
                    case 1: acc = accumulate_task<1>(colbase, ba, bb, sa, sb, N, gtab); break;
                    case 2: acc = accumulate_task<2>(colbase, ba, bb, sa, sb, N, gtab); break;
                    default: acc = accumulate_task<4>(colbase, ba, bb, sa, sb, N, gtab); break;
                }
                ps.S[t] = acc;
            }
            __syncwarp();
        }
    }
    __syncthreads();

    // ---- phase 5b: exact-zero fidelity ------------------------------------------------------------------
    // The entropy form above is exactly 0 when a channel is constant (as mi.cpp is), but for a table that is
    // exactly independent yet not constant it leaves rounding noise where mi.cpp may produce an exact 0 (every
    // p_xy/(p_x p_y) == 1.0, common for N = 8 with two bins). The P2 guards `den == 0 ? 0 : num/den` branch on
    // such zeros, so any pair whose MI is suspiciously small is recomputed cell by cell with the reference's own
    // expression and rounding (mi.cpp:66-87). Rare; the flagged lanes of a warp take turns using the warp's whole
    // histogram region as one contiguous u16 table.
    {
        const double dn = static_cast<double>(N);
        const double Stot = p.gpre[N];
        const double thr = 1e-4 / (dn * dn);
        if (warp < p.sm.mi_warps) {
            unsigned short *tab = reinterpret_cast<unsigned short *>(reinterpret_cast<uint32_t *>(big) +
                                                                      static_cast<size_t>(warp) * p.sm.wpl * 32);
            for (int grp = warp; grp < N_GROUP - 1; grp += p.sm.mi_warps) {
                const int t = grp * 32 + lane;
                int a, b;
                task_channels(t, a, b);
                // grouping keeps MI exactly 0 when either channel is constant (as mi.cpp yields exactly 0 there)
                double mi = ((ps.S[t] - ps.S[N_JOINT + a]) - (ps.S[N_JOINT + b] - Stot)) / dn;
                unsigned todo = __ballot_sync(0xffffffffu, mi != 0.0 && fabs(mi) < thr);
                while (todo) {
                    const int l = __ffs(todo) - 1;
                    todo &= todo - 1;
                    if (lane == l) {
                        const int ncell = nb * nb + nb;
                        for (int c = 0; c < ncell; ++c) tab[c] = 0;
                        const uint8_t *ba = bins + a, *bb = bins + b;
                        for (int j = 0; j < N; ++j) tab[ba[j * 20] * nb + bb[j * 20]] += 1;
                        for (int j = 0; j < nb; ++j) {  // column marginals
                            int sum = 0;
                            for (int i = 0; i < nb; ++i) sum += tab[i * nb + j];
                            tab[nb * nb + j] = static_cast<unsigned short>(sum);
                        }
                        mi = 0.0;
                        for (int i = 0; i < nb; ++i) {
                            int ni = 0;
                            for (int j = 0; j < nb; ++j) ni += tab[i * nb + j];
                            if (ni == 0) continue;
                            const double px = __ddiv_rn(static_cast<double>(ni), dn);
                            for (int j = 0; j < nb; ++j) {
                                const int n = tab[i * nb + j];
                                if (n == 0) continue;
                                const double pxy = __ddiv_rn(static_cast<double>(n), dn);
                                const double pp = __dmul_rn(px, __ddiv_rn(static_cast<double>(tab[nb * nb + j]), dn));
                                if (pp != 0.0) mi = __dadd_rn(mi, __dmul_rn(pxy, log(__ddiv_rn(pxy, pp))));
                            }
                        }
                    }
                    __syncwarp();
                }
                ps.mi[t] = mi;
            }
        }
    }
    __syncthreads();

    // ---- phase 6: dependencies -> alpha, beta, W_r_c (rpf.cpp:416-487) -----------------------------------
    // mi[] order: 4*i + {r0,r1,p0,p1} for feature i; 48 + 16*i + {r0,r1,p0,p1,f0..f11} for colour i.
    {
        const bool p2 = p.flags & FLAG_P2;
        const double eps = p.w_eps;
        auto ratio = [&](double num, double den) {
            if (eps != 0.0) den = den + eps;  // paper mode (rpf_set_paper_mode); 0 = the reference's ratios
            return (p2 && den == 0.0) ? 0.0 : num / den;
        };
        if (tid < NF) {
            const int i = tid;
            const double drf = (0.0 + ps.mi[4 * i]) + ps.mi[4 * i + 1];          // rpf.cpp:421
            const double dpf = (0.0 + ps.mi[4 * i + 2]) + ps.mi[4 * i + 3];      // rpf.cpp:425
            ps.wrf[i] = ratio(drf, drf + dpf);                                   // rpf.cpp:465
            ps.dcf[i] = ((0.0 + ps.mi[52 + i]) + ps.mi[68 + i]) + ps.mi[84 + i]; // P1: sum over colours of MI(c_i, f_j)
        } else if (tid >= 32 && tid < 32 + NC) {
            const int i = tid - 32, t = 48 + 16 * i;
            const double drc = (0.0 + ps.mi[t]) + ps.mi[t + 1];                  // rpf.cpp:432
            const double dpc = (0.0 + ps.mi[t + 2]) + ps.mi[t + 3];              // rpf.cpp:436
            double dfc = 0.0;
            for (int j = 0; j < NF; ++j) dfc += ps.mi[t + 4 + j];                // rpf.cpp:440
            ps.dsum[i] = dfc;
            ps.dsum[3 + i] = drc;
            ps.dsum[6 + i] = dpc;
            const double w = ratio(drc, drc + dpc);                              // rpf.cpp:470
            ps.wrck[i] = w;
            ps.alpha[i] = 1.0 - w;                                               // rpf.cpp:475
        }
        __syncthreads();
        if (tid < NF) {
            const int i = tid;
            const double Dfc = ((0.0 + ps.dsum[0]) + ps.dsum[1]) + ps.dsum[2];   // rpf.cpp:449-456
            const double Drc = ((0.0 + ps.dsum[3]) + ps.dsum[4]) + ps.dsum[5];
            const double Dpc = ((0.0 + ps.dsum[6]) + ps.dsum[7]) + ps.dsum[8];
            const double num = (p.flags & FLAG_P1) ? ps.dcf[i]
                                                   : (i < NC ? ps.dsum[i] : __longlong_as_double(0x7ff8000000000000ll));
            ps.beta[i] = (1.0 - ps.wrf[i]) * ratio(num, Dfc + Drc + Dpc);        // rpf.cpp:464,479
        } else if (tid == 32) {
            ps.wrc = (((0.0 + ps.wrck[0]) + ps.wrck[1]) + ps.wrck[2]) / NC;      // rpf.cpp:483-487
            atomicAdd(&p.counters[0], static_cast<unsigned long long>(N));
            atomicMin(&p.counters[2], static_cast<unsigned long long>(N));
            atomicMax(&p.counters[3], static_cast<unsigned long long>(N));
        }
        __syncthreads();
        if (p.probe.mi && static_cast<long long>(pix) == p.probe.pixel) {
            if (tid < N_JOINT) p.probe.mi[tid] = ps.mi[tid];
            if (tid == 0) { p.probe.mi[N_JOINT] = static_cast<double>(N); p.probe.mi[N_JOINT + 1] = static_cast<double>(nb); }
            if (tid < NCH) {  // marginal counts straight from the bin indices (the lane-private tables are gone by now)
                unsigned *m = p.probe.marg + tid * PROBE_BINS;
                for (int i = 0; i < PROBE_BINS; ++i) m[i] = 0u;
                for (int j = 0; j < N; ++j) m[bins[j * 20 + tid]] += 1u;
            }
        }
        if (p.dbg_n) {
            if (tid == 0) { p.dbg_n[pix] = N; p.dbg_wrc[pix] = ps.wrc; }
            if (tid < NC) p.dbg_alpha[pix * NC + tid] = ps.alpha[tid];
            if (tid < NF) p.dbg_beta[pix * NF + tid] = ps.beta[tid];
        }
    }

    // ---- phase 7: cross-bilateral weights and blend (rpf.cpp:627-717) -------------------------------------
    // w_ij = exp(-sum_p d^2/(2 sp2)) * exp(-sum_c alpha d^2/(2 s2)) * exp(-sum_f beta d^2/(2 s2)) on NORMALISED
    // values; here the three exponents are summed and each channel is pre-scaled by sqrt(coefficient)/sigma, so a
    // pair costs 17 subtract + 17 FMA + one exp. Differences to the three-exp product are O(1e-16) relative (and in
    // which denormal weights flush to 0). Each thread keeps its neighbour sample j in registers (read straight from
    // global memory); only the P own samples are shared, sample-major for 128-bit broadcast loads.
    {
        double *zi = reinterpret_cast<double *>(big);  // [P][18]
        if (tid < NWZ) {
            const int k = tid;
            const int c = k < 5 ? k : k + 2;
            const double sigma_p = static_cast<double>(B / 4);                       // rpf.cpp:531 integer division
            const double s2 = p.seed * p.seed / (1.0 - ps.wrc) / (1.0 - ps.wrc);     // rpf.cpp:662
            double coef;
            if (k < 2) coef = 1.0 / (2.0 * (sigma_p * sigma_p));
            else if (k < 5) coef = ps.alpha[k - 2] / (2.0 * s2);
            else coef = ps.beta[k - 5] / (2.0 * s2);
            const double sd = ps.SD[c];
            ps.scl[k] = sd == 0.0 ? 0.0 : sqrt(coef) / sd;
            ps.mz[k] = ps.M[c];
        }
        __syncthreads();
        auto load_z = [&](uint32_t g, double *z, double *col) {
#pragma unroll
            for (int k = 0; k < NWZ; ++k) {
                double x;
                if (k >= 2 && k < 5) {
                    x = p.col_in[static_cast<size_t>(k - 2) * S + g];
                    col[k - 2] = x;
                } else {
                    x = static_cast<double>(__ldg(p.f.planes + static_cast<size_t>(k < 2 ? k : k + 2) * S + g));
                }
                z[k] = scaled(x, ps.mz[k], ps.scl[k]);
            }
        };
        for (int n = tid; n < P; n += THREADS) {
            double z[NWZ], col[NC];
            load_z(idx[n], z, col);
#pragma unroll
            for (int k = 0; k < NWZ; ++k) zi[n * 18 + k] = z[k];
            zi[n * 18 + 17] = 0.0;
        }
        __syncthreads();

        constexpr int IC = 4;  // own samples per pass
        for (int i0 = 0; i0 < P; i0 += IC) {
            double sw[IC], swc[IC][NC];
#pragma unroll
            for (int ii = 0; ii < IC; ++ii) { sw[ii] = 0.0; swc[ii][0] = swc[ii][1] = swc[ii][2] = 0.0; }
            for (int j = tid; j < N; j += THREADS) {
                double zj[NWZ + 1], cj[NC];
                load_z(idx[j], zj, cj);
                zj[NWZ] = 0.0;
#pragma unroll
                for (int ii = 0; ii < IC; ++ii) {
                    if (i0 + ii < P) {
                        const double2 *zr = reinterpret_cast<const double2 *>(zi + (i0 + ii) * 18);
                        double arg = 0.0;
#pragma unroll
                        for (int k2 = 0; k2 < 9; ++k2) {
                            const double2 v = zr[k2];
                            const double d0 = v.x - zj[2 * k2], d1 = v.y - zj[2 * k2 + 1];
                            arg = fma(d0, d0, arg);
                            arg = fma(d1, d1, arg);
                        }
                        // exp(-arg) underflows to exactly 0 beyond ~745.2; NaN falls through to exp
                        const double w = arg > 745.2 ? 0.0 : exp(-arg);
                        sw[ii] += w;
                        swc[ii][0] = fma(w, cj[0], swc[ii][0]);
                        swc[ii][1] = fma(w, cj[1], swc[ii][1]);
                        swc[ii][2] = fma(w, cj[2], swc[ii][2]);
                    }
                }
            }
            // reduce the 16 partial sums over the CTA
            __syncthreads();  // previous pass finished with ps.red
#pragma unroll
            for (int ii = 0; ii < IC; ++ii) {
                const double a = warp_sum(sw[ii]), b0 = warp_sum(swc[ii][0]), b1 = warp_sum(swc[ii][1]),
                             b2 = warp_sum(swc[ii][2]);
                if (lane == 0) {
                    ps.red[warp][ii * 4 + 0] = a;
                    ps.red[warp][ii * 4 + 1] = b0;
                    ps.red[warp][ii * 4 + 2] = b1;
                    ps.red[warp][ii * 4 + 3] = b2;
                }
            }
            __syncthreads();
            if (tid < IC * NC) {
                const int ii = tid / NC, k = tid - ii * NC;
                if (i0 + ii < P) {
                    const double a = ((ps.red[0][ii * 4] + ps.red[1][ii * 4]) + ps.red[2][ii * 4]) + ps.red[3][ii * 4];
                    const double b = ((ps.red[0][ii * 4 + 1 + k] + ps.red[1][ii * 4 + 1 + k]) + ps.red[2][ii * 4 + 1 + k]) +
                                     ps.red[3][ii * 4 + 1 + k];
                    const double c = b / a;  // rpf.cpp:700
                    if (c != c) atomicAdd(&p.counters[1], 1ull);  // P4: count, never exit (rpf.cpp:702-705)
                    p.col_out[static_cast<size_t>(k) * S + base + i0 + ii] = c;
                }
            }
        }
    }
}

}  // namespace RPF_NS
