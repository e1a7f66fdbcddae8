// rpf_blend_wide.cuh -- cross-bilateral weights and blend (rpf.cpp:627-717) for many samples per pixel (spp >= 16; 27 % faster than blend_kernel at 16 spp, 2.7x at 64).
//
// blend_kernel (rpf_general.cuh) gives every thread one neighbour and loops over the own samples; its accumulators
// (four doubles per own sample) live in registers, so it takes the own samples eight at a time and re-gathers the whole
// neighbourhood for every eight: fine at 8 spp, eight gathers at 64 spp. Here the roles are swapped. The accepted
// samples stream through shared memory in chunks of 256: (1) every thread gathers ONE neighbour record (all the loads
// of a chunk are in flight together) and leaves its scaled channels in shared memory; (2) thread (i, q) owns own sample
// i (up to 64 at a time) and every TQ-th neighbour of the chunk, reads it from shared memory (a broadcast among the
// threads of one neighbour) and keeps four accumulators. 64 registers, four CTAs per SM. The sums over q are folded in
// a fixed order, so the result is reproducible. Same weight arithmetic as blend_kernel.
// Reference layout only (RPF_FULL in rpf_layouts.cuh): wider channel layouts take the general path.

namespace RPF_NS {

constexpr int KW_THREADS = 256, KW_SEG_WORDS = 128;
constexpr int KW_ZS = 21;  // doubles per staged neighbour: 17 scaled channels + 3 raw colours + 1 pad (odd: bank spread)
struct KWSmall {
    double scl[NWZ], mz[NWZ];
    int wsum[KW_THREADS / 32 + 1];
};

__global__ void __launch_bounds__(KW_THREADS, 4) blend_wide_kernel(const GeneralParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ KWSmall ks;
    __shared__ unsigned short segidx[KW_SEG_WORDS * 32];
    const int tid = threadIdx.x;
    const int W = p.f.W, P = p.f.P, B = p.box;
    const size_t S = p.f.S;
    const int pl = strip_order(blockIdx.x, p.f.W, p.nrows);
    const int py = chunk_row(pl / W, p.row0, p.gap_at, p.gap), px = pl - (pl / W) * W;
    const size_t pix = static_cast<size_t>(py) * W + px;
    const uint32_t base = static_cast<uint32_t>(pix * P);
    const uint32_t *mask = p.mask + static_cast<size_t>(pl) * p.pl.mw;
    const double *rec = p.rec + static_cast<size_t>(pl) * REC_DOUBLES;
    double *chunk = reinterpret_cast<double *>(smem);  // [256][KW_ZS]; also the reduction scratch [TQ][TI][4] at the end of a pass
    double *zi = chunk + KW_THREADS * KW_ZS;           // [P][18]: order colour(3), features(12), position(2), pad

    // channel order inside z: k = 0..2 colour, 3..14 features, 15..16 position (most selective first)
    if (tid < NWZ) {
        const int k = tid;
        const int c = k < 3 ? CH_COL + k : (k < 15 ? CH_FEAT + (k - 3) : CH_POS + (k - 15));
        const double wrc = rec[REC_WRC];
        const double sigma_p = static_cast<double>(B / 4);                       // rpf.cpp:531 integer division
        const double s2 = p.seed * p.seed / (1.0 - wrc) / (1.0 - wrc);           // rpf.cpp:662
        double coef;
        if (k < 3) coef = rec[REC_ALPHA + k] / (2.0 * s2);
        else if (k < 15) coef = rec[REC_BETA + (k - 3)] / (2.0 * s2);
        else coef = 1.0 / (2.0 * (sigma_p * sigma_p));
        const double sd = rec[REC_SD + c];
        ks.scl[k] = sd == 0.0 ? 0.0 : sqrt(coef) / sd;
        ks.mz[k] = rec[REC_M + c];
    }
    __syncthreads();
    // scaled channels of sample g -> dst[0..16] (z order: colour, features = slots 4..15, position = slots 0, 1)
    auto stage = [&](uint32_t g, double *dst, bool with_colours) {
        const float4 *r16 = p.rec16 + static_cast<size_t>(g) * 4;
        const float4 f0 = __ldg(r16), f1 = __ldg(r16 + 1), f2 = __ldg(r16 + 2), f3 = __ldg(r16 + 3);
        const double4 cc = p.col4[g];
        dst[0] = scaled(cc.x, ks.mz[0], ks.scl[0]);
        dst[1] = scaled(cc.y, ks.mz[1], ks.scl[1]);
        dst[2] = scaled(cc.z, ks.mz[2], ks.scl[2]);
        const float fs[16] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w, f2.x, f2.y, f2.z, f2.w, f3.x, f3.y, f3.z, f3.w};
#pragma unroll
        for (int k = 3; k < NWZ; ++k) {
            const double x = static_cast<double>(k < 15 ? fs[k + 1] : fs[k - 15]);
            dst[k] = scaled(x, ks.mz[k], ks.scl[k]);
        }
        if (with_colours) { dst[17] = cc.x; dst[18] = cc.y; dst[19] = cc.z; }
    };
    for (int n = tid; n < P; n += KW_THREADS) {
        stage(base + n, zi + n * 18, false);
        zi[n * 18 + 17] = 0.0;
    }

    // thread (i, q): TI own samples at a time, TQ = 256 / TI neighbour slices
    int TI = 1;
    while (TI < P && TI < 64) TI <<= 1;
    const int TQ = KW_THREADS / TI;
    const int ti = tid & (TI - 1), tq = tid / TI;
    const int nseg = (p.pl.mw + KW_SEG_WORDS - 1) / KW_SEG_WORDS;
    for (int i0 = 0; i0 < P; i0 += TI) {
        __syncthreads();  // zi ready; chunk free
        const int i = i0 + ti;
        const bool live = i < P;
        const double *zr = zi + (live ? i : 0) * 18;
        const double zc0 = zr[0], zc1 = zr[1], zc2 = zr[2];
        double sw = 0.0, sc0 = 0.0, sc1 = 0.0, sc2 = 0.0;
        for (int seg = 0; seg < nseg; ++seg) {
            const int w = seg * KW_SEG_WORDS + tid;
            unsigned bits = (tid < KW_SEG_WORDS && w < p.pl.mw) ? mask[w] : 0u;
            int total;
            int off = block_exclusive_scan<KW_THREADS>(__popc(bits), ks.wsum, &total);
            while (bits) {
                const int b = __ffs(bits) - 1;
                bits &= bits - 1;
                segidx[off++] = static_cast<unsigned short>(tid * 32 + b);
            }
            __syncthreads();
            for (int c0 = 0; c0 < total; c0 += KW_THREADS) {
                const int nc = min(KW_THREADS, total - c0);
                if (tid < nc) {
                    const int t = seg * (KW_SEG_WORDS * 32) + segidx[c0 + tid];
                    const uint32_t g = candidate_index(p.cd, t, px, py, B, P, W);
                    stage(g, chunk + tid * KW_ZS, true);
                }
                __syncthreads();
                if (live) {
                    for (int j = tq; j < nc; j += TQ) {
                        const double *zj = chunk + j * KW_ZS;
                        double d = zc0 - zj[0];
                        double arg = d * d;
                        d = zc1 - zj[1];
                        arg = fma(d, d, arg);
                        d = zc2 - zj[2];
                        arg = fma(d, d, arg);
                        // NaN falls through to exp; so does a neighbour with an Inf / NaN raw colour, which the reference
                        // blends in even with weight 0 (rpf.cpp:692: 0 * inf = NaN)
                        const double cs = (zj[17] + zj[18]) + zj[19];
                        if (!(arg > 745.2) || !(cs - cs == 0.0)) {
#pragma unroll
                            for (int k = 3; k < NWZ; ++k) {
                                const double e = zr[k] - zj[k];
                                arg = fma(e, e, arg);
                            }
                            const double wgt = arg > 745.2 ? 0.0 : exp(-arg);
                            sw += wgt;
                            sc0 = fma(wgt, zj[17], sc0);
                            sc1 = fma(wgt, zj[18], sc1);
                            sc2 = fma(wgt, zj[19], sc2);
                        }
                    }
                }
                __syncthreads();
            }
        }
        // fold the TQ partial sums of every own sample in a fixed order
        double *red = chunk;  // [TQ][TI][4]
        red[(tq * TI + ti) * 4 + 0] = sw;
        red[(tq * TI + ti) * 4 + 1] = sc0;
        red[(tq * TI + ti) * 4 + 2] = sc1;
        red[(tq * TI + ti) * 4 + 3] = sc2;
        __syncthreads();
        if (tid < TI * NC) {
            const int ii = tid / NC, k = tid - ii * NC;
            if (i0 + ii < P) {
                double a = 0.0, bsum = 0.0;
                for (int q = 0; q < TQ; ++q) {
                    a += red[(q * TI + ii) * 4];
                    bsum += red[(q * TI + ii) * 4 + 1 + k];
                }
                const double c = bsum / a;  // rpf.cpp:700
                if (c != c) atomicAdd(&p.counters[1], 1ull);  // P4: count, never exit (rpf.cpp:702-705)
                p.col_out[static_cast<size_t>(k) * S + base + i0 + ii] = c;
            }
        }
    }
}

}  // namespace RPF_NS
