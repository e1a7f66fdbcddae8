// rpf_kernels.cuh -- hand-written sm_100a kernels for the RPF sample filter.
//
// One CTA filters one pixel (everything in rpf.cpp:556-723 is per pixel and independent). All per-pixel
// state lives in shared memory; HBM is touched once for the gather (L1/L2 serve the box overlap between
// neighbouring CTAs) and once for the 3 output colours.
//
// Exactness: everything that feeds a discontinuity of the reference -- per-pixel mean/sigma and the 3-sigma
// test (rpf.cpp:575-583), neighbourhood mean/sigma (rpf.cpp:596-601, sequential sums like ops.h:111-144),
// normalisation (sd.h:224-235) and histogram binning (mi.cpp:14-16,27-38) -- is computed in FP64 with
// explicit round-to-nearest intrinsics so nvcc cannot contract a*b+c into FMA (gcc's x86-64 build of the
// reference has no FMA). Joint histograms are exact integer counts. Only log/exp and summation order in
// MI, weights and blend differ from the CPU at the 1e-15 level.
// RE-INCLUDABLE (once per channel layout, rpf_layouts.cuh); the layout-independent kernels (colour conversion, spike removal,
// Film reconstruction) are compiled for the reference layout only (RPF_FULL).

namespace RPF_NS {

#if RPF_FULL
// ---------------------------------------------------------------------------------------------------
// K0: float -> double colour planes (upload)
__global__ void colors_to_f64_kernel(const float *__restrict__ planes, size_t S, double *__restrict__ col) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= 3 * S) return;
    col[i] = static_cast<double>(planes[static_cast<size_t>(CH_COL) * S + i]);
}
#endif

// ---------------------------------------------------------------------------------------------------
// K1: per-pixel feature mean and 3*sigma (FillMeanAndStddev, rpf.cpp:302-353; getMean/getStdDev ops.h:111-144).
// One thread per (pixel, feature); sums run sequentially in sample order like the reference.
// Paper mode (rpf_set_paper_mode; not in the reference, both 0 by default): world-position features take pos_sigmas
// instead of 3, and the threshold never drops below min_sigma (low-variance guard). A NaN sigma stays NaN (accepts).
// The 3-sigma test of rpf.cpp:579-586 on a float feature f is  !(|fl(f - mu)| >= sd3)  in double arithmetic. fl(f - mu) is monotone
// in f, so for one pixel and feature the floats that pass form an INTERVAL [lo, hi] of floats: found here once per pixel by
// bisection over the ordered floats with the reference's own expression as the predicate, it turns the 24,200 tests per pixel of
// membership_kernel into two float compares each, exactly (pbound[pix] = lo[NF], hi[NF]). Written so that a candidate is rejected
// iff (f < lo || f > hi): a NaN feature passes (NaN >= x is false in the reference too); a NaN sigma gives [-inf, +inf] (everything
// passes, infinities included); otherwise infinities fail; sigma 0 gives the empty interval [+inf, -inf].
__device__ __forceinline__ float ordered_float(long long k) {  // k in [-0x7f800000, 0x7f7fffff] <-> finite floats, ascending
    const int i = static_cast<int>(k);
    return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff);
}
__device__ inline void three_sigma_interval(double mu, double sd3, float &lo, float &hi) {
    const float inf = __int_as_float(0x7f800000);
    if (sd3 != sd3) { lo = -inf; hi = inf; return; }
    const long long KMIN = -0x7f800000ll, KMAX = 0x7f7fffffll;
    // hi: the largest float with fl(f - mu) < sd3 (true below, false above)
    long long a = KMIN - 1, b = KMAX + 1;
    while (b - a > 1) {
        const long long m = a + (b - a) / 2;
        if (__dsub_rn(static_cast<double>(ordered_float(m)), mu) < sd3) a = m; else b = m;
    }
    hi = a >= KMIN ? ordered_float(a) : -inf;
    // lo: the smallest float with fl(f - mu) > -sd3 (false below, true above)
    a = KMIN - 1; b = KMAX + 1;
    while (b - a > 1) {
        const long long m = a + (b - a) / 2;
        if (__dsub_rn(static_cast<double>(ordered_float(m)), mu) > -sd3) b = m; else a = m;
    }
    lo = b <= KMAX ? ordered_float(b) : inf;
    if (!(lo <= hi)) { lo = inf; hi = -inf; }
}

__global__ void pixel_stats_kernel(FrameView f, int y0, int y1, double *__restrict__ pstat, double min_sigma, double pos_sigmas,
                                   float *__restrict__ pbound = nullptr) {
    const long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const int k = static_cast<int>(t % NF);
    const long long pix = t / NF + static_cast<long long>(y0) * f.W;
    if (pix >= static_cast<long long>(y1) * f.W) return;
    const float *p = f.planes + static_cast<size_t>(CH_FEAT + k) * f.S + static_cast<size_t>(pix) * f.P;
    double sum = 0.0, sq = 0.0;
    for (int s = 0; s < f.P; ++s) {
        const double v = static_cast<double>(p[s]);
        sum = __dadd_rn(sum, v);
        sq = __dadd_rn(sq, __dmul_rn(v, v));
    }
    const double n = static_cast<double>(f.P);
    const double mu = __ddiv_rn(sum, n);
    const double sd = __dsqrt_rn(__dsub_rn(__ddiv_rn(sq, n), __dmul_rn(mu, mu)));  // no clamp: NaN is meaningful
    pstat[pix * (2 * NF) + k] = mu;
    double sd3 = __dmul_rn(sd, 3.0);  // rpf.cpp:579 multiplyArray(stdDev, 3)
    const bool world_pos = feature_is_world_pos(k);  // p0, p1 of sd.h:62-94
    if (world_pos && pos_sigmas > 0.0) sd3 = __dmul_rn(sd, pos_sigmas);
    if (min_sigma > 0.0 && sd3 < min_sigma) sd3 = min_sigma;
    pstat[pix * (2 * NF) + NF + k] = sd3;
    if (pbound) {
        float lo, hi;
        three_sigma_interval(mu, sd3, lo, hi);
        pbound[pix * (2 * NF) + k] = lo;
        pbound[pix * (2 * NF) + NF + k] = hi;
    }
}

#if RPF_FULL
// Paper-mode spike removal (rpf_hdr_clamp; not in the reference): one thread per pixel of rows [y0,y1). Per channel the
// sequential mean and population sigma of the pixel's samples; a sample with |c - mean| > k * sigma in any channel
// takes the mean in all three channels. In place on colours [3][S].
__global__ void hdr_clamp_kernel(double *__restrict__ col, size_t S, int W, int P, int y0, int y1, double k_sigmas) {
    const long long pix = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x + static_cast<long long>(y0) * W;
    if (pix >= static_cast<long long>(y1) * W) return;
    const size_t base = static_cast<size_t>(pix) * P;
    double mu[NC], lim[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        const double *c = col + k * S + base;
        double sum = 0.0, sq = 0.0;
        for (int i = 0; i < P; ++i) {
            sum = __dadd_rn(sum, c[i]);
            sq = __dadd_rn(sq, __dmul_rn(c[i], c[i]));
        }
        const double n = static_cast<double>(P);
        mu[k] = __ddiv_rn(sum, n);
        const double var = __dsub_rn(__ddiv_rn(sq, n), __dmul_rn(mu[k], mu[k]));
        lim[k] = __dmul_rn(k_sigmas, __dsqrt_rn(var > 0.0 ? var : 0.0));
    }
    for (int i = 0; i < P; ++i) {
        bool spike = false;
#pragma unroll
        for (int k = 0; k < NC; ++k)
            if (fabs(__dsub_rn(col[k * S + base + i], mu[k])) > lim[k]) spike = true;
        if (spike) {
#pragma unroll
            for (int k = 0; k < NC; ++k) col[k * S + base + i] = mu[k];
        }
    }
}

#endif  // RPF_FULL

// ---------------------------------------------------------------------------------------------------
// Gather-friendly copies for the general path. The planes are structure-of-arrays (one coalesced stream per channel
// when neighbouring pixels are read together); the general path instead gathers scattered accepted samples, where a
// sample's channels should share sectors: rec16[g] = the 16 non-colour channels of sample g in slot order (64 B, built
// once per upload), col4[g] = {r, g, b, 0} of the current colours as doubles (32 B, rebuilt per pass).
__global__ void pack_features_kernel(const float *__restrict__ planes, size_t S, float4 *__restrict__ rec16) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= NQ4 * S) return;
    const size_t g = i / NQ4;
    const int q = static_cast<int>(i - g * NQ4);
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = 4 * q + k < NSLOT ? planes[static_cast<size_t>(slot_channel(4 * q + k)) * S + g] : 0.f;
    rec16[i] = make_float4(v[0], v[1], v[2], v[3]);
}
// rec19[g] = all 19 channels of sample g as doubles in channel order (+1 pad, 160 B): ordered_stats_kernel gives lane c
// channel c, and with this copy a gathered sample is ONE 8-byte load per lane on one code path (no float/double split).
__global__ void pack_colors_kernel(const double *__restrict__ col, size_t S, double4 *__restrict__ col4, double *__restrict__ rec19) {
    const size_t g = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (g >= S) return;
    const double r = col[g], gg = col[S + g], b = col[2 * S + g];
    col4[g] = make_double4(r, gg, b, 0.0);
    rec19[g * REC19 + CH_COL] = r;
    rec19[g * REC19 + CH_COL + 1] = gg;
    rec19[g * REC19 + CH_COL + 2] = b;
}
__global__ void pack_rec19_kernel(const float *__restrict__ planes, size_t S, double *__restrict__ rec19) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= S * REC19) return;
    const size_t g = i / REC19;
    const int c = static_cast<int>(i - g * REC19);
    if (c >= NCH) { rec19[i] = 0.0; return; }
    if (c >= CH_COL && c < CH_COL + NC) return;  // colours: pack_colors_kernel, every pass
    rec19[i] = static_cast<double>(planes[static_cast<size_t>(c) * S + g]);
}

// ---------------------------------------------------------------------------------------------------
// helpers

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double u = __shfl_xor_sync(0xffffffffu, v, o);
        if (u < v) v = u;
    }
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double u = __shfl_xor_sync(0xffffffffu, v, o);
        if (v < u) v = u;
    }
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// sd.h:224-235 with ops.h:44-51: (x - mean) / sigma, sigma == 0 -> 0
__device__ __forceinline__ double normalise(double x, double m, double sd) {
    return sd == 0.0 ? 0.0 : __ddiv_rn(__dsub_rn(x, m), sd);
}

// Weight-space coordinate of a raw value: (x - mean) * sqrt(coefficient) / sigma. A channel that is constant over the
// neighbourhood has scale 0 and coordinate 0 whatever x is (ops.h:44-51 returns 0 for sigma == 0 without dividing) -- also
// when x is Inf or NaN, which under P3 always comes with sigma 0 (the variance of a non-finite channel is NaN).
__device__ __forceinline__ double scaled(double x, double m, double scale) {
    return scale == 0.0 ? 0.0 : (x - m) * scale;
}

// Small per-pixel scalars in shared memory.
struct PixelSmall {
    double mu[NF], sd3[NF];          // centre pixel feature statistics
    double M[NCH], SD[NCH];          // neighbourhood mean / sigma
    double lo[NCH], hi[NCH];         // min / max of the normalised channel
    double rawlo[NCH], rawhi[NCH];   // min / max of the raw channel
    double sumx[NCH], sumsq[NCH];
    double S[N_TASK];                // sum over cells of n ln n, per task
    double mi[N_JOINT];              // mutual information per pair
    double alpha[NC], beta[NF], wrc;
    double wrf[NF], dcf[NF], dsum[9], wrck[NC];   // phase-6 intermediates
    double scl[NWZ], mz[NWZ];        // phase-7 per-channel scale and mean
    double red[4][16];               // cross-warp reduction scratch for the blend
    int wcnt[2][4];
    int N, nb;
};

}  // namespace RPF_NS

#if RPF_FULL
namespace RPF_NS {

// double -> float colours for the host (the Film takes Float RGB, sd.h:104-112)
__global__ void colors_to_f32_kernel(const double *__restrict__ col, size_t n, float *__restrict__ out) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i < n) out[i] = static_cast<float>(col[i]);
}

// Film reconstruction for the default box filter of radius 0.5 (filters/box.cpp:43-47): every sample lands in its own
// pixel with filter weight 1, so FilmTile::AddSample + Film::WriteImage (film.h:121-161, film.cpp:169-211) reduce
// to sum(L*rayWeight)/sum(1) per pixel, accumulated in Float in sample order. rayWeight is 1 for every camera
// in the reference's scenes; the RGB->XYZ->RGB round trip of the Film is not reproduced here.
__global__ void reconstruct_box_kernel(const double *__restrict__ col, size_t S, int npix, int P, float *__restrict__ rgb) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= npix * 3) return;
    const int pix = t / 3, k = t - pix * 3;
    const double *c = col + static_cast<size_t>(k) * S + static_cast<size_t>(pix) * P;
    float acc = 0.f;
    for (int s = 0; s < P; ++s) acc += static_cast<float>(c[s]);
    rgb[t] = acc / static_cast<float>(P);
}

// ---------------------------------------------------------------------------------------------------
// Film reconstruction, any pbrt filter: FilmTile::AddSample (film.h:121-161) + Film::MergeFilmTile
// (film.cpp:117-130) + Film::WriteImage (film.cpp:169-205) as a gather. One thread per output pixel walks the sample
// pixels that can reach it in the order the reference adds them (rpf.cpp:783-786: x outer, y inner, sample order) and
// accumulates in Float with the reference's operation order and no FMA contraction, so the image is bit-identical to
// the reference's single FilmTile. A sample is assumed to lie within one pixel of the pixel that stores it (pFilm =
// pixel + [0,1) offset in every pbrt sampler).
struct FilmParams {
    const float *planes;   // [19][S]; planes 0,1 = pFilm
    const double *col;     // [3][S] current colours
    const float *ray_weight;  // [S] or null (= 1)
    size_t S;
    int W, H, P;           // sample extent held by the context
    int sx0, sy0;          // raster coordinates of the context's sample pixel (0,0) (sampleBounds.pMin)
    int px0, py0, px1, py1;  // croppedPixelBounds
    float rx, ry, scale, max_lum;
    const float *table;    // [16*16] Film::filterTable
    float *rgb;            // [(py1-py0)*(px1-px0)][3]
};

__global__ void film_reconstruct_kernel(const FilmParams p) {
    const int pw = p.px1 - p.px0, ph = p.py1 - p.py0;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= pw * ph) return;
    const int x = p.px0 + t % pw, y = p.py0 + t / pw;
    const float irx = __fdiv_rn(1.f, p.rx), iry = __fdiv_rn(1.f, p.ry);  // filter.h:52
    float sum[3] = {0.f, 0.f, 0.f}, fws = 0.f;
    const int xs_lo = max(static_cast<int>(floorf(x + 0.5f - p.rx)) - 2 - p.sx0, 0);
    const int xs_hi = min(static_cast<int>(ceilf(x + 0.5f + p.rx)) + 1 - p.sx0, p.W - 1);
    const int ys_lo = max(static_cast<int>(floorf(y + 0.5f - p.ry)) - 2 - p.sy0, 0);
    const int ys_hi = min(static_cast<int>(ceilf(y + 0.5f + p.ry)) + 1 - p.sy0, p.H - 1);
    for (int xs = xs_lo; xs <= xs_hi; ++xs)
        for (int ys = ys_lo; ys <= ys_hi; ++ys) {
            const size_t g0 = (static_cast<size_t>(ys) * p.W + xs) * p.P;
            for (int s = 0; s < p.P; ++s) {
                const size_t g = g0 + s;
                const float dx = __fsub_rn(p.planes[g], 0.5f), dy = __fsub_rn(p.planes[p.S + g], 0.5f);  // pFilmDiscrete
                const int p0x = max(static_cast<int>(ceilf(__fsub_rn(dx, p.rx))), p.px0);
                const int p1x = min(static_cast<int>(floorf(__fadd_rn(dx, p.rx))) + 1, p.px1);
                const int p0y = max(static_cast<int>(ceilf(__fsub_rn(dy, p.ry))), p.py0);
                const int p1y = min(static_cast<int>(floorf(__fadd_rn(dy, p.ry))) + 1, p.py1);
                if (x < p0x || x >= p1x || y < p0y || y >= p1y) continue;
                float L[3] = {static_cast<float>(p.col[g]), static_cast<float>(p.col[p.S + g]), static_cast<float>(p.col[2 * p.S + g])};
                const float lum = __fadd_rn(__fadd_rn(__fmul_rn(0.212671f, L[0]), __fmul_rn(0.715160f, L[1])), __fmul_rn(0.072169f, L[2]));
                if (lum > p.max_lum) {  // film.h:125-126
                    const float a = __fdiv_rn(p.max_lum, lum);
#pragma unroll
                    for (int k = 0; k < 3; ++k) L[k] = __fmul_rn(L[k], a);
                }
                const float fx = fabsf(__fmul_rn(__fmul_rn(__fsub_rn(static_cast<float>(x), dx), irx), 16.f));
                const float fy = fabsf(__fmul_rn(__fmul_rn(__fsub_rn(static_cast<float>(y), dy), iry), 16.f));
                const int ifx = min(static_cast<int>(floorf(fx)), 15), ify = min(static_cast<int>(floorf(fy)), 15);
                const float fw = p.table[ify * 16 + ifx];
                const float sw = p.ray_weight ? p.ray_weight[g] : 1.f;
#pragma unroll
                for (int k = 0; k < 3; ++k) sum[k] = __fadd_rn(sum[k], __fmul_rn(__fmul_rn(L[k], sw), fw));  // film.h:157
                fws = __fadd_rn(fws, fw);
            }
        }
    // MergeFilmTile: contribSum -> XYZ (spectrum.h:62-66), added to a zero pixel; WriteImage: XYZ -> RGB (spectrum.h:56-60)
    auto dot3 = [](float a, float b, float c, float u, float v, float w) {
        return __fadd_rn(__fadd_rn(__fmul_rn(a, u), __fmul_rn(b, v)), __fmul_rn(c, w));
    };
    float xyz[3], rgb[3];
    xyz[0] = __fadd_rn(0.f, dot3(0.412453f, 0.357580f, 0.180423f, sum[0], sum[1], sum[2]));
    xyz[1] = __fadd_rn(0.f, dot3(0.212671f, 0.715160f, 0.072169f, sum[0], sum[1], sum[2]));
    xyz[2] = __fadd_rn(0.f, dot3(0.019334f, 0.119193f, 0.950227f, sum[0], sum[1], sum[2]));
    rgb[0] = __fsub_rn(__fsub_rn(__fmul_rn(3.240479f, xyz[0]), __fmul_rn(1.537150f, xyz[1])), __fmul_rn(0.498535f, xyz[2]));
    rgb[1] = __fadd_rn(__fadd_rn(__fmul_rn(-0.969256f, xyz[0]), __fmul_rn(1.875991f, xyz[1])), __fmul_rn(0.041556f, xyz[2]));
    rgb[2] = __fadd_rn(__fsub_rn(__fmul_rn(0.055648f, xyz[0]), __fmul_rn(0.204043f, xyz[1])), __fmul_rn(1.057311f, xyz[2]));
    if (fws != 0.f) {
        const float inv = __fdiv_rn(1.f, fws);
#pragma unroll
        for (int k = 0; k < 3; ++k) rgb[k] = fmaxf(0.f, __fmul_rn(rgb[k], inv));
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) p.rgb[static_cast<size_t>(t) * 3 + k] = __fmul_rn(__fadd_rn(rgb[k], 0.f), p.scale);  // splats are zero
}

}  // namespace RPF_NS
#endif  // RPF_FULL
