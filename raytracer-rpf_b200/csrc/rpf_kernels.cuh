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
#pragma once
#include "rpf_common.cuh"

namespace rpf {

// ---------------------------------------------------------------------------------------------------
// K0: float -> double colour planes (upload)
__global__ void colors_to_f64_kernel(const float *__restrict__ planes, size_t S, double *__restrict__ col) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= 3 * S) return;
    col[i] = static_cast<double>(planes[static_cast<size_t>(CH_COL) * S + i]);
}

// ---------------------------------------------------------------------------------------------------
// K1: per-pixel feature mean and 3*sigma (FillMeanAndStddev, rpf.cpp:302-353; getMean/getStdDev ops.h:111-144).
// One thread per (pixel, feature); sums run sequentially in sample order like the reference.
__global__ void pixel_stats_kernel(FrameView f, int y0, int y1, double *__restrict__ pstat) {
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
    pstat[pix * 24 + k] = mu;
    pstat[pix * 24 + NF + k] = __dmul_rn(sd, 3.0);  // rpf.cpp:579 multiplyArray(stdDev, 3)
}

// ---------------------------------------------------------------------------------------------------
// Gather-friendly copies for the general path. The planes are structure-of-arrays (one coalesced stream per channel
// when neighbouring pixels are read together); the general path instead gathers scattered accepted samples, where a
// sample's channels should share sectors: rec16[g] = the 16 non-colour channels of sample g in slot order (64 B, built
// once per upload), col4[g] = {r, g, b, 0} of the current colours as doubles (32 B, rebuilt per pass).
__global__ void pack_features_kernel(const float *__restrict__ planes, size_t S, float4 *__restrict__ rec16) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= 4 * S) return;
    const size_t g = i >> 2;
    const int q = static_cast<int>(i & 3);
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = planes[static_cast<size_t>(slot_channel(4 * q + k)) * S + g];
    rec16[i] = make_float4(v[0], v[1], v[2], v[3]);
}
__global__ void pack_colors_kernel(const double *__restrict__ col, size_t S, double4 *__restrict__ col4) {
    const size_t g = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (g >= S) return;
    col4[g] = make_double4(col[g], col[S + g], col[2 * S + g], 0.0);
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

}  // namespace rpf

namespace rpf {

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

}  // namespace rpf
