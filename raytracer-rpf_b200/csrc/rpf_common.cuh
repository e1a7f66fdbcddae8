// rpf_common.cuh -- shared definitions for the sm_100a RPF kernels.
// Channel order and slices follow SampleData::data (reference src/custom/sd.h:62-94,149-212).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rpf_subsample.h"

namespace rpf {

constexpr int NCH = 19;
constexpr int NF = 12, NC = 3, NR = 2, NP = 2;
constexpr int CH_POS = 0, CH_COL = 2, CH_RND = 5, CH_FEAT = 7;
constexpr int N_JOINT = 96;             // MI pairs of rpf.cpp:416-442
constexpr int N_TASK = N_JOINT + NCH;   // + one marginal histogram per channel
constexpr int N_GROUP = 4;              // task groups of 32 lanes: 3 joint groups + the marginals
constexpr int THREADS = 128;
constexpr int REC19 = 20;               // doubles per sample of the all-double gather record (19 channels + pad)
constexpr int NWZ = 17;                 // channels entering the weights: pos(2) + colour(3) + features(12)

constexpr int FLAG_P1 = 1, FLAG_P2 = 2, FLAG_P3 = 4;

// Optional neighbour subsampling (include/rpf_subsample.h): keep_q24 >= RPF_SUBSAMPLE_ALL switches it off.
struct SubsampleParams {
    unsigned long long seed;
    unsigned keep_q24;
    int row_g0;   // global row of this context's row 0 (row bands)
};

// Component-level probe of ONE pixel (rpf_set_debug_pixel): its 96 mutual informations in task order (task_channels below),
// N, the bin count, and the 19 marginal histograms (mi.cpp:5-21) -- what SURVEY.md 8(c) pins as "histogram counts exact".
constexpr int PROBE_BINS = 1024;  // sqrt(2^20 candidates)
struct DebugProbe {
    long long pixel;      // y*W+x in context coordinates, -1: off
    double *mi;           // [96] MI values, [96] = N, [97] = bins
    unsigned *marg;       // [19][PROBE_BINS]
};

// Device view of the frame: 19 float planes, each S = W*H*P samples, index (y*W+x)*P+s.
struct FrameView {
    const float *planes;
    size_t S;
    int W, H, P;
};

// Byte offsets into dynamic shared memory for the per-pixel kernel, computed on the host.
struct SmemLayout {
    int ncap;      // box*box*spp, capacity of the neighbourhood
    int nstride;   // odd-padded ncap: stride of channel-major staging rows (bank-conflict free)
    int wpl;       // 32-bit words per lane of one joint histogram column (two u16 cells per word)
    int mi_warps;  // warps that own a histogram region (1..4)
    int off_idx, off_bins, off_gtab, off_small, off_big;
    int total;
};

struct FilterParams {
    FrameView f;
    const double *col_in;   // [3][S] current colours
    double *col_out;        // [3][S] next colours
    const double *pstat;    // [W*H][24]: per-pixel feature mean[12], 3*sigma[12]
    const double *gtab;     // g[k] = (k+1)ln(k+1) - k ln k
    const double *gpre;     // gpre[n] = fl(g[0]+...+g[n-1]) summed sequentially
    int box, y0, y1, flags;
    double seed;
    SmemLayout sm;
    int *dbg_n;             // optional per-pixel outputs (may be null)
    double *dbg_alpha, *dbg_beta, *dbg_wrc;
    unsigned long long *counters;  // [0] sum N, [1] NaN channels, [2] min N, [3] max N
    SubsampleParams sub;
    double w_eps;           // paper mode: epsilon added to the denominators of the W ratios (0 = reference)
    DebugProbe probe;
};

// Non-colour channels are staged as floats in slots 0..15: slot -> channel.
__host__ __device__ inline int slot_channel(int s) { return s < 2 ? s : s + 3; }
__host__ __device__ inline int channel_slot(int c) { return c < 2 ? c : c - 3; }

// Task t -> (row channel a, column channel b). Order of rpf.cpp:416-442:
//   t = 4*i + {0,1: r_0,r_1 ; 2,3: p_0,p_1} for feature i (t < 48)
//   t = 48 + 16*i + {0,1: r ; 2,3: p ; 4..15: f_0..f_11} for colour i
//   t = 96 + c: marginal of channel c.
__host__ __device__ inline void task_channels(int t, int &a, int &b) {
    if (t < 48) {
        a = CH_FEAT + (t >> 2);
        const int q = t & 3;
        b = q < 2 ? CH_RND + q : CH_POS + (q - 2);
    } else if (t < N_JOINT) {
        const int u = t - 48;
        a = CH_COL + (u >> 4);
        const int q = u & 15;
        b = q < 2 ? CH_RND + q : (q < 4 ? CH_POS + (q - 2) : CH_FEAT + (q - 4));
    } else {
        a = t - N_JOINT;
        b = -1;
    }
}

}  // namespace rpf
