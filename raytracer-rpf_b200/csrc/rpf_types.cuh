// rpf_types.cuh -- layout-independent types of the RPF kernels (included once; rpf_common.cuh and the kernel headers are
// compiled once per channel layout, see rpf_layouts.cuh).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rpf_subsample.h"

namespace rpf_types {

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
    double *mi;           // [N_JOINT] MI values, then N, then bins
    unsigned *marg;       // [NCH][PROBE_BINS]
};

// Device view of the frame: NCH float planes, each S = W*H*P samples, index (y*W+x)*P+s.
struct FrameView {
    const float *planes;   // [NCH][S]
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

}  // namespace rpf_types
