// rpf_common.cuh -- shared definitions for the sm_100a RPF kernels.
// Channel order and slices follow SampleData::data (reference src/custom/sd.h:62-94,149-212).
//
// RE-INCLUDABLE: the kernel headers are compiled once per CHANNEL LAYOUT (rpf_layouts.cuh), each time inside its own
// namespace RPF_NS with RPF_NR random parameters and RPF_NF features -- the run-time channel table of SURVEY.md 8(f) N3:
//     position (2) | colour (3) | random parameters (RPF_NR) | features (RPF_NF)
// The reference's record is the layout (2, 12): pLens.xy | n0, p0, n1, p1 (sd.h:46,62-94). Wider layouts carry what
// north_star lists beyond it (time and light samples as random parameters, a texture colour as three more features).
// Nothing in here may be guarded by #pragma once; layout-independent types live in rpf_types.cuh.
#include "rpf_types.cuh"

namespace RPF_NS {

using namespace rpf_types;

constexpr int NR = RPF_NR, NF = RPF_NF, NC = 3, NP = 2;
constexpr int NCH = NP + NC + NR + NF;
static_assert(NR >= 1 && NF >= 1 && NCH <= 32, "one warp lane per channel (ordered_stats_kernel): at most 32 channels");
constexpr int CH_POS = 0, CH_COL = 2, CH_RND = 5, CH_FEAT = 5 + NR;
constexpr int NPCOL = NR + NP;               // MI partners of one feature: every random parameter, both positions (rpf.cpp:416-427)
constexpr int NCCOL = NR + NP + NF;          // MI partners of one colour: those and every feature (rpf.cpp:429-442)
constexpr int T0_COL = NF * NPCOL;           // first task of colour 0
constexpr int N_JOINT = NF * NPCOL + NC * NCCOL;  // MI pairs of rpf.cpp:416-442 (96 for the reference's layout)
constexpr int N_TASK = N_JOINT + NCH;   // + one marginal histogram per channel
constexpr int N_GROUP = (N_TASK + 31) / 32;  // task groups of 32 lanes (fused kernel)
constexpr int THREADS = 128;
constexpr int REC19 = (NCH + 2) & ~1;   // doubles per sample of the all-double gather record (channels + pad to an even count)
constexpr int NWZ = NP + NC + NF;       // channels entering the weights: pos(2) + colour(3) + features
constexpr int ZS = (NWZ + 2) & ~1;      // stride of one sample's weight-space coordinates in shared memory (even, >= NWZ + 1)
constexpr int NSLOT = NCH - NC;         // non-colour channels, staged as floats in slot order
constexpr int NQ4 = 2 * ((NSLOT + 7) / 8);  // float4s per sample of the gather record (an even count: 256-bit loads)

constexpr int FLAG_P1 = 1, FLAG_P2 = 2, FLAG_P3 = 4;





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

// Non-colour channels are staged as floats in slots 0..NSLOT-1: slot -> channel.
__host__ __device__ inline int slot_channel(int s) { return s < 2 ? s : s + 3; }
__host__ __device__ inline int channel_slot(int c) { return c < 2 ? c : c - 3; }

// Task t -> (row channel a, column channel b). Order of rpf.cpp:416-442 (reference layout: 48 + 48 tasks):
//   t = NPCOL*i + {q < NR: r_q ; then p_0, p_1} for feature i (t < T0_COL)
//   t = T0_COL + NCCOL*i + {q < NR: r_q ; p_0, p_1 ; f_0..f_(NF-1)} for colour i
//   t = N_JOINT + c: marginal of channel c.
__host__ __device__ inline void task_channels(int t, int &a, int &b) {
    if (t < T0_COL) {
        const int i = t / NPCOL, q = t - i * NPCOL;
        a = CH_FEAT + i;
        b = q < NR ? CH_RND + q : CH_POS + (q - NR);
    } else if (t < N_JOINT) {
        const int u = t - T0_COL, i = u / NCCOL, q = u - i * NCCOL;
        a = CH_COL + i;
        b = q < NR ? CH_RND + q : (q < NPCOL ? CH_POS + (q - NR) : CH_FEAT + (q - NPCOL));
    } else {
        a = t - N_JOINT;
        b = -1;
    }
}

// World positions among the features (p0, p1 of sd.h:62-94: n0 | p0 | n1 | p1 | anything wider): paper mode's sigma factor.
__host__ __device__ inline bool feature_is_world_pos(int k) { return (k >= 3 && k < 6) || (k >= 9 && k < 12); }

}  // namespace RPF_NS
