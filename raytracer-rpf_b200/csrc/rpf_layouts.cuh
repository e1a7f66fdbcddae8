// rpf_layouts.cuh -- the channel layouts this build carries (a run-time channel table).
//
// A sample record is  position (2) | colour (3) | random parameters (NR) | features (NF). The reference's SampleData is
// (NR, NF) = (2, 12): pLens.xy | n0, p0, n1, p1 (sd.h:46,62-94). north_star's feature buffer is wider -- time and light
// samples are random parameters too, a texture colour is three more features -- so every kernel header is written against
// NR / NF and compiled once per layout into its own namespace; rpf_config.n_random / n_features pick the layout of a
// context at run time (rpf_api.cu: layout_ops). The reference layout gets every kernel (fused small-neighbourhood kernel,
// large-neighbourhood kernel, wide blend); the wider ones the general path (box*box*spp <= 65,535, spp < 16).
#pragma once
#include "rpf_types.cuh"

// ---- (2, 12): the reference's record, all kernels -------------------------------------------------------------------
#define RPF_NS rpf
#define RPF_NR 2
#define RPF_NF 12
#define RPF_FULL 1
#include "rpf_common.cuh"
#include "rpf_kernels.cuh"
#include "rpf_pixel_kernel.cuh"
#include "rpf_general.cuh"
#include "rpf_membership.cuh"
#include "rpf_large.cuh"
#include "rpf_blend_wide.cuh"
#undef RPF_NS
#undef RPF_NR
#undef RPF_NF
#undef RPF_FULL

// ---- (5, 12): + time, light sample u / v as random parameters -------------------------------------------------------
#define RPF_NS rpf_r5f12
#define RPF_NR 5
#define RPF_NF 12
#define RPF_FULL 0
#include "rpf_common.cuh"
#include "rpf_kernels.cuh"
#include "rpf_general.cuh"
#include "rpf_membership.cuh"
#undef RPF_NS
#undef RPF_NR
#undef RPF_NF
#undef RPF_FULL

// ---- (5, 15): + a texture colour as three more features ---------------------------------------------------------------
#define RPF_NS rpf_r5f15
#define RPF_NR 5
#define RPF_NF 15
#define RPF_FULL 0
#include "rpf_common.cuh"
#include "rpf_kernels.cuh"
#include "rpf_general.cuh"
#include "rpf_membership.cuh"
#undef RPF_NS
#undef RPF_NR
#undef RPF_NF
#undef RPF_FULL
