/* include/rpf_subsample.h -- the shared counter-based RNG behind neighbour subsampling (first step).
 *
 * Sen & Darabi's filter draws a random subset of the samples in the box instead of taking all of them; the reference
 * (src/custom/rpf.cpp:556-586) takes all and says so in README.md:27. With rpf_set_subsampling() a neighbour sample
 * that passed the 3-sigma test stays in the neighbourhood only if this function says so. It is a pure function of
 * (seed, box size, centre pixel, neighbour pixel, sample number) in GLOBAL pixel coordinates, so the CPU oracle, every
 * GPU kernel and every row band of a sharded frame draw identical index streams without exchanging anything.
 * Plain C, also compiled by nvcc for the device.
 */
#ifndef RPF_SUBSAMPLE_H
#define RPF_SUBSAMPLE_H
#include <stdint.h>

#if defined(__CUDACC__)
#define RPF_HD __host__ __device__ __forceinline__
#else
#define RPF_HD static inline
#endif

#define RPF_SUBSAMPLE_ALL (1u << 24) /* keep_q24 value meaning "no subsampling" */

RPF_HD uint64_t rpf_mix64(uint64_t z) { /* splitmix64 finaliser */
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

/* key of a (pass, centre pixel): hoisted out of the candidate loops */
RPF_HD uint64_t rpf_subsample_centre_key(uint64_t seed, int box, int cx, int cy) {
    return rpf_mix64(seed ^ ((uint64_t)(uint32_t)box << 48) ^ rpf_mix64(((uint64_t)(uint32_t)cx << 32) | (uint32_t)cy));
}

/* 1 if sample s of neighbour pixel (nx, ny) stays; keep_q24 = floor(keep probability * 2^24) */
RPF_HD int rpf_subsample_keep(uint64_t centre_key, uint32_t keep_q24, int nx, int ny, int s) {
    const uint64_t k = rpf_mix64((centre_key ^ (((uint64_t)(uint32_t)nx << 32) | (uint32_t)ny)) + 0x632be59bd9b4e019ull * (uint64_t)(s + 1));
    return (uint32_t)(k >> 40) < keep_q24;
}

#endif
