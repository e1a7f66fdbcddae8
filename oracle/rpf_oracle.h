/* oracle/rpf_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement ("REF_DEFINED") of the Random Parameter
 * Filtering sample filter of tux550/RayTracer-RPF:
 *   src/custom/rpf.cpp:302-353  FillMeanAndStddev
 *   src/custom/rpf.cpp:356-488  ComputeCFWeights
 *   src/custom/rpf.cpp:497-733  ApplyRPFFilter
 *   src/custom/mi.cpp:5-90      histograms + MutualInformation
 *   src/custom/ops.h:44-51,100-144  divideArrays / allLessThan / getMean / getStdDev
 *   src/custom/sd.h:224-235     SampleData::normalized
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker / timed CPU baseline.
 *
 * Pinning status: the reference has no golden vectors for this path.
 * The restatement is pinned against the reference's own sources compiled in place
 * (oracle/_ref/libref_rpf.so, built by oracle/Makefile from /root/reference) and
 * against fixtures generated from that library (tests/golden/).
 */
#ifndef RPF_ORACLE_H
#define RPF_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* Behaviour patches relative to the verbatim reference. */
#define RPF_ORACLE_P1 1 /* beta uses D_c_f[j] = sum_i MI(c_i,f_j) instead of the out-of-bounds D_f_ck[j] (rpf.cpp:464) */
#define RPF_ORACLE_P2 2 /* 0/0 -> 0 in W_c_fk, W_r_fk, W_r_ck (rpf.cpp:464,465,470) */
#define RPF_ORACLE_P3 4 /* neighbourhood sigma = sqrt(max(0, var)) (ops.h:141, neighbourhood statistics only) */
#define RPF_ORACLE_DEFINED 7

#define RPF_NCH 19

typedef struct rpf_oracle_args {
    int width, height, spp;
    const float *planes;     /* [19][W*H*spp], channel order = SampleData::data (sd.h:62-94); index (y*W+x)*spp+s */
    double *colors;          /* in/out [3][W*H*spp]; if NULL on input the colours are taken from planes[2..4] */
    const int *box_sizes;    /* one per iteration (rpf.cpp:767) */
    int n_iter;
    double sigma_fc_seed;    /* rpf.cpp:533 literal 0.002 */
    int flags;               /* RPF_ORACLE_P* */
    int n_threads;           /* <=0: hardware concurrency */
    int y0, y1;              /* rows filtered: [y0,y1); y1<=0 means height. Rows outside keep their colours. */
    int x0, x1;              /* columns filtered: [x0,x1); x1<=0 means width */
    /* optional per-pixel debug outputs of the LAST iteration (may be NULL), index y*W+x */
    int *nbh_size;           /* N */
    double *alpha;           /* [W*H][3]  */
    double *beta;            /* [W*H][12] */
    double *wrc;             /* [W*H]     */
    long long nan_count;     /* out: number of NaN colour channels written (P4: counted, never fatal) */
    /* Neighbour subsampling (not in the reference; include/rpf_subsample.h is the RNG both sides call):
     * keep_q24 = floor(keep probability * 2^24), 0 or >= 2^24 = keep every sample that passes the 3-sigma test. */
    unsigned int subsample_keep_q24;
    unsigned long long subsample_seed;
    int global_row0;         /* global row of local row 0 (row bands) */
    /* Paper mode (Sen & Darabi 2012). None of it is in the reference; all zero = reference semantics.
     *   w_epsilon         added to the denominators of W_c_fk, W_r_fk, W_r_ck (rpf.cpp:464,465,470)
     *   min_sigma         low-variance guard of the 3-sigma test: threshold max(factor * sigma, min_sigma) (rpf.cpp:578-583)
     *   world_pos_sigmas  factor for the world-position features p0, p1 (30 in the paper); 0 = 3 like every other feature
     *   sigma_seeds       [n_iter] sigma_fc_seed of each iteration (NULL: sigma_fc_seed for all, rpf.cpp:533)
     *   hdr_clamp_sigmas  after the last iteration a sample with a colour channel further than this many pixel sigmas from
     *                     its pixel's mean is replaced by the mean (spike removal); 0 = off */
    double w_epsilon, min_sigma, world_pos_sigmas;
    const double *sigma_seeds;
    double hdr_clamp_sigmas;
    /* Channel layout: position (2) | colour (3) | n_random random parameters | n_features features;
     * planes holds 5 + n_random + n_features of them, beta has n_features entries per pixel. 0 / 0 = the reference's record
     * (2 / 12, sd.h:40-46), the only layout with a reference output to pin against. */
    int n_random, n_features;
} rpf_oracle_args;

/* returns 0 on success, <0 on bad arguments */
int rpf_oracle_filter(rpf_oracle_args *a);

/* mi.cpp:45-90 restated; bins = max(1,(int)sqrt(n)) */
double rpf_oracle_mi(const double *x, const double *y, int n);

/* ops.h:111-144 restated for one channel: sequential mean and population sigma (no clamp) */
void rpf_oracle_mean_std(const double *v, int n, double *mean, double *sd);

/* rpf.cpp:356-488 restated: normalised neighbourhood z[n][19] -> alpha[3], beta[12], wrc */
void rpf_oracle_cf_weights(const double *z, int n, int flags, double *alpha, double *beta, double *wrc);

/* Paper-mode spike removal on colours [3][W*H*spp], rows [y0,y1): per pixel and channel the sequential mean and population
 * sigma of the spp samples; a sample with |c - mean| > k_sigmas * sigma in any channel takes the mean in all three. */
void rpf_oracle_hdr_clamp(double *colors, int width, int height, int spp, int y0, int y1, double k_sigmas);

/* Component view of one pixel of one pass (box_sizes[0]): the MutualInformation values (96 for the reference layout) in the order of rpf.cpp:416-442,
 * N, bins (mi.cpp:53-58), the 19 marginal histograms [19][bins_cap] (mi.cpp:5-21) and optionally the normalised
 * neighbourhood z [n][19] (rpf.cpp:603-612) so a test can hand its columns to the reference's own MutualInformation. */
int rpf_oracle_pixel_probe(const rpf_oracle_args *a, int px, int py, double *mi96, int *n_out, int *bins_out, int *marg,
                           int bins_cap, double *z_out, int n_cap);

#ifdef __cplusplus
}
#endif
#endif
