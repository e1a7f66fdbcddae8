/* include/rpf_b200.h -- C ABI of the B200-native Random Parameter Filtering sample filter.
 *
 * This is the drop-in boundary for the one hot path of tux550/RayTracer-RPF:
 *   void RPFIntegrator::ApplyRPFFilter(SamplingFilm&, const int tileSize, int box_size)
 *       reference: src/custom/rpf.h:91-95, src/custom/rpf.cpp:497-733, called per box size from
 *       RPFIntegrator::Render at src/custom/rpf.cpp:767-775.
 * The replacement host integrator (raytracer-rpf_b200/host/rpf_b200_host.cpp) packs the SamplingFilm
 * into structure-of-arrays planes and calls the entry points below instead of ApplyRPFFilter.
 *
 * Conventions: plain pointers and sizes only; every call returns RPF_OK (0) or a negative rpf_status and
 * leaves a message retrievable with rpf_last_error(). A context is bound to one CUDA device -- or, with
 * rpf_config.n_gpus > 1, to one row band per device -- and is not re-entrant. There is no CPU
 * fallback: without a usable sm_100 device rpf_create fails.
 *
 * Sample layout: channel c of sample s of pixel (x,y) lives at plane[c][(y*width + x)*spp + s].
 * Channel order is SampleData::data (src/custom/sd.h:62-94):
 *   0-1 pFilm.xy | 2-4 L.rgb | 5-6 pLens.xy | 7-9 n0 | 10-12 p0 | 13-15 n1 | 16-18 p1
 */
#ifndef RPF_B200_H
#define RPF_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RPF_NUM_CHANNELS 19
#define RPF_MAX_GPUS 8

typedef enum rpf_status {
    RPF_OK = 0,
    RPF_ERR_INVALID = -1,     /* bad argument */
    RPF_ERR_CUDA = -2,        /* CUDA runtime error (message has the CUDA string) */
    RPF_ERR_NO_DEVICE = -3,   /* no CUDA device / not an sm_100 part */
    RPF_ERR_UNSUPPORTED = -4, /* box*box*spp above rpf_max_neighbourhood() (see message) */
    RPF_ERR_STATE = -5        /* call order (e.g. filter before upload) */
} rpf_status;

/* Behaviour switches relative to the verbatim reference ("REF_DEFINED" = all three):
 * the reference reads out of bounds at rpf.cpp:464 and divides 0/0 at rpf.cpp:464-470. */
#define RPF_FLAG_P1_BETA_AXIS 1   /* W_c_fk[j] uses D_c_f[j] = sum_i MI(c_i,f_j) (rpf.cpp:459 comment, :464 bug) */
#define RPF_FLAG_P2_ZERO_GUARD 2  /* den==0 ? 0 : num/den in W_c_fk, W_r_fk, W_r_ck (rpf.cpp:464,465,470) */
#define RPF_FLAG_P3_CLAMP_VAR 4   /* neighbourhood sigma = sqrt(max(0,var)) (ops.h:141) */
#define RPF_FLAGS_DEFINED 7
#define RPF_FLAGS_VERBATIM 0x100    /* none of P1-P3: beta[3..11] is NaN (the reference value is undefined there) */
/* Testing / tuning switch, may be OR-ed into flags: route every box size through the general three-kernel path
 * (membership mask in HBM, shared-atomic histograms) even where the fused shared-memory kernel would serve it. */
#define RPF_FLAG_FORCE_GENERAL 0x200
/* Same, and use the large-neighbourhood dependency kernel (u16 bins, u32 cells, slabbed tables: what box*box*spp > 65,535
 * takes, e.g. 64 spp with boxes 35 / 55) for every box size. Implies RPF_FLAG_FORCE_GENERAL. */
#define RPF_FLAG_FORCE_LARGE 0x400

typedef struct rpf_ctx rpf_ctx;

typedef struct rpf_config {
    int width;   /* sampleBounds extent x (rpf.cpp:503-504) */
    int height;  /* rows held by THIS context (a row band plus its halo rows when sharded) */
    int spp;     /* samples per pixel, constant */
    int device;  /* CUDA device ordinal */
    int flags;   /* RPF_FLAG_* bits; 0 selects RPF_FLAGS_DEFINED, RPF_FLAGS_VERBATIM selects none */
    int debug;   /* 1: keep per-pixel N, alpha, beta, W_r_c of the last iteration for rpf_download_debug */
    /* Row-band sharding inside ONE process (the role of ParallelFor2D over tiles, rpf.cpp:542, parallel.cpp:247-299):
     * n_gpus 0 or 1 = the single device `device`; 2..8 = the frame is split into n_gpus row bands of equal WORK
     * (rpf_band_plan), band g on device_ids[g]. The context then owns one band buffer, one compute stream and one
     * copy stream per device; rpf_filter exchanges the colour halo rows between passes with peer copies over NVLink
     * that overlap the interior rows. Results are bit-identical to n_gpus = 1. */
    int n_gpus;
    int device_ids[RPF_MAX_GPUS];
    /* Channel layout: a sample record is position (2) | colour (3) | n_random random parameters |
     * n_features features, and rpf_upload takes 5 + n_random + n_features planes in that order. 0 / 0 = the reference's
     * record, 2 random parameters (pLens.xy) and 12 features (n0, p0, n1, p1; sd.h:46,62-94). This build also carries
     * 5 / 12 (+ time and the light sample u, v as random parameters) and 5 / 15 (+ a texture colour as three more features);
     * rpf_layout_supported tells. The wider layouts take the general path: box*box*spp <= 65,535 and spp < 16. Every formula
     * is the reference's with the loop bounds SD_N_RANDOM / SD_N_FEATURES (sd.h:40-43) lifted to the layout's counts. */
    int n_random;
    int n_features;
    int max_box;  /* n_gpus > 1 only: largest box size rpf_filter will be given; every band keeps b_delta(max_box) halo
                     rows of all planes. 0 = 55 (north_star's schedule 55/35/17/7). */
} rpf_config;

typedef struct rpf_stats_t {
    long long pixels_filtered;   /* pixel-iterations since create/reset */
    long long nbh_sum;           /* sum of neighbourhood sizes N over those pixel-iterations ("RPF/Neighborhood Size", rpf.cpp:588) */
    int nbh_min, nbh_max;
    long long nan_channels;      /* NaN colour channels written (rpf.cpp:702-705 would exit(1); here: counted) */
    long long kernel_launches;   /* device kernels launched by this context */
    double last_filter_ms;       /* device time of the last rpf_filter / rpf_filter_rows call (CUDA events; the slowest band when sharded) */
    long long halo_bytes;        /* colour bytes sent to neighbouring row bands since create/reset (0 on an unsharded context) */
} rpf_stats_t;

/* Replaces the SamplingFilm container (sample_film.h:28-45): allocate device planes for width*height*spp samples. */
int rpf_create(rpf_ctx **out, const rpf_config *cfg);
void rpf_destroy(rpf_ctx *ctx);
const char *rpf_last_error(const rpf_ctx *ctx); /* ctx may be NULL: last error of a failed rpf_create */

/* Replaces SamplingFilm::AddSample / MergeSamplingTile (sample_film.cpp:48-66): copy 19 host planes
 * (each width*height*spp floats) to the device. planes[2..4] seed the working colours. */
int rpf_upload(rpf_ctx *ctx, const float *const planes[RPF_NUM_CHANNELS]);
/* Same, from device-resident planes (synthetic benchmarks, multi-GPU bands). */
int rpf_upload_device(rpf_ctx *ctx, const float *const d_planes[RPF_NUM_CHANNELS]);
/* Working colours <- planes[2..4] of the last upload, on the device (stream ordered): filter the resident frame again. */
int rpf_reset_colors(rpf_ctx *ctx);
/* Optional: overwrite the working colours with doubles (the reference keeps filtered colours in double
 * between iterations, sd.h:205-207). rgb = [3][width*height*spp]. */
int rpf_upload_colors_f64(rpf_ctx *ctx, const double *rgb);

/* Replaces the loop at rpf.cpp:767-775: one ApplyRPFFilter pass per entry of box_sizes over the whole
 * context, Jacobi style (rpf.cpp:527,732). sigma_fc_seed is the literal of rpf.cpp:533 (0.002). */
int rpf_filter(rpf_ctx *ctx, const int *box_sizes, int n_iter, double sigma_fc_seed);

/* One ApplyRPFFilter pass restricted to pixel rows [y0,y1) of this context (row-band sharding). Reads the
 * current colours, writes the next colours for those rows; rpf_swap_colors makes them current. Rows outside
 * [y0,y1) of the next buffer are whatever the caller puts there (halo exchange) or stale. */
int rpf_filter_rows(rpf_ctx *ctx, int box_size, double sigma_fc_seed, int y0, int y1);
/* Same for two row ranges at once, [y0,y1) and [y2,y3) with y2 >= y1, in ONE set of kernel launches: the rows next to the
 * upper and the lower edge of a band, which must be finished (and on their way to the neighbouring bands) before the
 * interior rows -- two thin slabs launched separately would leave most of the GPU idle. */
int rpf_filter_rows2(rpf_ctx *ctx, int box_size, double sigma_fc_seed, int y0, int y1, int y2, int y3);
int rpf_swap_colors(rpf_ctx *ctx);
/* Optional, not in the reference (which takes every sample of the box, rpf.cpp:556-586, README.md:27): random neighbour
 * subsampling as in Sen & Darabi. A neighbour sample that passed the 3-sigma test stays with probability
 * keep_probability, decided by the counter-based RNG of include/rpf_subsample.h from (rng_seed, box size, centre pixel,
 * neighbour pixel, sample number) in GLOBAL coordinates -- the CPU oracle and every row band draw the same stream.
 * global_row0 = global row of this context's row 0. keep_probability >= 1 switches it off (the default). */
int rpf_set_subsampling(rpf_ctx *ctx, double keep_probability, unsigned long long rng_seed, int global_row0);
/* Paper mode (Sen & Darabi 2012). None of it is in the reference (contrast rpf.cpp:531-533,575-583,
 * 464-470,767): every field 0 -- the state after rpf_create, and what a NULL pointer restores -- is the reference's
 * arithmetic bit for bit. The CPU restatement of each option is in oracle/rpf_oracle.cpp.
 *   w_epsilon         added to the denominators of W_c_fk, W_r_fk and W_r_ck (rpf.cpp:464,465,470): no 0/0, and
 *                     dependencies too small to measure give weights near 0 instead of noise ratios
 *   min_sigma         low-variance guard of the 3-sigma test: the threshold is max(factor * sigma, min_sigma), so a
 *                     pixel whose samples agree exactly (sigma = 0) still accepts neighbours that differ by rounding
 *   world_pos_sigmas  factor for the world-position features p0, p1 (30 in the paper); 0 = 3 like every feature
 * Takes effect from the next pass; the per-pixel thresholds are rebuilt when min_sigma / world_pos_sigmas change. */
typedef struct rpf_paper_opts {
    double w_epsilon;
    double min_sigma;
    double world_pos_sigmas;
} rpf_paper_opts;
int rpf_set_paper_mode(rpf_ctx *ctx, const rpf_paper_opts *opts);
/* rpf_filter with one sigma_fc_seed per iteration (the paper lowers the variance from pass to pass; the reference
 * uses the literal 0.002 of rpf.cpp:533 for every box size). Same as n_iter calls of rpf_filter with one box each. */
int rpf_filter_sigmas(rpf_ctx *ctx, const int *box_sizes, const double *sigma_seeds, int n_iter);
/* Paper-mode spike removal on the CURRENT colours, rows [y0,y1) (y1 <= 0: height), in place and stream ordered: per
 * pixel and channel the sequential mean and population sigma of the spp samples; a sample with |c - mean| >
 * k_sigmas * sigma in any channel takes the pixel mean in all three (the paper uses k = 1 after the last pass).
 * Local to a pixel: row bands call it on their own rows, no exchange. */
int rpf_hdr_clamp(rpf_ctx *ctx, double k_sigmas, int y0, int y1);
/* Device pointers to the colour planes, [3][width*height*spp] doubles: which = 0 current, 1 next.
 * For halo exchange by the host (NCCL send/recv or peer copies between row bands). */
int rpf_color_device_ptr(rpf_ctx *ctx, int which, double **out);
/* The CUDA stream (cudaStream_t as void*) the context launches on, so callers can order their copies. */
int rpf_stream(rpf_ctx *ctx, void **out);
int rpf_synchronize(rpf_ctx *ctx);

/* Replaces reading samplingFilm.samples back at rpf.cpp:780-786. Any pointer may be NULL.
 * rpf_download_colors, rpf_reconstruct_box and rpf_reconstruct_film use the NEXT colour buffer as conversion scratch: in the
 * row protocol (rpf_filter_rows ... rpf_swap_colors) call them only after rpf_swap_colors, never between a rpf_filter_rows and
 * its swap, or the freshly filtered rows are overwritten. rpf_download_colors_f64 has no such restriction.
 * rpf_upload / rpf_upload_colors_f64 return once the copies are ENQUEUED on the context's stream: from pageable memory that
 * is after the data has left the host buffer, from page-locked memory (rpf_host_alloc) the buffer must stay untouched until
 * rpf_synchronize or any download has returned. */
int rpf_download_colors(rpf_ctx *ctx, float *r, float *g, float *b);
int rpf_download_colors_f64(rpf_ctx *ctx, double *rgb); /* [3][width*height*spp] */
/* Per-pixel (index y*width+x) diagnostics of the last pass; needs cfg.debug. Any pointer may be NULL.
 * nbh = N (rpf.cpp:588), alpha [..][3] (rpf.cpp:475), beta [..][12] (rpf.cpp:479), wrc (rpf.cpp:483-487). */
int rpf_download_debug(rpf_ctx *ctx, int *nbh, double *alpha, double *beta, double *wrc);

/* Component-level probe of one pixel (x, y in context coordinates; negative: off): the next passes leave the 96
 * MutualInformation values of that pixel (mi.cpp:45-90; order of rpf.cpp:416-442: feature i against r0, r1, p0, p1 at
 * 4*i + {0..3}, then colour i against r0, r1, p0, p1, f0..f11 at 48 + 16*i + {0..15}), its N, its bin count
 * (mi.cpp:53-58) and the 19 marginal histograms (mi.cpp:5-21; integer counts, exact). marginals: [19][bins_cap]. */
int rpf_set_debug_pixel(rpf_ctx *ctx, int x, int y);
int rpf_download_debug_mi(rpf_ctx *ctx, double *mi /* [96] */, int *n, int *bins, unsigned *marginals, int bins_cap);

/* Film reconstruction for pbrt's default box filter of radius 0.5 (filters/box.cpp:43-47, film.h:121-161):
 * per-pixel mean of the current colours weighted by nothing but rayWeight==1; rgb = [height*width][3] floats. */
int rpf_reconstruct_box(rpf_ctx *ctx, float *rgb);

/* Film reconstruction for ANY pbrt filter: replaces the serial tail of RPFIntegrator::Render,
 * rpf.cpp:776-804 = Film::GetFilmTile + FilmTile::AddSample per sample (core/film.h:121-161) + Film::MergeFilmTile
 * (core/film.cpp:117-130) + the arithmetic of Film::WriteImage (core/film.cpp:169-205: XYZ->RGB, / filterWeightSum,
 * max(0,.), * scale), all in Float with the reference's operation order: bit-identical to the reference image.
 * The caller hands over what the Film holds: bounds, filter radius, the 16x16 filter table, scale, luminance clamp. */
typedef struct rpf_film_desc {
    int sample_x0, sample_y0;   /* raster coordinates of the context's sample pixel (0,0): Film::GetSampleBounds().pMin */
    int pixel_x0, pixel_y0, pixel_x1, pixel_y1; /* Film::croppedPixelBounds */
    float radius_x, radius_y;   /* filter->radius */
    float filter_table[256];    /* Film::filterTable (core/film.cpp:69-77), see rpf_film_filter_table */
    float scale;                /* Film::scale */
    float max_sample_luminance; /* Film::maxSampleLuminance (INFINITY by default) */
} rpf_film_desc;
/* ray_weight: host [width*height*spp] (SampleData::rayWeight, sd.h:59) or NULL for 1. rgb: host
 * [(pixel_y1-pixel_y0)*(pixel_x1-pixel_x0)][3] floats, the buffer Film::WriteImage hands to the image writer. */
int rpf_reconstruct_film(rpf_ctx *ctx, const rpf_film_desc *film, const float *ray_weight, float *rgb);
/* Host helper: the Film constructor's filter table (core/film.cpp:69-77) for pbrt's box (kind 0, filters/box.cpp:41)
 * or gaussian (kind 1, filters/gaussian.h:52-66, gaussian.cpp:41-43) filter. Other filters: evaluate them yourself. */
int rpf_film_filter_table(int kind, float radius_x, float radius_y, float alpha, float table[256]);

/* ---- row bands: pixels are independent within a pass (Jacobi, rpf.cpp:527,732), so the frame splits
 * along y; features never change (rpf.cpp:715 writes colours only), so a band holds its rows plus b_delta(largest box)
 * halo rows of all 19 planes once, and between passes only the 3 colour planes of the b_delta(next box) rows next to a
 * band edge move to the neighbouring band.
 *
 * rpf_band_plan (host only, no device): first owned row of each of n_bands bands, own0[0..n_bands] with own0[n_bands] =
 * height, so that every band carries about the same WORK: a row's weight is sum over passes of box * (0.9 box + 0.1 rows
 * of the box inside the frame) -- rows near the top and bottom edge see truncated boxes and are slightly cheaper. Every band owns
 * at least b_delta(largest box after the first pass) rows, so the exchange after a pass talks to adjacent bands only;
 * RPF_ERR_INVALID if the frame is too short for that. */
int rpf_band_plan(int height, int n_bands, const int *box_sizes, int n_iter, int *own0 /* [n_bands + 1] */);

/* Same with a MEASURED weight per row (row_weight[height], e.g. a band's step time divided by its rows, from one untimed
 * calibration step with the geometric plan) instead of the geometric model; NULL = rpf_band_plan. */
int rpf_band_plan_weighted(int height, int n_bands, const int *box_sizes, int n_iter, const double *row_weight, int *own0);

/* One band per PROCESS (torchrun-style launch, one process per GPU): this context holds global rows [buf0, buf0 +
 * cfg.height) of a frame of global_height rows and owns rows [own0, own1). After rpf_band_attach, rpf_filter /
 * rpf_filter_sigmas run the band schedule: per pass the rows next to a band edge are filtered first, their colours go
 * to the neighbouring ranks with ncclSend / ncclRecv on the context's copy stream while the interior rows are filtered,
 * and no host synchronisation happens inside the schedule. NCCL is resolved at run time (dlopen of libnccl.so.2: the
 * one the process already loaded, e.g. torch's); nccl_unique_id = the 128 bytes of ncclUniqueId, created with
 * rpf_band_unique_id on rank 0 and handed to every rank by the launcher. rpf_download_* return the whole local buffer;
 * only the owned rows are meaningful. */
typedef struct rpf_band {
    int global_height;
    int buf0;        /* global row of this context's row 0 */
    int own0, own1;  /* owned global rows */
    int rank, world;
} rpf_band;
int rpf_band_unique_id(void *out128);
int rpf_band_attach(rpf_ctx *ctx, const rpf_band *band, const void *nccl_unique_id);

/* Page-locked host memory for the planes handed to rpf_upload and the buffers rpf_download_* fill: uploads from it are
 * asynchronous single DMA transfers (the reference's host container is nested std::vectors, sample_film.h:28-45). */
int rpf_host_alloc(void **out, size_t bytes);
void rpf_host_free(void *p);

int rpf_get_stats(rpf_ctx *ctx, rpf_stats_t *out);
int rpf_reset_stats(rpf_ctx *ctx);

/* Per-kernel device time (the reference bills the whole filter to one pbrt profiler category).
 * With profiling on, every kernel launch is bracketed by CUDA events on the context's stream; totals accumulate
 * until rpf_reset_stats. Kernels: per-pixel feature statistics (rpf.cpp:302-353), the fused small-neighbourhood
 * kernel, and the four kernels of the general path (membership, ordered statistics, dependency estimation, blend). */
enum { RPF_K_PIXEL_STATS = 0, RPF_K_FUSED = 1, RPF_K_MEMBERSHIP = 2, RPF_K_ORDERED_STATS = 3, RPF_K_DEPENDENCY = 4, RPF_K_BLEND = 5,
       RPF_NUM_KERNELS = 6 };
typedef struct rpf_kernel_times_t {
    double ms[RPF_NUM_KERNELS];
    long long launches[RPF_NUM_KERNELS];
} rpf_kernel_times_t;
int rpf_set_profiling(rpf_ctx *ctx, int on);
int rpf_get_kernel_times(rpf_ctx *ctx, rpf_kernel_times_t *out);

/* Channel layout of a context: planes rpf_upload expects (19 for the reference's record), MutualInformation pairs per pixel
 * (n_features * (n_random + 2) + 3 * (n_random + 2 + n_features): 96 for the reference, rpf.cpp:416-442). ctx may be NULL:
 * the reference layout. rpf_layout_supported: 1 if this build has kernels for the layout (0 / 0 = the reference's). */
int rpf_num_channels(const rpf_ctx *ctx);
int rpf_num_pairs(const rpf_ctx *ctx);
int rpf_layout_supported(int n_random, int n_features);

/* Largest neighbourhood (box*box*spp candidate samples) this build can filter, for error messages and tests. */
int rpf_max_neighbourhood(void);
/* Largest box*box*spp served by the fused shared-memory kernel (one CTA per pixel); larger neighbourhoods take the
 * general path. Default 512 (measured crossover on B200), at most 1279; the environment variable RPF_FUSED_MAX
 * overrides. Both paths implement rpf.cpp:556-723 with the same results. */
int rpf_fused_max_candidates(void);

/* One filter pass that ALLOCATES NOTHING -- no counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it). It is ApplyRPFFilter's per-pixel body (rpf.cpp:527-723) for ONE box size through the fused kernel, with all
 * device scratch (per-pixel feature mean / 3 sigma, the n ln n tables, the counters) in file-scope static __device__
 * arrays of the library: 12 MiB that every GPU which loads the library holds until the process exits.
 *   d_planes      [19][width*height*spp] float32, DEVICE memory of the CURRENT device, the layout of rpf_upload
 *   d_colors_in   [3][width*height*spp] float64, the colours entering the pass (first pass: planes 2..4 as doubles)
 *   d_colors_out  [3][width*height*spp] float64, the colours leaving it (must not alias d_colors_in: Jacobi pass)
 *   flags         as rpf_config.flags (0 = RPF_FLAGS_DEFINED);  cuda_stream: a cudaStream_t (NULL = the default stream)
 * The call only enqueues three kernels on cuda_stream and returns; results equal rpf_filter's for the same box bit for bit.
 * Limits: width*height <= RPF_NOALLOC_MAX_PIXELS and box*box*spp <= rpf_fused_max_candidates(); beyond either the call
 * returns RPF_ERR_UNSUPPORTED with a message (rpf_last_error(NULL)) and launches nothing. It calls no cudaMalloc,
 * cudaFree or allocating library routine; the first call on each device also fetches the arrays' device addresses,
 * uploads two constant tables (21 KiB, synchronous) and sets the kernel's shared-memory attribute -- make that call
 * outside a graph capture. All calls on one device share the arrays, which keep the last call's contents (the call
 * rewrites what it reads): calls that OVERLAP on the GPU (different streams without an ordering event) race. */
#define RPF_NOALLOC_MAX_PIXELS 65536
int rpf_filter_pass_noalloc(const float *d_planes, const double *d_colors_in, double *d_colors_out, int width, int height,
                            int spp, int box, double sigma_fc_seed, int flags, void *cuda_stream);

/* Deterministic synthetic feature buffer, host side, float32 values.
 * kind 0: all-noisy; kind 1: degenerate (flat normals, black background, duplicated samples).
 * (x_off,y_off) shift the generator so a row band of a larger frame can be generated on its own. */
int rpf_synth_planes(int width, int height, int spp, unsigned long long seed, int kind, int x_off, int y_off,
                     float *planes /* [19][width*height*spp] */);

/* Same for a wider channel layout (rpf_config.n_random / n_features): planes [5 + n_random + n_features][width*height*spp].
 * The 19 reference channels are those of rpf_synth_planes. kind 2 = kind 0 with every extra random parameter constant. */
int rpf_synth_planes_layout(int width, int height, int spp, unsigned long long seed, int kind, int x_off, int y_off,
                            int n_random, int n_features, float *planes);

/* Feature-buffer files: lossless dump / load of what the integrator emits per sample, replacing
 * the lossy debug EXRs of visualizeSF (rpf.cpp:37-101). Host side only; layout in raytracer-rpf_b200/csrc/rpf_io.cpp. */
typedef struct rpf_buffer_info {
    int width, height, spp;
    int sample_x0, sample_y0; /* sampleBounds.pMin */
    int has_ray_weight;       /* out of rpf_buffer_info_read; ignored by rpf_buffer_write */
    int n_random, n_features; /* channel layout of the file: 5 + n_random + n_features planes (0 / 0 on write: the reference's 2 / 12) */
} rpf_buffer_info;
int rpf_buffer_write(const char *path, const rpf_buffer_info *info, const float *const planes[RPF_NUM_CHANNELS],
                     const float *ray_weight /* may be NULL */);
int rpf_buffer_info_read(const char *path, rpf_buffer_info *info);
/* planes: [5 + n_random + n_features][width*height*spp] contiguous (19 planes for the reference's record); ray_weight: [width*height*spp] or NULL (filled with 1 if the file has none) */
int rpf_buffer_read(const char *path, float *planes, float *ray_weight);

#ifdef __cplusplus
}
#endif
#endif
