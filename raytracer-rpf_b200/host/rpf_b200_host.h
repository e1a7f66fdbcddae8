// rpf_b200_host.h -- host side (C++) of the B200 RPF filter: the drop-in for the loop
//     for (int box_size : box_sizes) ApplyRPFFilter(samplingFilm, tileSize, box_size);
// of RPFIntegrator::Render (reference src/custom/rpf.cpp:767-775). It packs the reference's SamplingFilm
// (src/custom/sample_film.h:28-45, samples[x - pMin.x][y - pMin.y] = vector<SampleData>) into the structure-of-arrays
// planes of include/rpf_b200.h, runs the passes on the GPU through the C ABI and writes the filtered colours back
// into the same SamplingFilm, which is what ApplyRPFFilter does in place (rpf.cpp:732).
// Compiles against the reference's own headers; INTEGRATION.md shows the three-line change in rpf.cpp.
#ifndef RPF_B200_HOST_H
#define RPF_B200_HOST_H

#include <cstdint>
#include <functional>
#include <string>
#include <vector>

#include "custom/sample_film.h"
#include "rpf_b200.h"

namespace pbrt {

// What the reference's 19-double SampleData (sd.h:46) has no room for: more random parameters (time, the
// light sample) and more features (a texture colour) per sample, kept beside the SamplingFilm. The device record becomes
// position | colour | pLens, extra random parameters | n0, p0, n1, p1, extra features (rpf_config.n_random / n_features).
struct RPFB200WideRecord {
    int nRandom = 0, nFeatures = 0;   // extra random parameters / extra features per sample
    std::vector<float> values;        // [(y*W + x)*spp + s][nRandom + nFeatures], random parameters first; (x, y) relative to sampleBounds.pMin
};

struct RPFB200Options {
    std::vector<int> boxSizes{7};  // rpf.cpp:767 ships {7}; the comment there lists {55, 35, 17, 7}
    double sigmaSeed = 0.002;      // rpf.cpp:533
    int device = 0;                // CUDA device ordinal (one GPU)
    std::vector<int> devices;      // two or more ordinals: one row band per device inside this process (rpf_config.n_gpus)
    bool debugStats = false;       // keep the per-pixel neighbourhood sizes of the last pass (RPFB200Report::neighbourhood)
    // Runs body(i) for i in [0, count) on the caller's thread pool -- the integrator passes pbrt's ParallelFor
    // (core/parallel.h:101), the pool FillSampleFilm used (rpf.cpp:235). Empty: a std::thread per hardware thread.
    std::function<void(const std::function<void(int64_t)> &, int64_t)> parallelFor;
    int flags = 0;                 // RPF_FLAG_* of include/rpf_b200.h; 0 = REF_DEFINED
    double keepProbability = 1.0;  // < 1: random neighbour subsampling (rpf_set_subsampling; not the reference's semantics)
    unsigned long long rngSeed = 0;
    // Paper mode (rpf_set_paper_mode / rpf_filter_sigmas / rpf_hdr_clamp; none of it is the reference's semantics):
    rpf_paper_opts paper{0.0, 0.0, 0.0};  // epsilon of the W ratios, low-variance guard, sigma factor of p0 / p1
    std::vector<double> sigmaSeeds;       // one per box size when not empty, instead of sigmaSeed for all
    double hdrClampSigmas = 0.0;          // > 0: spike removal after the last pass
    const RPFB200WideRecord *wide = nullptr;  // not null: filter the wider record (the library must carry that layout)
};

struct RPFB200Report {
    long long nanChannels = 0;       // rpf.cpp:702-705 would exit(1); here they are counted
    double meanNeighbourhood = 0;    // "RPF/Neighborhood Size" (rpf.cpp:588)
    int minNeighbourhood = 0, maxNeighbourhood = 0;
    double filterMs = 0;             // device time of all passes
    // host wall-clock milliseconds around the device work: SamplingFilm -> pinned structure-of-arrays planes, the uploads,
    // the passes (filter call to completion), the download, planes -> SamplingFilm
    double packMs = 0, uploadMs = 0, filterWallMs = 0, downloadMs = 0, unpackMs = 0;
    int gpus = 1;
    std::vector<int> neighbourhood;  // per pixel, index y*W+x, last pass (only with RPFB200Options::debugStats)
};

// Returns true on success. On failure nothing in samplingFilm is modified and *error holds the reason
// (no CUDA device, ragged sample counts, neighbourhood above rpf_max_neighbourhood(), CUDA errors).
bool ApplyRPFFilterB200(SamplingFilm &samplingFilm, const RPFB200Options &options, RPFB200Report *report,
                        std::string *error);

// Same, and the Film reconstruction that follows in Render (rpf.cpp:776-804: FilmTile::AddSample per sample with
// SampleData::rayWeight, MergeFilmTile, WriteImage arithmetic) also runs on the GPU while the samples are resident:
// *rgb receives the buffer Film::WriteImage would hand to pbrt::WriteImage (core/film.cpp:210), bit for bit,
// [(pixel_y1-pixel_y0)*(pixel_x1-pixel_x0)][3] floats. The caller then calls pbrt::WriteImage(film->filename, rgb->data(),
// film->croppedPixelBounds, film->fullResolution) itself.
bool ApplyRPFFilterAndFilmB200(SamplingFilm &samplingFilm, const RPFB200Options &options, const rpf_film_desc &film,
                               std::vector<float> *rgb, RPFB200Report *report, std::string *error);

// Dump hook: writes the SamplingFilm as a feature-buffer file (include/rpf_b200.h rpf_buffer_write;
// lossless, unlike the debug EXRs of visualizeSF, rpf.cpp:37-101) so a frame rendered once can be filtered anywhere.
bool DumpSamplingFilmB200(const SamplingFilm &samplingFilm, int sampleX0, int sampleY0, const std::string &path, std::string *error,
                          const RPFB200WideRecord *wide = nullptr);

// What the GPU needs to know about a pbrt Film, from its public parts: the filter (core/film.h:78, filter.h:47-58), the
// bounds, and the two private scalars the caller passes in (Film "scale" and "maxsampleluminance" scene parameters,
// core/film.cpp:213-251). The table is the Film constructor's (core/film.cpp:69-77). A template so that this header does
// not depend on pbrt's core/filter.h: FilterT = pbrt::Filter.
template <class FilterT, class Bounds2iT>
rpf_film_desc MakeFilmDescB200(const FilterT &filter, const Bounds2iT &sampleBounds, const Bounds2iT &croppedPixelBounds, float scale,
                               float maxSampleLuminance) {
    rpf_film_desc d{};
    d.sample_x0 = sampleBounds.pMin.x; d.sample_y0 = sampleBounds.pMin.y;
    d.pixel_x0 = croppedPixelBounds.pMin.x; d.pixel_y0 = croppedPixelBounds.pMin.y;
    d.pixel_x1 = croppedPixelBounds.pMax.x; d.pixel_y1 = croppedPixelBounds.pMax.y;
    d.radius_x = filter.radius.x; d.radius_y = filter.radius.y;
    int offset = 0;
    for (int y = 0; y < 16; ++y)
        for (int x = 0; x < 16; ++x, ++offset) {
            decltype(filter.radius.x) px = (x + 0.5f) * filter.radius.x / 16, py = (y + 0.5f) * filter.radius.y / 16;
            d.filter_table[offset] = filter.Evaluate({px, py});
        }
    d.scale = scale;
    d.max_sample_luminance = maxSampleLuminance;
    return d;
}

}  // namespace pbrt
#endif
