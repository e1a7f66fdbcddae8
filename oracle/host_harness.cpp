// oracle/host_harness.cpp -- TEST INFRASTRUCTURE. Drives the product's host-side C++ (raytracer-rpf_b200/host/
// rpf_b200_host.cpp) exactly as the reference's Render would: through a real SamplingFilm built from the reference's own
// sample_film.cpp / sd.h (compiled from /root/reference behind oracle/ref_stubs). Built into oracle/_ref/ by the Makefile.
#include "pbrt.h"

#include "rpf_b200_host.h"

using namespace pbrt;

extern "C" int host_apply_rpf_b200(int W, int H, int P, const float *planes, double *colors, const int *boxes, int n_boxes,
                                   double seed, int flags, double *mean_nbh, long long *nan_channels, char *err, int err_len) {
    Bounds2i bounds(Point2i(0, 0), Point2i(W, H));
    SamplingFilm sf(bounds);
    const size_t S = static_cast<size_t>(W) * H * P;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x)
            for (int s = 0; s < P; ++s) {
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                SampleData sd;
                for (int c = 0; c < 19; ++c) sd.data[c] = planes[c * S + i];
                for (int k = 0; k < 3; ++k) sd.data[2 + k] = colors[k * S + i];
                sd.rayWeight = 1;
                sf.AddSample(Point2i(x, y), sd);
            }
    RPFB200Options opt;
    opt.boxSizes.assign(boxes, boxes + n_boxes);
    opt.sigmaSeed = seed;
    opt.flags = flags;
    RPFB200Report rep;
    std::string error;
    if (!ApplyRPFFilterB200(sf, opt, &rep, &error)) {
        if (err && err_len > 0) { std::snprintf(err, err_len, "%s", error.c_str()); }
        return -1;
    }
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x)
            for (int s = 0; s < P; ++s) {
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                for (int k = 0; k < 3; ++k) colors[k * S + i] = sf.samples[x][y][s].data[2 + k];
            }
    if (mean_nbh) *mean_nbh = rep.meanNeighbourhood;
    if (nan_channels) *nan_channels = rep.nanChannels;
    return 0;
}

// Filter + Film reconstruction through the host C++: a gaussian pbrt filter described to MakeFilmDescB200 by a stand-in
// with the interface of pbrt::Filter (radius, Evaluate), rayWeight carried by the SampleData records.
namespace {
struct HarnessGaussian {  // interface of core/filter.h:47-58; values supplied by the test (the reference filter's table)
    struct { float x, y; } radius;
    const float *table;
    mutable int calls = 0;
    float Evaluate(const Point2f &) const { return table[calls++]; }
};
}  // namespace

extern "C" int host_apply_rpf_film_b200(int W, int H, int P, const float *planes, const float *ray_weight, const int *boxes, int n_boxes,
                                        double seed, float rx, float ry, const float *table, float scale, float max_lum,
                                        float *rgb /* [H*W][3] */, char *err, int err_len) {
    Bounds2i bounds(Point2i(0, 0), Point2i(W, H));
    SamplingFilm sf(bounds);
    const size_t S = static_cast<size_t>(W) * H * P;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x)
            for (int s = 0; s < P; ++s) {
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                SampleData sd;
                for (int c = 0; c < 19; ++c) sd.data[c] = planes[c * S + i];
                sd.rayWeight = ray_weight ? ray_weight[i] : 1;
                sf.AddSample(Point2i(x, y), sd);
            }
    RPFB200Options opt;
    opt.boxSizes.assign(boxes, boxes + n_boxes);
    opt.sigmaSeed = seed;
    HarnessGaussian filt;
    filt.radius.x = rx; filt.radius.y = ry; filt.table = table;
    const rpf_film_desc fd = MakeFilmDescB200(filt, bounds, bounds, scale, max_lum);
    std::vector<float> out;
    std::string error;
    if (!ApplyRPFFilterAndFilmB200(sf, opt, fd, &out, nullptr, &error)) {
        if (err && err_len > 0) { std::snprintf(err, err_len, "%s", error.c_str()); }
        return -1;
    }
    std::copy(out.begin(), out.end(), rgb);
    return 0;
}

// The dump hook over a real reference SamplingFilm (host code only, no GPU).
extern "C" int host_dump_sampling_film(int W, int H, int P, const float *planes, const float *ray_weight, const char *path) {
    Bounds2i bounds(Point2i(0, 0), Point2i(W, H));
    SamplingFilm sf(bounds);
    const size_t S = static_cast<size_t>(W) * H * P;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x)
            for (int s = 0; s < P; ++s) {
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                SampleData sd;
                for (int c = 0; c < 19; ++c) sd.data[c] = planes[c * S + i];
                sd.rayWeight = ray_weight[i];
                sf.AddSample(Point2i(x, y), sd);
            }
    std::string error;
    return DumpSamplingFilmB200(sf, 0, 0, path, &error) ? 0 : -1;
}
