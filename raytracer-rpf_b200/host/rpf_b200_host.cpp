// rpf_b200_host.cpp -- see rpf_b200_host.h. Host C++ only; all device work happens behind include/rpf_b200.h.
#include "rpf_b200_host.h"

#include <algorithm>
#include <chrono>
#include <memory>
#include <thread>

#include "rpf_b200.h"

namespace pbrt {

namespace {
struct CtxDeleter {
    void operator()(rpf_ctx *c) const { rpf_destroy(c); }
};
struct PinnedDeleter {
    void operator()(void *p) const { rpf_host_free(p); }
};
template <class T>
using Pinned = std::unique_ptr<T, PinnedDeleter>;

template <class T>
Pinned<T> pinned(size_t n, std::string *error) {
    void *p = nullptr;
    if (rpf_host_alloc(&p, n * sizeof(T)) != RPF_OK) {
        if (error) *error = std::string("rpf_host_alloc: ") + rpf_last_error(nullptr);
        return Pinned<T>();
    }
    return Pinned<T>(static_cast<T *>(p));
}

bool failed(std::string *error, const std::string &what, const rpf_ctx *ctx) {
    if (error) *error = what + ": " + rpf_last_error(ctx);
    return false;
}

double ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

// body(i) for i in [0, count): the caller's pool, or one std::thread per hardware thread over interleaved indices
void parallel_for(const RPFB200Options &opt, int64_t count, const std::function<void(int64_t)> &body) {
    if (opt.parallelFor) {
        opt.parallelFor(body, count);
        return;
    }
    const int nt = static_cast<int>(std::min<int64_t>(std::max(1u, std::thread::hardware_concurrency()), count));
    if (nt <= 1) {
        for (int64_t i = 0; i < count; ++i) body(i);
        return;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t)
        pool.emplace_back([&, t] { for (int64_t i = t; i < count; i += nt) body(i); });
    for (auto &th : pool) th.join();
}
}  // namespace

static bool run(SamplingFilm &film, const RPFB200Options &opt, const rpf_film_desc *fdesc, std::vector<float> *rgb,
                RPFB200Report *report, std::string *error) {
    const int W = film.getWidth(), H = film.getHeight();
    if (W <= 0 || H <= 0 || film.samples.empty() || film.samples[0].empty()) {
        if (error) *error = "empty SamplingFilm";
        return false;
    }
    // every sampler of the reference yields a constant count per pixel; the device layout needs it
    const size_t spp = film.samples[0][0].size();
    for (int x = 0; x < W; ++x)
        for (int y = 0; y < H; ++y)
            if (film.samples[x][y].size() != spp || spp == 0) {
                if (error) *error = "SamplingFilm has a ragged or zero sample count at pixel (" + std::to_string(x) + "," +
                                    std::to_string(y) + ")";
                return false;
            }
    const size_t S = static_cast<size_t>(W) * H * spp;
    const int xr = opt.wide ? opt.wide->nRandom : 0, xf = opt.wide ? opt.wide->nFeatures : 0, xs = xr + xf;
    const int nch = RPF_NUM_CHANNELS + xs;
    if (opt.wide && (xr < 0 || xf < 0 || opt.wide->values.size() != S * xs)) {
        if (error) *error = "RPFB200WideRecord: values must hold (extra random + extra features) floats for every sample";
        return false;
    }
    if (opt.wide && !rpf_layout_supported(2 + xr, 12 + xf)) {
        if (error) *error = "this build of librpf_b200 has no kernels for " + std::to_string(2 + xr) + " random parameters and " +
                            std::to_string(12 + xf) + " features";
        return false;
    }
    if (fdesc && opt.devices.size() > 1) {
        if (error) *error = "Film reconstruction on the device needs a single-GPU context (RPFB200Options::devices)";
        return false;
    }

    // the context first: without a usable device nothing else needs to happen (there is no CPU path)
    rpf_config cfg{};
    cfg.width = W;
    cfg.height = H;
    cfg.spp = static_cast<int>(spp);
    cfg.device = opt.device;
    cfg.flags = opt.flags;
    cfg.debug = opt.debugStats ? 1 : 0;
    if (opt.wide) { cfg.n_random = 2 + xr; cfg.n_features = 12 + xf; }
    if (opt.devices.size() > 1) {
        if (opt.devices.size() > RPF_MAX_GPUS) {
            if (error) *error = "RPFB200Options::devices lists more than RPF_MAX_GPUS devices";
            return false;
        }
        cfg.n_gpus = static_cast<int>(opt.devices.size());
        for (size_t g = 0; g < opt.devices.size(); ++g) cfg.device_ids[g] = opt.devices[g];
        cfg.max_box = opt.boxSizes.empty() ? 0 : *std::max_element(opt.boxSizes.begin(), opt.boxSizes.end());
    } else if (opt.devices.size() == 1)
        cfg.device = opt.devices[0];
    rpf_ctx *raw = nullptr;
    if (rpf_create(&raw, &cfg) != RPF_OK) return failed(error, "rpf_create", nullptr);
    std::unique_ptr<rpf_ctx, CtxDeleter> ctx(raw);


    // pack: plane[c][(y*W + x)*spp + s]; SampleData::data order (sd.h:62-94). The producer writes pbrt Floats into the
    // doubles (sd.h:62-94), so narrowing to float is exact for a fresh film. The container is samples[x][y] (x-major,
    // sample_film.cpp:16-22); one task per image row writes contiguous runs of every plane into page-locked memory.
    auto t0 = std::chrono::steady_clock::now();
    Pinned<float> planes = pinned<float>(static_cast<size_t>(nch) * S, error);
    Pinned<double> colours = pinned<double>(3 * S, error);
    if (!planes || !colours) return false;
    float *pl = planes.get();
    double *co = colours.get();
    const float *wv = opt.wide ? opt.wide->values.data() : nullptr;
    parallel_for(opt, H, [&](int64_t y) {
        for (int x = 0; x < W; ++x) {
            const SampleDataSet &px = film.samples[x][y];
            const size_t base = (static_cast<size_t>(y) * W + x) * spp;
            for (size_t s = 0; s < spp; ++s) {
                const auto &d = px[s].data;
                // position, colour, pLens stay where they are; the extra random parameters follow pLens, the extra
                // features follow p1
                for (int c = 0; c < 7; ++c) pl[c * S + base + s] = static_cast<float>(d[c]);
                for (int c = 7; c < RPF_NUM_CHANNELS; ++c) pl[(c + xr) * S + base + s] = static_cast<float>(d[c]);
                if (xs) {
                    const float *e = wv + (base + s) * xs;
                    for (int q = 0; q < xr; ++q) pl[(7 + q) * S + base + s] = e[q];
                    for (int q = 0; q < xf; ++q) pl[(RPF_NUM_CHANNELS + xr + q) * S + base + s] = e[xr + q];
                }
                for (int k = 0; k < 3; ++k) co[k * S + base + s] = d[2 + k];
            }
        }
    });
    const double packMs = ms_since(t0);

    if (opt.keepProbability < 1.0 && rpf_set_subsampling(ctx.get(), opt.keepProbability, opt.rngSeed, 0) != RPF_OK)
        return failed(error, "rpf_set_subsampling", ctx.get());
    t0 = std::chrono::steady_clock::now();
    std::vector<const float *> ptrs(nch);
    for (int c = 0; c < nch; ++c) ptrs[c] = pl + c * S;
    if (rpf_upload(ctx.get(), ptrs.data()) != RPF_OK) return failed(error, "rpf_upload", ctx.get());
    // colours re-entering the filter after an earlier pass are doubles (sd.h:205-207): keep every bit
    if (rpf_upload_colors_f64(ctx.get(), co) != RPF_OK) return failed(error, "rpf_upload_colors_f64", ctx.get());
    if (rpf_synchronize(ctx.get()) != RPF_OK) return failed(error, "rpf_synchronize", ctx.get());
    const double uploadMs = ms_since(t0);
    if ((opt.paper.w_epsilon != 0.0 || opt.paper.min_sigma != 0.0 || opt.paper.world_pos_sigmas != 0.0) &&
        rpf_set_paper_mode(ctx.get(), &opt.paper) != RPF_OK)
        return failed(error, "rpf_set_paper_mode", ctx.get());
    t0 = std::chrono::steady_clock::now();
    if (!opt.sigmaSeeds.empty()) {
        if (opt.sigmaSeeds.size() != opt.boxSizes.size()) {
            if (error) *error = "RPFB200Options: sigmaSeeds must hold one value per box size";
            return false;
        }
        if (rpf_filter_sigmas(ctx.get(), opt.boxSizes.data(), opt.sigmaSeeds.data(), static_cast<int>(opt.boxSizes.size())) != RPF_OK)
            return failed(error, "rpf_filter_sigmas", ctx.get());
    } else if (rpf_filter(ctx.get(), opt.boxSizes.data(), static_cast<int>(opt.boxSizes.size()), opt.sigmaSeed) != RPF_OK)
        return failed(error, "rpf_filter", ctx.get());
    if (opt.hdrClampSigmas > 0.0 && rpf_hdr_clamp(ctx.get(), opt.hdrClampSigmas, 0, 0) != RPF_OK)
        return failed(error, "rpf_hdr_clamp", ctx.get());
    if (rpf_synchronize(ctx.get()) != RPF_OK) return failed(error, "rpf_synchronize", ctx.get());
    const double filterWallMs = ms_since(t0);
    t0 = std::chrono::steady_clock::now();
    if (rpf_download_colors_f64(ctx.get(), co) != RPF_OK) return failed(error, "rpf_download_colors_f64", ctx.get());
    const double downloadMs = ms_since(t0);
    rpf_stats_t st{};
    if (rpf_get_stats(ctx.get(), &st) != RPF_OK) return failed(error, "rpf_get_stats", ctx.get());
    std::vector<int> nbh;
    if (opt.debugStats && report) {
        nbh.resize(static_cast<size_t>(W) * H);
        if (rpf_download_debug(ctx.get(), nbh.data(), nullptr, nullptr, nullptr) != RPF_OK) return failed(error, "rpf_download_debug", ctx.get());
    }
    if (fdesc && rgb) {  // rpf.cpp:776-804 on the device, samples still resident
        Pinned<float> rayWeight = pinned<float>(S, error);
        if (!rayWeight) return false;
        float *rw = rayWeight.get();
        parallel_for(opt, H, [&](int64_t y) {
            for (int x = 0; x < W; ++x)
                for (size_t s = 0; s < spp; ++s) rw[(static_cast<size_t>(y) * W + x) * spp + s] = film.samples[x][y][s].rayWeight;
        });
        rgb->resize(3 * static_cast<size_t>(fdesc->pixel_x1 - fdesc->pixel_x0) * (fdesc->pixel_y1 - fdesc->pixel_y0));
        if (rpf_reconstruct_film(ctx.get(), fdesc, rw, rgb->data()) != RPF_OK)
            return failed(error, "rpf_reconstruct_film", ctx.get());
    }

    // rpf.cpp:715,732: only the three colour slots change
    t0 = std::chrono::steady_clock::now();
    parallel_for(opt, W, [&](int64_t x) {
        for (int y = 0; y < H; ++y) {
            SampleDataSet &px = film.samples[x][y];
            const size_t base = (static_cast<size_t>(y) * W + x) * spp;
            for (size_t s = 0; s < spp; ++s)
                for (int k = 0; k < 3; ++k) px[s].data[2 + k] = co[k * S + base + s];
        }
    });
    const double unpackMs = ms_since(t0);
    if (report) {
        report->nanChannels = st.nan_channels;
        report->meanNeighbourhood = st.pixels_filtered ? static_cast<double>(st.nbh_sum) / st.pixels_filtered : 0.0;
        report->minNeighbourhood = st.nbh_min;
        report->maxNeighbourhood = st.nbh_max;
        report->filterMs = st.last_filter_ms;
        report->packMs = packMs; report->uploadMs = uploadMs; report->filterWallMs = filterWallMs;
        report->downloadMs = downloadMs; report->unpackMs = unpackMs;
        report->gpus = opt.devices.size() > 1 ? static_cast<int>(opt.devices.size()) : 1;
        report->neighbourhood.swap(nbh);
    }
    return true;
}

bool DumpSamplingFilmB200(const SamplingFilm &film, int sampleX0, int sampleY0, const std::string &path, std::string *error,
                          const RPFB200WideRecord *wide) {
    const int W = film.getWidth(), H = film.getHeight();
    const size_t spp = (W > 0 && H > 0 && !film.samples.empty() && !film.samples[0].empty()) ? film.samples[0][0].size() : 0;
    if (spp == 0) {
        if (error) *error = "empty SamplingFilm";
        return false;
    }
    const size_t S = static_cast<size_t>(W) * H * spp;
    const int xr = wide ? wide->nRandom : 0, xf = wide ? wide->nFeatures : 0, xs = xr + xf, nch = RPF_NUM_CHANNELS + xs;
    if (wide && wide->values.size() != S * xs) {
        if (error) *error = "RPFB200WideRecord does not match the SamplingFilm";
        return false;
    }
    std::vector<float> planes(static_cast<size_t>(nch) * S), rayWeight(S);
    for (int x = 0; x < W; ++x)
        for (int y = 0; y < H; ++y) {
            const SampleDataSet &px = film.samples[x][y];
            if (px.size() != spp) {
                if (error) *error = "SamplingFilm has a ragged sample count";
                return false;
            }
            const size_t base = (static_cast<size_t>(y) * W + x) * spp;
            for (size_t s = 0; s < spp; ++s) {
                for (int c = 0; c < 7; ++c) planes[c * S + base + s] = static_cast<float>(px[s].data[c]);
                for (int c = 7; c < RPF_NUM_CHANNELS; ++c) planes[(c + xr) * S + base + s] = static_cast<float>(px[s].data[c]);
                for (int q = 0; q < xr; ++q) planes[(7 + q) * S + base + s] = wide->values[(base + s) * xs + q];
                for (int q = 0; q < xf; ++q) planes[(RPF_NUM_CHANNELS + xr + q) * S + base + s] = wide->values[(base + s) * xs + xr + q];
                rayWeight[base + s] = px[s].rayWeight;
            }
        }
    std::vector<const float *> ptrs(nch);
    for (int c = 0; c < nch; ++c) ptrs[c] = planes.data() + c * S;
    rpf_buffer_info info{W, H, static_cast<int>(spp), sampleX0, sampleY0, 1, 2 + xr, 12 + xf};
    if (rpf_buffer_write(path.c_str(), &info, ptrs.data(), rayWeight.data()) != RPF_OK) {
        if (error) *error = "cannot write " + path;
        return false;
    }
    return true;
}

bool ApplyRPFFilterB200(SamplingFilm &film, const RPFB200Options &opt, RPFB200Report *report, std::string *error) {
    return run(film, opt, nullptr, nullptr, report, error);
}

bool ApplyRPFFilterAndFilmB200(SamplingFilm &film, const RPFB200Options &opt, const rpf_film_desc &fdesc, std::vector<float> *rgb,
                               RPFB200Report *report, std::string *error) {
    if (!rgb) {
        if (error) *error = "ApplyRPFFilterAndFilmB200: rgb is null";
        return false;
    }
    return run(film, opt, &fdesc, rgb, report, error);
}

}  // namespace pbrt
