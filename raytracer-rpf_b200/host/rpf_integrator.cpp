// rpf_integrator.cpp -- the plugin surface of `Integrator "rpf"` with the B200 filter behind it.
//
// Drop-in replacement for the reference's src/custom/rpf.cpp: it defines the external symbols that file defines
//     pbrt::RPFIntegrator::{RPFIntegrator, Preprocess, Render, Li, ...}     (class declared in src/custom/rpf.h:28-96)
//     pbrt::CreateRPFIntegrator                                            (rpf.h:98-102, rpf.cpp:941-966)
//     pbrt::visualizeSF                                                    (rpf.cpp:37-101)
// against the UNCHANGED reference header, so the unmodified reference api.cpp (#include at api.cpp:63, factory call at
// api.cpp:1696-1697) links against this object. Put it on the link line ahead of libpbrt.a: every symbol of the archive
// member rpf.cpp.o is then already defined and the member is never pulled (host/Makefile does exactly that).
//
// What stays on the CPU, as in the reference: the feature producer (FillSampleFilm + Li: camera rays, path tracing,
// per-sample features, rpf.cpp:220-299,808-939) and pbrt's Film (rpf.cpp:776-804). What moves to the GPU: the loop
// `for (box_size : box_sizes) ApplyRPFFilter(...)` of rpf.cpp:767-775, through host/rpf_b200_host.cpp and the C ABI of
// include/rpf_b200.h. ApplyRPFFilter / ComputeCFWeights / FillMeanAndStddev themselves are not defined here: nothing
// outside rpf.cpp ever referenced them (private members, rpf.h:56-95).
//
// Scene-file parameters: the reference's four (maxdepth, pixelbounds, rrthreshold, lightsamplestrategy; rpf.cpp:946-963)
// plus optional ones whose defaults are the reference's literals, all looked up so that ParamSet::ReportUnused
// (api.cpp:1713) stays quiet:
//     "integer boxsizes" [7]          rpf.cpp:767 ({55 35 17 7} is the schedule in that line's comment)
//     "float sigmaseed" 0.002         rpf.cpp:533
//     "integer ngpus" 1               row bands over the first n CUDA devices (rpf_config.n_gpus)
//     "integer gpudevices" [..]       explicit device ordinals instead
//     "string rpfmode" "defined"      "defined" = REF_DEFINED patch list; "verbatim" = the reference's
//                                     formulas including the division 0/0 (NaN where the reference is undefined);
//                                     "paper" = REF_DEFINED plus Sen & Darabi's epsilon / low-variance guard / x30 world
//                                     positions / sigma schedule / spike removal (rpf_set_paper_mode)
//     "string rpfdump" ""             write the feature buffer (all 19 channels + ray weights, lossless) to this file
//                                     before filtering -- what visualizeSF's six debug EXRs were for
//     "bool rpfdumponly" false        stop after the dump: no filtering, the unfiltered samples go to the Film
//                                     (for machines without a GPU)
//     "string rpfrecord" "reference"  "reference" = the 19 channels of SampleData (sd.h:46-94); "wide" = north_star's wider
//                                     feature buffer: the sample's time and the first light sample
//                                     (u, v) as three more random parameters, the first-hit texture colour (the BSDF's
//                                     reflectance) as three more features -- 25 channels, kept beside the SamplingFilm
//                                     because rpf.h leaves no room in it. The sampler requests stay in the reference's
//                                     order: the first 19 channels of a wide frame are the reference frame bit for bit.
#include "custom/rpf.h"

#include <chrono>
#include <map>
#include <mutex>

#include "bssrdf.h"
#include "camera.h"
#include "film.h"
#include "interaction.h"
#include "parallel.h"
#include "paramset.h"
#include "progressreporter.h"
#include "sampler.h"
#include "scene.h"
#include "stats.h"

#include "rpf_b200_host.h"

namespace pbrt {

// the reference's statistics that still have a meaning here, under the reference's names (rpf.cpp:20-33)
STAT_COUNTER("Integrator/Camera rays traced", nCameraRays);
STAT_INT_DISTRIBUTION("Integrator/Samples Per Pixel", samplesPerPixel);
STAT_INT_DISTRIBUTION("RPF/Neighborhood Size", neighborhoodSize);
STAT_FLOAT_DISTRIBUTION("RPF/Output color", outputColorDistribution);
STAT_PERCENT("Integrator/Zero-radiance paths", zeroRadiancePaths, totalPaths);
STAT_INT_DISTRIBUTION("Integrator/Path length", pathLength);
STAT_COUNTER("RPF/NaN colour channels", nanColourChannels);

namespace {

// rpf.h is unchanged, so the integrator object has no room for new members: what the extra scene parameters say is kept
// beside it, keyed by the object.
struct PluginOptions {
    RPFB200Options filter;
    std::string dumpPath;
    bool dumpOnly = false;
    bool wideRecord = false;
};

// The six values per sample of the wide record: time, light sample u, v | texture colour r, g, b.
constexpr int kWideRandom = 3, kWideFeatures = 3, kWide = kWideRandom + kWideFeatures;
std::mutex g_optionsMutex;
std::map<const RPFIntegrator *, PluginOptions> g_options;

PluginOptions optionsOf(const RPFIntegrator *self) {
    std::lock_guard<std::mutex> lock(g_optionsMutex);
    auto it = g_options.find(self);
    return it == g_options.end() ? PluginOptions() : it->second;
}

// FillSampleFilm's signature is the reference's (rpf.h:76-79): where the wide record of the frame being sampled goes is
// handed over here by Render (one render at a time, as in pbrt).
RPFB200WideRecord *g_fillWide = nullptr;

std::string withoutExtension(const std::string &name) {
    const std::string::size_type dot = name.rfind('.');
    return dot == std::string::npos ? name : name.substr(0, dot);
}

}  // namespace

RPFIntegrator::RPFIntegrator(int maxDepth, std::shared_ptr<const Camera> camera, std::shared_ptr<Sampler> sampler,
                             const Bounds2i &pixelBounds, Float rrThreshold, const std::string &lightSampleStrategy)
    : camera(camera), maxDepth(maxDepth), rrThreshold(rrThreshold), lightSampleStrategy(lightSampleStrategy),
      sampler(sampler), pixelBounds(pixelBounds) {}

void RPFIntegrator::Preprocess(const Scene &scene, Sampler &) {
    lightDistribution = CreateLightSampleDistribution(lightSampleStrategy, scene);
}

// The feature buffer as a file. The reference's version averages six features per pixel into debug
// EXRs; this one keeps every sample of every channel: <film name without extension>.rpfb unless the name already ends
// in .rpfb.
void visualizeSF(const SamplingFilm &samplingFilm, const std::string &filename) {
    const std::string ext = ".rpfb";
    const bool named = filename.size() > ext.size() && filename.compare(filename.size() - ext.size(), ext.size(), ext) == 0;
    const std::string path = named ? filename : withoutExtension(filename) + ext;
    std::string err;
    if (!DumpSamplingFilmB200(samplingFilm, samplingFilm.pixelBounds.pMin.x, samplingFilm.pixelBounds.pMin.y, path, &err))
        Error("RPF feature buffer dump to \"%s\" failed: %s", path.c_str(), err.c_str());
}

// Path tracing with feature capture: what the reference's Li does (rpf.cpp:808-939, itself pbrt's PathIntegrator::Li
// minus the light distribution for direct lighting, rpf.cpp:871). The order of sampler requests is the reference's, so a
// given (pixel, sample index) yields the same path and the same features.
namespace {
// Direct lighting from one light chosen uniformly: core/integrator.cpp:85-106 (UniformSampleOneLight without a light
// distribution, as rpf.cpp:871 calls it) with the same sampler requests in the same order and the same arithmetic, so the
// estimate is the reference's bit for bit -- restated here only to get at the light sample it draws.
Spectrum sampleOneLightRecording(const Interaction &it, const Scene &scene, MemoryArena &arena, Sampler &sampler, Point2f *uLightOut) {
    ProfilePhase p(Prof::DirectLighting);
    const int nLights = int(scene.lights.size());
    if (nLights == 0) return Spectrum(0.f);
    const int lightNum = std::min((int)(sampler.Get1D() * nLights), nLights - 1);
    const Float lightPdf = Float(1) / nLights;
    const std::shared_ptr<Light> &light = scene.lights[lightNum];
    const Point2f uLight = sampler.Get2D();
    const Point2f uScattering = sampler.Get2D();
    if (uLightOut) *uLightOut = uLight;
    return EstimateDirect(it, uScattering, *light, uLight, scene, sampler, arena, false) / lightPdf;
}
}  // namespace

// `wide` (optional, 6 floats): [1], [2] = the light sample of the first vertex that samples a light, [3..5] = the
// reflectance of the first vertex's BSDF ("texture colour"); [0] (time) is the caller's.
static void traceWithFeatures(const RayDifferential &cameraRay, const Scene &scene, Sampler &sampler, MemoryArena &arena,
                              SampleData &sample, int maxDepth, Float rrThreshold, const LightDistribution *lightDistribution, float *wide) {
    ProfilePhase phase(Prof::SamplerIntegratorLi);
    bool haveLightSample = false, haveTexture = false;
    Spectrum radiance(0.f), throughput(1.f);
    RayDifferential ray(cameraRay);
    bool lastWasSpecular = false;
    Float etaScale = 1;
    int bounces = 0;
    for (;; ++bounces) {
        SurfaceInteraction isect;
        const bool hit = scene.Intersect(ray, &isect);

        // emission is only picked up where light sampling could not have: camera rays and specular chains
        if (bounces == 0 || lastWasSpecular) {
            if (hit)
                radiance += throughput * isect.Le(-ray.d);
            else
                for (const auto &light : scene.infiniteLights) radiance += throughput * light->Le(ray);
        }
        if (!hit || bounces >= maxDepth) break;

        // the features RPF clusters on: geometric normal and world position of the first two vertices (rpf.cpp:851-857).
        // A path that ends at its first vertex leaves n1 / p1 at zero.
        if (bounces == 0) {
            sample.setN0(isect.n);
            sample.setP0(isect.p);
        } else if (bounces == 1) {
            sample.setN1(isect.n);
            sample.setP1(isect.p);
        }

        isect.ComputeScatteringFunctions(ray, arena, true);
        if (!isect.bsdf) {  // medium boundary: continue straight on, the vertex does not count (and is recorded again)
            ray = isect.SpawnRay(ray.d);
            --bounces;
            continue;
        }
        if (wide && !haveTexture) {
            // hemispherical-directional reflectance of the first surface: the texture's colour for a matte surface
            // (LambertianReflection::rho returns R, reflection.h:394), an estimate over four fixed directions otherwise.
            // No sampler request: the sample stream stays the reference's.
            static const Point2f kRhoSamples[4] = {Point2f(0.125f, 0.375f), Point2f(0.625f, 0.875f), Point2f(0.375f, 0.625f), Point2f(0.875f, 0.125f)};
            Float rgb[3];
            isect.bsdf->rho(-ray.d, 4, kRhoSamples, BSDF_ALL).ToRGB(rgb);
            wide[3] = rgb[0]; wide[4] = rgb[1]; wide[5] = rgb[2];
            haveTexture = true;
        }

        if (isect.bsdf->NumComponents(BxDFType(BSDF_ALL & ~BSDF_SPECULAR)) > 0) {
            ++totalPaths;
            Point2f uLight;
            const Spectrum direct = throughput * sampleOneLightRecording(isect, scene, arena, sampler, &uLight);
            if (wide && !haveLightSample) {
                wide[1] = uLight.x;
                wide[2] = uLight.y;
                haveLightSample = true;
            }
            if (direct.IsBlack()) ++zeroRadiancePaths;
            CHECK_GE(direct.y(), 0.f);
            radiance += direct;
        }

        const Vector3f wo = -ray.d;
        Vector3f wi;
        Float pdf;
        BxDFType sampled;
        const Spectrum f = isect.bsdf->Sample_f(wo, &wi, sampler.Get2D(), &pdf, BSDF_ALL, &sampled);
        if (f.IsBlack() || pdf == 0.f) break;
        throughput *= f * AbsDot(wi, isect.shading.n) / pdf;
        CHECK_GE(throughput.y(), 0.f);
        lastWasSpecular = (sampled & BSDF_SPECULAR) != 0;
        if ((sampled & BSDF_SPECULAR) && (sampled & BSDF_TRANSMISSION)) {
            const Float eta = isect.bsdf->eta;
            etaScale *= (Dot(wo, isect.n) > 0) ? (eta * eta) : 1 / (eta * eta);
        }
        ray = isect.SpawnRay(wi);

        if (isect.bssrdf && (sampled & BSDF_TRANSMISSION)) {  // subsurface scattering: leave the surface somewhere else
            SurfaceInteraction exitPoint;
            const Spectrum S = isect.bssrdf->Sample_S(scene, sampler.Get1D(), sampler.Get2D(), arena, &exitPoint, &pdf);
            if (S.IsBlack() || pdf == 0) break;
            throughput *= S / pdf;
            radiance += throughput * UniformSampleOneLight(exitPoint, scene, arena, sampler, false,
                                                           lightDistribution ? lightDistribution->Lookup(exitPoint.p) : nullptr);
            const Spectrum fExit = exitPoint.bsdf->Sample_f(exitPoint.wo, &wi, sampler.Get2D(), &pdf, BSDF_ALL, &sampled);
            if (fExit.IsBlack() || pdf == 0) break;
            throughput *= fExit * AbsDot(wi, exitPoint.shading.n) / pdf;
            lastWasSpecular = (sampled & BSDF_SPECULAR) != 0;
            ray = exitPoint.SpawnRay(wi);
        }

        // Russian roulette on the throughput with the refraction scaling factored out
        const Spectrum rr = throughput * etaScale;
        if (rr.MaxComponentValue() < rrThreshold && bounces > 3) {
            const Float q = std::max((Float).05, 1 - rr.MaxComponentValue());
            if (sampler.Get1D() < q) break;
            throughput /= 1 - q;
        }
    }
    ReportValue(pathLength, bounces);
    sample.setL(radiance);
}

void RPFIntegrator::Li(const RayDifferential &cameraRay, const Scene &scene, Sampler &sampler, MemoryArena &arena,
                       SampleData &sample, int /*depth*/) const {
    traceWithFeatures(cameraRay, scene, sampler, arena, sample, maxDepth, rrThreshold, lightDistribution.get(), nullptr);
}

// rpf.cpp:220-299: one task per 16x16 tile on pbrt's pool, a sampler clone per tile seeded with the tile number, every
// sample of every pixel through Li, the tile merged into the SamplingFilm under its mutex.
void RPFIntegrator::FillSampleFilm(SamplingFilm &samplingFilm, const Scene &scene, const int tileSize) {
    const Bounds2i sampleBounds = camera->film->GetSampleBounds();
    RPFB200WideRecord *wideRecord = g_fillWide;
    const int fullW = sampleBounds.pMax.x - sampleBounds.pMin.x;
    const int64_t spp = sampler->samplesPerPixel;
    const Vector2i extent = pixelBounds.Diagonal();  // as the reference: tiles are counted over pixelBounds (rpf.cpp:226)
    const Point2i nTiles((extent.x + tileSize - 1) / tileSize, (extent.y + tileSize - 1) / tileSize);
    ProgressReporter progress(nTiles.x * nTiles.y, "Sampling");
    ParallelFor2D([&](Point2i tile) {
        MemoryArena arena;
        std::unique_ptr<Sampler> tileSampler = sampler->Clone(tile.y * nTiles.x + tile.x);
        const int x0 = sampleBounds.pMin.x + tile.x * tileSize, x1 = std::min(x0 + tileSize, sampleBounds.pMax.x);
        const int y0 = sampleBounds.pMin.y + tile.y * tileSize, y1 = std::min(y0 + tileSize, sampleBounds.pMax.y);
        const Bounds2i tileBounds(Point2i(x0, y0), Point2i(x1, y1));
        std::unique_ptr<SamplingTile> samplingTile = samplingFilm.GetSamplingTile(tileBounds);
        for (Point2i pixel : tileBounds) {
            {
                ProfilePhase pp(Prof::StartPixel);
                tileSampler->StartPixel(pixel);
            }
            if (!InsideExclusive(pixel, pixelBounds)) continue;
            do {
                const CameraSample cameraSample = tileSampler->GetCameraSample(pixel);
                RayDifferential ray;
                const Float rayWeight = camera->GenerateRayDifferential(cameraSample, &ray);
                ray.ScaleDifferentials(1 / std::sqrt((Float)tileSampler->samplesPerPixel));
                ++nCameraRays;
                SampleData sample(cameraSample.pFilm, cameraSample.pLens, 0.f, rayWeight);
                if (!wideRecord) {
                    if (rayWeight > 0) Li(ray, scene, *tileSampler, arena, sample);
                } else {
                    // the values the 19-double record has no room for, at this sample's place in the frame
                    const int64_t at = ((int64_t)(pixel.y - sampleBounds.pMin.y) * fullW + (pixel.x - sampleBounds.pMin.x)) * spp +
                                       tileSampler->CurrentSampleNumber();
                    float *w = wideRecord->values.data() + at * kWide;
                    w[0] = cameraSample.time;
                    if (rayWeight > 0)
                        traceWithFeatures(ray, scene, *tileSampler, arena, sample, maxDepth, rrThreshold, lightDistribution.get(), w);
                }
                samplingTile->addSample(pixel, sample);
                arena.Reset();
            } while (tileSampler->StartNextSample());
        }
        samplingFilm.MergeSamplingTile(std::move(samplingTile));
        progress.Update();
    }, nTiles);
    progress.Done();
}

// rpf.cpp:737-805 with the filter loop on the GPU.
void RPFIntegrator::Render(const Scene &scene) {
    const PluginOptions opt = optionsOf(this);
    Film *film = camera->film;
    const Bounds2i sampleBounds = film->GetSampleBounds();
    const Vector2i sampleExtent = sampleBounds.Diagonal();
    const int tileSize = 16;  // rpf.cpp:742

    SamplingFilm samplingFilm(sampleBounds);
    RPFB200WideRecord wide;
    if (opt.wideRecord) {
        wide.nRandom = kWideRandom;
        wide.nFeatures = kWideFeatures;
        wide.values.assign(static_cast<size_t>(sampleExtent.x) * sampleExtent.y * sampler->samplesPerPixel * kWide, 0.f);
        g_fillWide = &wide;
    }
    FillSampleFilm(samplingFilm, scene, tileSize);
    g_fillWide = nullptr;
    for (const auto &column : samplingFilm.samples)
        for (const auto &pixel : column) ReportValue(samplesPerPixel, static_cast<int64_t>(pixel.size()));

    if (!opt.dumpPath.empty()) {
        if (!opt.wideRecord)
            visualizeSF(samplingFilm, opt.dumpPath);
        else {
            std::string err;
            if (!DumpSamplingFilmB200(samplingFilm, sampleBounds.pMin.x, sampleBounds.pMin.y, opt.dumpPath, &err, &wide))
                Error("RPF feature buffer dump to \"%s\" failed: %s", opt.dumpPath.c_str(), err.c_str());
        }
    }

    if (opt.dumpOnly) {
        Warning("RPF: \"rpfdumponly\" is set -- the samples go to the Film unfiltered");
    } else {
        RPFB200Options fo = opt.filter;
        fo.debugStats = true;
        if (opt.wideRecord) fo.wide = &wide;
        fo.parallelFor = [](const std::function<void(int64_t)> &body, int64_t count) { ParallelFor(body, count); };
        RPFB200Report rep;
        std::string err;
        if (!ApplyRPFFilterB200(samplingFilm, fo, &rep, &err)) {
            // no CPU filter exists in this build: say so and stop, as the reference stops on its own fatal condition
            // (rpf.cpp:702-705)
            Error("RPF B200 filter failed: %s", err.c_str());
            fflush(stdout);
            fflush(stderr);
            ParallelCleanup();  // pbrt's workers must be joined first or exit() hangs in static destruction
            exit(1);
        }
        for (int n : rep.neighbourhood) ReportValue(neighborhoodSize, static_cast<int64_t>(n));
        nanColourChannels += rep.nanChannels;
        if (rep.nanChannels)
            Warning("RPF: %lld filtered colour channels are NaN (the reference exits here, rpf.cpp:702-705)", static_cast<long long>(rep.nanChannels));
        if (!PbrtOptions.quiet) {
            printf("RPF on %d B200 GPU(s): %zu pass(es), mean neighbourhood %.1f [%d - %d]; pack %.1f ms, upload %.1f ms, filter %.1f ms "
                   "(device %.1f ms), download %.1f ms, unpack %.1f ms\n",
                   rep.gpus, fo.boxSizes.size(), rep.meanNeighbourhood, rep.minNeighbourhood, rep.maxNeighbourhood, rep.packMs, rep.uploadMs,
                   rep.filterWallMs, rep.filterMs, rep.downloadMs, rep.unpackMs);
            fflush(stdout);
        }
    }

    // rpf.cpp:776-804: every sample through one FilmTile in the reference's order (x outer, y inner, sample order), so the
    // Float accumulation and therefore the image are the reference's for the same samples
    std::unique_ptr<FilmTile> filmTile = film->GetFilmTile(sampleBounds);
    for (int x = 0; x < sampleExtent.x; ++x)
        for (int y = 0; y < sampleExtent.y; ++y)
            for (const SampleData &s : samplingFilm.samples[x][y]) {
                filmTile->AddSample(s.getPFilm(), s.getL(), s.rayWeight);
                for (size_t k = 0; k < SD_N_COLOR; ++k) ReportValue(outputColorDistribution, s.getColorI(k));
            }
    film->MergeFilmTile(std::move(filmTile));
    film->WriteImage();
    {
        std::lock_guard<std::mutex> lock(g_optionsMutex);
        g_options.erase(this);
    }
}

RPFIntegrator *CreateRPFIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler, std::shared_ptr<const Camera> camera) {
    // the reference's four parameters (rpf.cpp:946-963)
    const int maxDepth = params.FindOneInt("maxdepth", 5);
    int np;
    const int *pb = params.FindInt("pixelbounds", &np);
    Bounds2i pixelBounds = camera->film->GetSampleBounds();
    if (pb) {
        if (np != 4)
            Error("Expected four values for \"pixelbounds\" parameter. Got %d.", np);
        else {
            pixelBounds = Intersect(pixelBounds, Bounds2i{{pb[0], pb[2]}, {pb[1], pb[3]}});
            if (pixelBounds.Area() == 0) Error("Degenerate \"pixelbounds\" specified.");
        }
    }
    const Float rrThreshold = params.FindOneFloat("rrthreshold", 1.);
    const std::string lightStrategy = params.FindOneString("lightsamplestrategy", "spatial");

    // the filter's own parameters; every default is the reference's literal
    PluginOptions opt;
    int nb = 0;
    if (const int *boxes = params.FindInt("boxsizes", &nb)) {
        opt.filter.boxSizes.assign(boxes, boxes + nb);
        for (int b : opt.filter.boxSizes)
            if (b < 1 || b % 2 == 0) {
                Error("\"boxsizes\" must be odd and positive (rpf.cpp:561); got %d. Using [7].", b);
                opt.filter.boxSizes = {7};
                break;
            }
    }
    opt.filter.sigmaSeed = params.FindOneFloat("sigmaseed", 0.002f);
    if (!(opt.filter.sigmaSeed > 0)) {
        Error("\"sigmaseed\" must be positive; using 0.002.");
        opt.filter.sigmaSeed = 0.002;
    }
    const int nGpus = params.FindOneInt("ngpus", 1);
    int nd = 0;
    if (const int *devs = params.FindInt("gpudevices", &nd))
        opt.filter.devices.assign(devs, devs + nd);
    else if (nGpus > 1)
        for (int g = 0; g < nGpus; ++g) opt.filter.devices.push_back(g);
    if (opt.filter.devices.size() > RPF_MAX_GPUS) {
        Error("\"ngpus\" / \"gpudevices\": at most %d devices; using the first %d.", RPF_MAX_GPUS, RPF_MAX_GPUS);
        opt.filter.devices.resize(RPF_MAX_GPUS);
    }
    const std::string mode = params.FindOneString("rpfmode", "defined");
    if (mode == "verbatim")
        opt.filter.flags = RPF_FLAGS_VERBATIM;
    else if (mode == "paper") {
        opt.filter.paper = rpf_paper_opts{1e-7, 0.1, 30.0};
        opt.filter.hdrClampSigmas = 1.0;
    } else if (mode != "defined")
        Error("\"rpfmode\" must be \"defined\", \"verbatim\" or \"paper\"; got \"%s\". Using \"defined\".", mode.c_str());
    const std::string record = params.FindOneString("rpfrecord", "reference");
    if (record == "wide")
        opt.wideRecord = true;
    else if (record != "reference")
        Error("\"rpfrecord\" must be \"reference\" or \"wide\"; got \"%s\". Using \"reference\".", record.c_str());
    opt.dumpPath = params.FindOneString("rpfdump", "");
    opt.dumpOnly = params.FindOneBool("rpfdumponly", false);

    RPFIntegrator *integrator = new RPFIntegrator(maxDepth, camera, sampler, pixelBounds, rrThreshold, lightStrategy);
    {
        std::lock_guard<std::mutex> lock(g_optionsMutex);
        g_options[integrator] = opt;
    }
    return integrator;
}

}  // namespace pbrt
