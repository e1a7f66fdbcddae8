// oracle/ref_film_shim.cpp -- TEST INFRASTRUCTURE. C-ABI harness around the REFERENCE'S OWN Film reconstruction
// (SURVEY.md 8(f) N1), compiled where the sources lie under /root/reference (never copied into this repo):
//   src/core/spectrum.h:56-66        XYZToRGB / RGBToXYZ                      (REF_SPEC_FUNCS)
//   src/core/spectrum.h:455,462-465  RGBSpectrum::ToXYZ, RGBSpectrum::y       (REF_SPEC_MEMBERS)
//   src/core/film.h:51-55            FilmTilePixel                            (REF_TILEPIXEL)
//   src/core/film.h:107-187          class FilmTile incl. AddSample           (REF_FILMTILE)
//   src/core/film.cpp:69-77          filter table of the Film constructor     (REF_FILTERTABLE)
//   src/core/film.cpp:117-130        Film::MergeFilmTile                      (REF_MERGE)
//   src/core/film.cpp:169-211        Film::WriteImage                         (REF_WRITEIMAGE)
//   src/filters/box.{h,cpp}, gaussian.{h,cpp}  whole files, separate TUs, against ref_film_stubs/{filter,paramset}.h
// The Makefile streams exactly those line ranges into temp files. pbrt's geometry / logging / profiling types are
// replaced by the plumbing below (no filter or film arithmetic is restated here). pbrt::WriteImage (imageio.cpp,
// OpenEXR) is replaced by a capture of the RGB buffer it is handed.
#include "filter.h"  // ref_film_stubs/filter.h: Float, Point2f, Vector2f, Point2i, Bounds2i, Filter

#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "filters/box.h"
#include "filters/gaussian.h"

namespace pbrt {

#include REF_SPEC_FUNCS

struct Spectrum {  // RGBSpectrum: container + the operators AddSample uses; arithmetic members come from the extract
    Float c[3];
    Spectrum(Float v = 0.f) { c[0] = c[1] = c[2] = v; }
    static Spectrum FromRGB(const Float rgb[3]) { Spectrum s; s.c[0] = rgb[0]; s.c[1] = rgb[1]; s.c[2] = rgb[2]; return s; }
    Spectrum &operator*=(Float a) { for (int i = 0; i < 3; ++i) c[i] *= a; return *this; }       // spectrum.h:122-126
    Spectrum operator*(Float a) const { Spectrum r = *this; for (int i = 0; i < 3; ++i) r.c[i] *= a; return r; }  // :116-121
    Spectrum &operator+=(const Spectrum &s) { for (int i = 0; i < 3; ++i) c[i] += s.c[i]; return *this; }  // :71-75
#include REF_SPEC_MEMBERS
};

struct ProfilePhase { template <class T> ProfilePhase(T) {} };
enum class Prof { AddFilmSample, MergeFilmTile };
struct NullLog { template <typename T> NullLog &operator<<(const T &) { return *this; } };
inline NullLog &operator<<(NullLog &l, const Bounds2i &) { return l; }
#define LOG(severity) ::pbrt::NullLog()
#define VLOG(level) ::pbrt::NullLog()
#define CHECK(cond) ((void)0)
#define ALLOCA(TYPE, COUNT) (TYPE *) alloca((COUNT) * sizeof(TYPE))

#include REF_TILEPIXEL
class Film;
#include REF_FILMTILE

static std::vector<Float> g_written;
static void WriteImage(const std::string &, const Float *rgb, const Bounds2i &b, const Point2i &) {
    g_written.assign(rgb, rgb + 3 * static_cast<size_t>(b.Area()));
}

// Declaration mirror of film.h:58-105 restricted to what the extracted definitions touch.
class Film {
  public:
    Film(const Point2i &resolution, const Bounds2i &cropped, std::unique_ptr<Filter> filt, Float scale, Float maxLum)
        : fullResolution(resolution), filter(std::move(filt)), croppedPixelBounds(cropped), scale(scale), maxSampleLuminance(maxLum) {
        pixels = std::unique_ptr<Pixel[]>(new Pixel[croppedPixelBounds.Area()]);
#include REF_FILTERTABLE
    }
    std::unique_ptr<FilmTile> GetFilmTile(const Bounds2i &tilePixelBounds) {  // film.cpp:103-105 with the bounds given
        return std::unique_ptr<FilmTile>(new FilmTile(tilePixelBounds, filter->radius, filterTable, filterTableWidth, maxSampleLuminance));
    }
    void MergeFilmTile(std::unique_ptr<FilmTile> tile);
    void WriteImage(Float splatScale = 1);
    const Point2i fullResolution;
    std::unique_ptr<Filter> filter;
    const std::string filename;
    Bounds2i croppedPixelBounds;
    struct Pixel {  // film.h:83-89 (AtomicFloat splats are never written on this path: plain floats at 0)
        Pixel() { xyz[0] = xyz[1] = xyz[2] = filterWeightSum = 0; splatXYZ[0] = splatXYZ[1] = splatXYZ[2] = 0; }
        Float xyz[3];
        Float filterWeightSum;
        Float splatXYZ[3];
        Float pad;
    };
    std::unique_ptr<Pixel[]> pixels;
    static constexpr int filterTableWidth = 16;
    Float filterTable[filterTableWidth * filterTableWidth];
    std::mutex mutex;
    const Float scale;
    const Float maxSampleLuminance;
    Pixel &GetPixel(const Point2i &p) {  // film.h:97-104
        int width = croppedPixelBounds.pMax.x - croppedPixelBounds.pMin.x;
        int offset = (p.x - croppedPixelBounds.pMin.x) + (p.y - croppedPixelBounds.pMin.y) * width;
        return pixels[offset];
    }
};

#include REF_MERGE
#include REF_WRITEIMAGE

}  // namespace pbrt

using namespace pbrt;

extern "C" {

// The tail of RPFIntegrator::Render (rpf.cpp:776-804) on a frame given as planes: one FilmTile over the cropped pixel
// bounds, AddSample for every sample in the reference's order (x outer, y inner, sample order), MergeFilmTile, WriteImage.
// W x H x P samples whose sample bounds start at (sx0, sy0); pfilm [2][S] float, colours [3][S] double (cast to Float
// as sd.h:102-110 does), ray_weight [S] or null (= 1). filter_kind 0 box, 1 gaussian(alpha). Pixel bounds [px0,px1) x
// [py0,py1). rgb_out [(py1-py0)*(px1-px0)][3]; table_out [256] (may be null).
int ref_film_reconstruct(int W, int H, int P, int sx0, int sy0, const float *pfilm, const double *colors, const float *ray_weight,
                         int filter_kind, float rx, float ry, float alpha, int px0, int py0, int px1, int py1, float scale,
                         float max_lum, float *rgb_out, float *table_out) {
    std::unique_ptr<Filter> filt;
    if (filter_kind == 0) filt.reset(new BoxFilter(Vector2f(rx, ry)));
    else filt.reset(new GaussianFilter(Vector2f(rx, ry), alpha));
    const Bounds2i cropped(Point2i(px0, py0), Point2i(px1, py1));
    Film film(Point2i(px1, py1), cropped, std::move(filt), scale, max_lum);
    if (table_out) std::memcpy(table_out, film.filterTable, sizeof film.filterTable);
    std::unique_ptr<FilmTile> tile = film.GetFilmTile(cropped);
    const size_t S = static_cast<size_t>(W) * H * P;
    for (int x = 0; x < W; ++x)
        for (int y = 0; y < H; ++y)
            for (int s = 0; s < P; ++s) {
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                Float rgb[3] = {static_cast<Float>(colors[i]), static_cast<Float>(colors[S + i]), static_cast<Float>(colors[2 * S + i])};
                tile->AddSample(Point2f(pfilm[i], pfilm[S + i]), Spectrum::FromRGB(rgb), ray_weight ? ray_weight[i] : 1.f);
            }
    (void)sx0; (void)sy0;
    film.MergeFilmTile(std::move(tile));
    film.WriteImage();
    std::memcpy(rgb_out, g_written.data(), g_written.size() * sizeof(Float));
    return 0;
}
}
