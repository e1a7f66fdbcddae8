// oracle/ref_stubs/pbrt.h -- TEST INFRASTRUCTURE. Minimal stand-ins for the pbrt core types that
// the reference's src/custom/{sd.h,sample_film.h,sample_film.cpp} and rpf.cpp:302-733 touch, so those
// reference sources compile in place (from /root/reference) without pbrt's CMake build, glog or OpenEXR.
// Nothing here restates the filter; it is only the scaffolding the filter code expects
// (core/pbrt.h, core/geometry.h, core/spectrum.h, core/parallel.h, core/progressreporter.h, core/stats.h).
#ifndef RPF_REF_STUB_PBRT_H
#define RPF_REF_STUB_PBRT_H
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <functional>
#include <iostream>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace pbrt {
typedef float Float;  // CMakeLists.txt:12 PBRT_FLOAT_AS_DOUBLE is off by default

struct Point2f { Float x, y; Point2f(Float x = 0, Float y = 0) : x(x), y(y) {} };
struct Point3f { Float x, y, z; Point3f(Float x = 0, Float y = 0, Float z = 0) : x(x), y(y), z(z) {} };
struct Normal3f { Float x, y, z; Normal3f(Float x = 0, Float y = 0, Float z = 0) : x(x), y(y), z(z) {} };
struct Vector2i { int x, y; Vector2i(int x = 0, int y = 0) : x(x), y(y) {} };
struct Point2i {
    int x, y;
    Point2i(int x = 0, int y = 0) : x(x), y(y) {}
    bool operator==(const Point2i &o) const { return x == o.x && y == o.y; }
    bool operator!=(const Point2i &o) const { return !(*this == o); }
};

struct Spectrum {  // RGBSpectrum subset used by sd.h:66-70,104-112
    Float c[3];
    Spectrum(Float v = 0) { c[0] = c[1] = c[2] = v; }
    static Spectrum FromRGB(const Float rgb[3]) { Spectrum s; s.c[0] = rgb[0]; s.c[1] = rgb[1]; s.c[2] = rgb[2]; return s; }
    Float operator[](int i) const { return c[i]; }
    Float &operator[](int i) { return c[i]; }
};

struct Bounds2i;
struct Bounds2iIterator {  // core/geometry.h: x fastest, then y
    Point2i p; const Bounds2i *b;
    Bounds2iIterator(const Bounds2i &b, const Point2i &p) : p(p), b(&b) {}
    inline Bounds2iIterator &operator++();
    bool operator!=(const Bounds2iIterator &o) const { return p != o.p || b != o.b; }
    Point2i operator*() const { return p; }
};
struct Bounds2i {
    Point2i pMin, pMax;
    Bounds2i() {}
    Bounds2i(const Point2i &a, const Point2i &b) : pMin(std::min(a.x, b.x), std::min(a.y, b.y)), pMax(std::max(a.x, b.x), std::max(a.y, b.y)) {}
    Vector2i Diagonal() const { return Vector2i(pMax.x - pMin.x, pMax.y - pMin.y); }
};
inline Bounds2iIterator &Bounds2iIterator::operator++() {
    ++p.x;
    if (p.x == b->pMax.x) { p.x = b->pMin.x; ++p.y; }
    return *this;
}
inline Bounds2iIterator begin(const Bounds2i &b) { return Bounds2iIterator(b, b.pMin); }
inline Bounds2iIterator end(const Bounds2i &b) {
    Point2i pe(b.pMin.x, b.pMax.y);
    if (b.pMin.x >= b.pMax.x || b.pMin.y >= b.pMax.y) pe = b.pMin;
    return Bounds2iIterator(b, pe);
}
inline Bounds2i Intersect(const Bounds2i &a, const Bounds2i &b) {
    Bounds2i r;
    r.pMin = Point2i(std::max(a.pMin.x, b.pMin.x), std::max(a.pMin.y, b.pMin.y));
    r.pMax = Point2i(std::min(a.pMax.x, b.pMax.x), std::min(a.pMax.y, b.pMax.y));
    return r;
}

// core/parallel.h: ParallelFor2D hands out (x,y) tiles to a thread pool
extern int ref_stub_threads;
inline void ParallelFor2D(std::function<void(Point2i)> func, const Point2i &count) {
    const int total = count.x * count.y;
    std::atomic<int> next{0};
    auto work = [&]() { for (int t; (t = next.fetch_add(1)) < total;) func(Point2i(t % count.x, t / count.x)); };
    std::vector<std::thread> th;
    for (int i = 1; i < ref_stub_threads; ++i) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
}
struct ProgressReporter { ProgressReporter(int64_t, const std::string &) {} void Done() {} };
struct NullLog { template <typename T> NullLog &operator<<(const T &) { return *this; } };
#define LOG(severity) ::pbrt::NullLog()
#define ReportValue(stat, value) ((void)0)

struct Film { Bounds2i sampleBounds; std::string filename; Bounds2i GetSampleBounds() const { return sampleBounds; } };
struct Camera { Film *film; };
}  // namespace pbrt
#endif
