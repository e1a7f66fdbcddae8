// oracle/ref_film_stubs/filter.h -- TEST INFRASTRUCTURE. Stand-in for pbrt's core/{pbrt,geometry,filter}.h so that the
// reference's src/filters/{box,gaussian}.{h,cpp} and the extracted Film / FilmTile code (ref_film_shim.cpp) compile in
// place without pbrt's CMake build. Plumbing only: 2-D points / vectors / bounds with the operators that code uses
// (core/geometry.h) and the Filter base class interface (core/filter.h:47-58). No film arithmetic lives here.
#ifndef RPF_REF_FILM_STUB_FILTER_H
#define RPF_REF_FILM_STUB_FILTER_H
#include <alloca.h>
#include <algorithm>
#include <cmath>
#include <limits>
#include <memory>

namespace pbrt {
typedef float Float;  // PBRT_FLOAT_AS_DOUBLE is off by default
static constexpr Float Infinity = std::numeric_limits<Float>::infinity();
class ParamSet;

struct Vector2f { Float x, y; Vector2f(Float x = 0, Float y = 0) : x(x), y(y) {} };
struct Point2i;
struct Point2f {
    Float x, y;
    Point2f(Float x = 0, Float y = 0) : x(x), y(y) {}
    Point2f operator-(const Vector2f &v) const { return Point2f(x - v.x, y - v.y); }
    Point2f operator+(const Vector2f &v) const { return Point2f(x + v.x, y + v.y); }
};
struct Point2i {
    int x, y;
    Point2i(int x = 0, int y = 0) : x(x), y(y) {}
    explicit Point2i(const Point2f &p) : x((int)p.x), y((int)p.y) {}
    Point2i operator+(const Point2i &o) const { return Point2i(x + o.x, y + o.y); }
    bool operator==(const Point2i &o) const { return x == o.x && y == o.y; }
    bool operator!=(const Point2i &o) const { return !(*this == o); }
};
inline Point2f Ceil(const Point2f &p) { return Point2f(std::ceil(p.x), std::ceil(p.y)); }
inline Point2f Floor(const Point2f &p) { return Point2f(std::floor(p.x), std::floor(p.y)); }
inline Point2i Max(const Point2i &a, const Point2i &b) { return Point2i(std::max(a.x, b.x), std::max(a.y, b.y)); }
inline Point2i Min(const Point2i &a, const Point2i &b) { return Point2i(std::min(a.x, b.x), std::min(a.y, b.y)); }

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
    int Area() const { return (pMax.x - pMin.x) * (pMax.y - pMin.y); }
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
inline bool InsideExclusive(const Point2i &p, const Bounds2i &b) { return p.x >= b.pMin.x && p.x < b.pMax.x && p.y >= b.pMin.y && p.y < b.pMax.y; }

class Filter {  // core/filter.h:47-58
  public:
    virtual ~Filter() {}
    Filter(const Vector2f &radius) : radius(radius), invRadius(Vector2f(1 / radius.x, 1 / radius.y)) {}
    virtual Float Evaluate(const Point2f &p) const = 0;
    const Vector2f radius, invRadius;
};
}  // namespace pbrt
#endif
