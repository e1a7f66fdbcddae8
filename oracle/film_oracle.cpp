// oracle/film_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// CPU restatement of the Film reconstruction that follows the RPF filter:
//   src/custom/rpf.cpp:776-804      tail of RPFIntegrator::Render: one FilmTile, AddSample per sample (x outer, y inner)
//   src/core/film.h:121-161         FilmTile::AddSample
//   src/core/film.cpp:69-77         filter table; :117-130 MergeFilmTile; :169-205 WriteImage arithmetic
//   src/core/spectrum.h:56-66,455,462-465   XYZToRGB, RGBToXYZ, RGBSpectrum::ToXYZ / y
//   src/filters/box.cpp:41, gaussian.h:52-66, gaussian.cpp:41-43
// Pinned against the reference's own lines compiled in place (oracle/_ref/libref_film.so, ref_film_shim.cpp) and the
// fixtures generated from it (tests/golden/film_case*.npz). Float arithmetic, -ffp-contract=off (oracle/Makefile).
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <vector>

typedef float Float;

extern "C" {

// film.cpp:69-77 with the box / gaussian Evaluate
void rpf_oracle_film_table(int kind, Float rx, Float ry, Float alpha, Float *table) {
    const Float expX = std::exp(-alpha * rx * rx), expY = std::exp(-alpha * ry * ry);  // gaussian.h:56-57
    int offset = 0;
    for (int y = 0; y < 16; ++y)
        for (int x = 0; x < 16; ++x, ++offset) {
            const Float px = (x + 0.5f) * rx / 16, py = (y + 0.5f) * ry / 16;
            if (kind == 0) { table[offset] = 1.f; continue; }  // box.cpp:41
            const Float gx = std::max((Float)0, Float(std::exp(-alpha * px * px) - expX));  // gaussian.h:63-65
            const Float gy = std::max((Float)0, Float(std::exp(-alpha * py * py) - expY));
            table[offset] = gx * gy;  // gaussian.cpp:42
        }
}

// Whole tail: samples (pfilm [2][S] float, colours [3][S] double, ray_weight [S] float or null) -> rgb [(py1-py0)*(px1-px0)][3]
void rpf_oracle_film(int W, int H, int P, const float *pfilm, const double *colors, const float *ray_weight, const Float *table,
                     Float rx, Float ry, int px0, int py0, int px1, int py1, Float scale, Float max_lum, Float *rgb_out) {
    const int pw = px1 - px0, ph = py1 - py0;
    const size_t S = static_cast<size_t>(W) * H * P;
    std::vector<Float> sum(3 * static_cast<size_t>(pw) * ph, 0.f), fws(static_cast<size_t>(pw) * ph, 0.f);
    const Float irx = 1 / rx, iry = 1 / ry;  // filter.h:52
    for (int x = 0; x < W; ++x)          // rpf.cpp:783-786 order
        for (int y = 0; y < H; ++y)
            for (int s = 0; s < P; ++s) {
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                Float L[3] = {static_cast<Float>(colors[i]), static_cast<Float>(colors[S + i]), static_cast<Float>(colors[2 * S + i])};
                const Float sw = ray_weight ? ray_weight[i] : 1.f;
                const Float lum = 0.212671f * L[0] + 0.715160f * L[1] + 0.072169f * L[2];  // spectrum.h:462-465
                if (lum > max_lum) { const Float a = max_lum / lum; for (int k = 0; k < 3; ++k) L[k] *= a; }  // film.h:125-126
                const Float dx = pfilm[i] - 0.5f, dy = pfilm[S + i] - 0.5f;                 // film.h:128
                int p0x = (int)std::ceil(dx - rx), p0y = (int)std::ceil(dy - ry);           // film.h:129-131
                int p1x = (int)std::floor(dx + rx) + 1, p1y = (int)std::floor(dy + ry) + 1;
                p0x = std::max(p0x, px0); p0y = std::max(p0y, py0);                         // film.h:132-133
                p1x = std::min(p1x, px1); p1y = std::min(p1y, py1);
                for (int yy = p0y; yy < p1y; ++yy) {
                    const Float fy = std::abs((yy - dy) * iry * 16);                        // film.h:146-149
                    const int ify = std::min((int)std::floor(fy), 15);
                    for (int xx = p0x; xx < p1x; ++xx) {
                        const Float fx = std::abs((xx - dx) * irx * 16);                    // film.h:139-143
                        const int ifx = std::min((int)std::floor(fx), 15);
                        const Float fw = table[ify * 16 + ifx];                             // film.h:153-154
                        const size_t o = static_cast<size_t>(yy - py0) * pw + (xx - px0);
                        for (int k = 0; k < 3; ++k) sum[3 * o + k] += L[k] * sw * fw;       // film.h:157
                        fws[o] += fw;                                                       // film.h:158
                    }
                }
            }
    for (size_t o = 0; o < static_cast<size_t>(pw) * ph; ++o) {
        const Float *c = &sum[3 * o];
        Float xyz[3], rgb[3];
        xyz[0] = 0.f + (0.412453f * c[0] + 0.357580f * c[1] + 0.180423f * c[2]);  // film.cpp:125-127, spectrum.h:62-66
        xyz[1] = 0.f + (0.212671f * c[0] + 0.715160f * c[1] + 0.072169f * c[2]);
        xyz[2] = 0.f + (0.019334f * c[0] + 0.119193f * c[1] + 0.950227f * c[2]);
        rgb[0] = 3.240479f * xyz[0] - 1.537150f * xyz[1] - 0.498535f * xyz[2];    // film.cpp:178, spectrum.h:56-60
        rgb[1] = -0.969256f * xyz[0] + 1.875991f * xyz[1] + 0.041556f * xyz[2];
        rgb[2] = 0.055648f * xyz[0] - 0.204043f * xyz[1] + 1.057311f * xyz[2];
        if (fws[o] != 0) {                                                        // film.cpp:181-189
            const Float inv = (Float)1 / fws[o];
            for (int k = 0; k < 3; ++k) rgb[k] = std::max((Float)0, rgb[k] * inv);
        }
        for (int k = 0; k < 3; ++k) rgb_out[3 * o + k] = (rgb[k] + 1.f * 0.f) * scale;  // film.cpp:192-203 (no splats on this path)
    }
}
}
