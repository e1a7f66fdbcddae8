// rpf_synth.cpp -- deterministic synthetic feature buffers, host side.
// A counter-based hash of (seed, x, y, s, stream) gives every value, so any row band of a larger frame can be
// generated on its own (multi-GPU sharding) and the CPU oracle and the GPU see identical float32 inputs.
// The 19 planes follow SampleData::data (reference src/custom/sd.h:62-94).
#include "../../include/rpf_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <thread>
#include <vector>

namespace {

inline uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

struct Rng {
    uint64_t key;
    Rng(uint64_t seed, int x, int y, int s) {
        key = mix64(seed ^ mix64(static_cast<uint64_t>(static_cast<uint32_t>(x)) << 32 | static_cast<uint32_t>(y)) ^
                    mix64(0x5851f42d4c957f2dull * static_cast<uint64_t>(s + 1)));
    }
    double uniform(int stream) const {  // [0,1)
        return static_cast<double>(mix64(key + 0x632be59bd9b4e019ull * static_cast<uint64_t>(stream + 1)) >> 11) *
               (1.0 / 9007199254740992.0);
    }
    double normal(int stream) const {  // Box-Muller
        const double u1 = uniform(2 * stream + 100), u2 = uniform(2 * stream + 101);
        return std::sqrt(-2.0 * std::log(1.0 - u1)) * std::cos(6.283185307179586 * u2);
    }
};

void sample_noisy(const Rng &r, int gx, int gy, float out[19]) {
    const double lx = r.uniform(2), ly = r.uniform(3);
    out[0] = static_cast<float>(gx + r.uniform(0));
    out[1] = static_cast<float>(gy + r.uniform(1));
    out[5] = static_cast<float>(lx);
    out[6] = static_cast<float>(ly);
    for (int k = 0; k < 12; ++k)
        out[7 + k] = static_cast<float>(std::sin(0.05 * gx + 0.3 * k) + std::cos(0.04 * gy - 0.2 * k) + 0.3 * r.normal(k) +
                                        0.2 * ((k & 1) ? ly : lx));
    for (int c = 0; c < 3; ++c)
        out[2 + c] = static_cast<float>(std::fabs(0.5 + 0.3 * std::sin(0.07 * gx + c) + 0.3 * std::cos(0.06 * gy) +
                                                  0.4 * lx * ly + 0.1 * r.normal(20 + c)));
}

// Degenerate content of real renders: axis-aligned normals, a black background with zeroed second-hit
// features (rpf.cpp:851-857 leaves n1/p1 at 0 when the path ends), planar positions, duplicated samples.
void sample_degenerate(const Rng &r, int gx, int gy, int s, float out[19]) {
    const bool background = ((gx / 6) + (gy / 5)) % 4 == 0;
    const bool wall = (gx / 9) % 2 == 0;
    const double u0 = (s & 1) ? 0.25 : r.uniform(0), u1 = (s & 1) ? 0.75 : r.uniform(1);  // odd samples share a position
    const double lx = (s & 2) ? 0.5 : r.uniform(2), ly = r.uniform(3);
    out[0] = static_cast<float>(gx + u0);
    out[1] = static_cast<float>(gy + u1);
    out[5] = static_cast<float>(lx);
    out[6] = static_cast<float>(ly);
    for (int k = 0; k < 12; ++k) out[7 + k] = 0.f;
    if (!background) {
        out[7 + (wall ? 0 : 2)] = 1.f;                                  // n0
        out[10] = static_cast<float>(0.1 * (gx + u0));                  // p0
        out[11] = static_cast<float>(0.1 * (gy + u1));
        out[12] = wall ? 2.f : static_cast<float>(0.05 * gx);
        if (s % 3 != 0) {                                               // second hit exists
            out[13 + (wall ? 1 : 0)] = -1.f;
            out[16] = static_cast<float>(lx);
            out[17] = static_cast<float>(0.3 * ly + 0.01 * gx);
            out[18] = 1.f;
        }
    }
    for (int c = 0; c < 3; ++c)
        out[2 + c] = background ? 0.f : static_cast<float>(std::fabs(0.2 + 0.1 * c + 0.5 * lx + 0.05 * r.normal(20 + c)));
}

}  // namespace

// Wider records: the 19 reference channels are generated exactly as above and land in their slots of
// the layout position | colour | random | features; the extra random parameters (time, light sample u / v, ...) are fresh
// uniforms, the extra features (a texture colour, ...) smooth functions of the pixel plus noise and a dependence on the
// extra random parameters, like the reference features. kind 2: like kind 0 but every EXTRA random parameter is the
// constant 0.5 -- a channel that carries no information, so the filter's result must equal the reference layout's.
extern "C" int rpf_synth_planes_layout(int width, int height, int spp, unsigned long long seed, int kind, int x_off, int y_off,
                                       int n_random, int n_features, float *planes) {
    if (n_random == 0) n_random = 2;
    if (n_features == 0) n_features = 12;
    if (width <= 0 || height <= 0 || spp <= 0 || !planes || kind < 0 || kind > 2 || n_random < 2 || n_features < 12 || n_random + n_features > 27)
        return RPF_ERR_INVALID;
    const int nch = 5 + n_random + n_features, ch_feat = 5 + n_random;
    const size_t S = static_cast<size_t>(width) * height * spp;
    unsigned nt = std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 64) nt = 64;
    std::vector<std::thread> th;
    auto work = [&](int t) {
        for (int y = t; y < height; y += static_cast<int>(nt))
            for (int x = 0; x < width; ++x)
                for (int s = 0; s < spp; ++s) {
                    float v[19], w[32];
                    const int gx = x + x_off, gy = y + y_off;
                    Rng r(seed, gx, gy, s);
                    if (kind == 1) sample_degenerate(r, gx, gy, s, v); else sample_noisy(r, gx, gy, v);
                    for (int c = 0; c < 7; ++c) w[c] = v[c];                       // position, colour, pLens
                    for (int k = 0; k < 12; ++k) w[ch_feat + k] = v[7 + k];        // n0, p0, n1, p1
                    double extra = 0.0;
                    for (int q = 2; q < n_random; ++q) {
                        const double u = kind == 2 ? 0.5 : r.uniform(40 + q);
                        w[5 + q] = static_cast<float>(u);
                        extra += u;
                    }
                    for (int k = 12; k < n_features; ++k)
                        w[ch_feat + k] = kind == 1 ? (v[7] != 0.f ? 0.25f * (k - 11) : 0.f)
                                                   : static_cast<float>(0.5 + 0.4 * std::sin(0.09 * gx + 0.7 * k) * std::cos(0.05 * gy) + 0.1 * r.normal(30 + k) +
                                                                        (kind == 2 ? 0.0 : 0.15 * extra));
                    const size_t i = (static_cast<size_t>(y) * width + x) * spp + s;
                    for (int c = 0; c < nch; ++c) planes[c * S + i] = w[c];
                }
    };
    for (unsigned t = 1; t < nt; ++t) th.emplace_back(work, static_cast<int>(t));
    work(0);
    for (auto &t : th) t.join();
    return RPF_OK;
}

extern "C" int rpf_synth_planes(int width, int height, int spp, unsigned long long seed, int kind, int x_off, int y_off,
                                float *planes) {
    if (kind < 0 || kind > 1) return RPF_ERR_INVALID;
    return rpf_synth_planes_layout(width, height, spp, seed, kind, x_off, y_off, 2, 12, planes);
}

// Film constructor's filter table (reference core/film.cpp:69-77) for the box and gaussian filters
// (filters/box.cpp:41, filters/gaussian.h:52-66, filters/gaussian.cpp:41-43). Float arithmetic like the reference.
extern "C" int rpf_film_filter_table(int kind, float radius_x, float radius_y, float alpha, float table[256]) {
    if (!table || (kind != 0 && kind != 1) || !(radius_x > 0.f) || !(radius_y > 0.f)) return RPF_ERR_INVALID;
    const float expX = std::exp(-alpha * radius_x * radius_x), expY = std::exp(-alpha * radius_y * radius_y);
    auto gaussian = [&](float d, float expv) { return std::max(0.f, float(std::exp(-alpha * d * d) - expv)); };
    int offset = 0;
    for (int y = 0; y < 16; ++y)
        for (int x = 0; x < 16; ++x, ++offset) {
            const float px = (x + 0.5f) * radius_x / 16, py = (y + 0.5f) * radius_y / 16;
            table[offset] = kind == 0 ? 1.f : gaussian(px, expX) * gaussian(py, expY);
        }
    return RPF_OK;
}
