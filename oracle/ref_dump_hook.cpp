// oracle/ref_dump_hook.cpp -- TEST INFRASTRUCTURE. Pins the product's feature producer (host/rpf_integrator.cpp:
// FillSampleFilm + Li) against the reference's own. Linked into the UNMODIFIED reference pbrt (every object of
// libpbrt.a, rpf.cpp.o included) ahead of the archive with --allow-multiple-definition, this definition of
// pbrt::visualizeSF replaces the reference's debug-EXR writer (rpf.cpp:37-101). RPFIntegrator::Render calls it right
// after its own FillSampleFilm (rpf.cpp:759-762), so the hook sees exactly the SamplingFilm the reference's producer
// built; it writes it out losslessly (rpf_buffer_write, the product's feature-buffer format) and ends the process
// before the reference's CPU filter starts.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "custom/sample_film.h"
#include "rpf_b200.h"

namespace pbrt {
void visualizeSF(const SamplingFilm &film, const std::string &filename) {
    const int W = film.getWidth(), H = film.getHeight();
    const size_t P = film.samples[0][0].size(), S = static_cast<size_t>(W) * H * P;
    std::vector<float> planes(19 * S), rw(S);
    for (int x = 0; x < W; ++x)
        for (int y = 0; y < H; ++y)
            for (size_t s = 0; s < P; ++s) {
                const SampleData &d = film.samples[x][y][s];
                const size_t i = (static_cast<size_t>(y) * W + x) * P + s;
                for (int c = 0; c < 19; ++c) planes[c * S + i] = static_cast<float>(d.data[c]);
                rw[i] = d.rayWeight;
            }
    const float *ptrs[19];
    for (int c = 0; c < 19; ++c) ptrs[c] = planes.data() + c * S;
    std::string path = filename;
    const std::string::size_type dot = path.rfind('.');
    if (dot != std::string::npos) path = path.substr(0, dot);
    path += ".ref.rpfb";
    rpf_buffer_info info{W, H, static_cast<int>(P), film.pixelBounds.pMin.x, film.pixelBounds.pMin.y, 1};
    const int rc = rpf_buffer_write(path.c_str(), &info, ptrs, rw.data());
    std::printf("ref_dump_hook: %s (%dx%dx%zu) rc=%d\n", path.c_str(), W, H, P, rc);
    std::fflush(stdout);
    std::_Exit(rc == 0 ? 0 : 3);
}
}  // namespace pbrt
