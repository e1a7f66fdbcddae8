// rpf_b200_host.h -- host side (C++) of the B200 RPF filter: the drop-in for the loop
//     for (int box_size : box_sizes) ApplyRPFFilter(samplingFilm, tileSize, box_size);
// of RPFIntegrator::Render (reference src/custom/rpf.cpp:767-775). It packs the reference's SamplingFilm
// (src/custom/sample_film.h:28-45, samples[x - pMin.x][y - pMin.y] = vector<SampleData>) into the structure-of-arrays
// planes of include/rpf_b200.h, runs the passes on the GPU through the C ABI and writes the filtered colours back
// into the same SamplingFilm, which is what ApplyRPFFilter does in place (rpf.cpp:732).
// Compiles against the reference's own headers; INTEGRATION.md shows the three-line change in rpf.cpp.
#ifndef RPF_B200_HOST_H
#define RPF_B200_HOST_H

#include <string>
#include <vector>

#include "custom/sample_film.h"

namespace pbrt {

struct RPFB200Options {
    std::vector<int> boxSizes{7};  // rpf.cpp:767 ships {7}; the comment there lists {55, 35, 17, 7}
    double sigmaSeed = 0.002;      // rpf.cpp:533
    int device = 0;                // CUDA device ordinal
    int flags = 0;                 // RPF_FLAG_* of include/rpf_b200.h; 0 = REF_DEFINED
};

struct RPFB200Report {
    long long nanChannels = 0;       // rpf.cpp:702-705 would exit(1); here they are counted
    double meanNeighbourhood = 0;    // "RPF/Neighborhood Size" (rpf.cpp:588)
    int minNeighbourhood = 0, maxNeighbourhood = 0;
    double filterMs = 0;             // device time of all passes
};

// Returns true on success. On failure nothing in samplingFilm is modified and *error holds the reason
// (no CUDA device, ragged sample counts, neighbourhood above rpf_max_neighbourhood(), CUDA errors).
bool ApplyRPFFilterB200(SamplingFilm &samplingFilm, const RPFB200Options &options, RPFB200Report *report,
                        std::string *error);

}  // namespace pbrt
#endif
