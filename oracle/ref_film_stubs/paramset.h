// oracle/ref_film_stubs/paramset.h -- TEST INFRASTRUCTURE. The Create*Filter factories of the reference's filter
// sources read their parameters from a ParamSet; the harness constructs filters directly, so lookups return defaults.
#ifndef RPF_REF_FILM_STUB_PARAMSET_H
#define RPF_REF_FILM_STUB_PARAMSET_H
#include <string>
#include "filter.h"
namespace pbrt {
class ParamSet {
  public:
    Float FindOneFloat(const std::string &, Float d) const { return d; }
};
}  // namespace pbrt
#endif
