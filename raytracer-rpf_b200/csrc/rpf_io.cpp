// rpf_io.cpp -- feature-buffer file format, host side.
// The reference has no way to keep a SamplingFilm: its only dump is visualizeSF (src/custom/rpf.cpp:37-101), which
// writes lossy per-channel debug EXRs. This is the lossless counterpart: the 19 structure-of-arrays planes of
// include/rpf_b200.h (SampleData::data order, sd.h:62-94) plus SampleData::rayWeight (sd.h:59), so a buffer rendered
// once on a CPU box can be filtered, benchmarked and compared anywhere.
//
// Layout (little endian): 64-byte header {char magic[4]="RPFB"; u32 version=1; i32 width, height, spp; i32 sample_x0,
// sample_y0 (sampleBounds.pMin); i32 channels; i32 dtype=0 (float32); i32 has_ray_weight; u64 samples; u32 n_random,
// n_features; u32 pad[2]}, then `channels` planes of `samples` floats (index (y*width+x)*spp+s), then the ray-weight plane
// when present. channels = 5 + n_random + n_features; a file written before the channel table existed has 0 / 0 there and
// 19 channels: the reference's record, 2 random parameters and 12 features (sd.h:40-46).
#include "../../include/rpf_b200.h"

#include <cstdint>
#include <cstdio>
#include <cstring>

namespace {
struct Header {
    char magic[4];
    uint32_t version;
    int32_t width, height, spp, sample_x0, sample_y0, channels, dtype, has_ray_weight;
    uint64_t samples;
    uint32_t n_random, n_features;
    uint32_t pad[2];
};
static_assert(sizeof(Header) == 64, "header is 64 bytes");

int read_header(FILE *f, Header &h) {
    if (fread(&h, sizeof h, 1, f) != 1) return RPF_ERR_INVALID;
    if (memcmp(h.magic, "RPFB", 4) != 0 || h.version != 1 || h.dtype != 0) return RPF_ERR_INVALID;
    if (h.n_random == 0 && h.n_features == 0) { h.n_random = 2; h.n_features = 12; }
    if (h.n_random < 1 || h.n_features < 1 || h.channels != static_cast<int32_t>(5 + h.n_random + h.n_features) || h.channels > 32) return RPF_ERR_INVALID;
    if (h.width <= 0 || h.height <= 0 || h.spp <= 0) return RPF_ERR_INVALID;
    if (h.samples != static_cast<uint64_t>(h.width) * h.height * h.spp) return RPF_ERR_INVALID;
    return RPF_OK;
}
}  // namespace

extern "C" int rpf_buffer_write(const char *path, const rpf_buffer_info *info, const float *const planes[RPF_NUM_CHANNELS],
                                const float *ray_weight) {
    if (!path || !info || !planes || info->width <= 0 || info->height <= 0 || info->spp <= 0) return RPF_ERR_INVALID;
    const int nr = info->n_random > 0 ? info->n_random : 2, nf = info->n_features > 0 ? info->n_features : 12, nch = 5 + nr + nf;
    if (nch > 32) return RPF_ERR_INVALID;
    for (int c = 0; c < nch; ++c)
        if (!planes[c]) return RPF_ERR_INVALID;
    Header h{};
    memcpy(h.magic, "RPFB", 4);
    h.version = 1;
    h.width = info->width; h.height = info->height; h.spp = info->spp;
    h.sample_x0 = info->sample_x0; h.sample_y0 = info->sample_y0;
    h.channels = nch;
    h.n_random = static_cast<uint32_t>(nr);
    h.n_features = static_cast<uint32_t>(nf);
    h.dtype = 0;
    h.has_ray_weight = ray_weight ? 1 : 0;
    h.samples = static_cast<uint64_t>(info->width) * info->height * info->spp;
    FILE *f = fopen(path, "wb");
    if (!f) return RPF_ERR_INVALID;
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    for (int c = 0; ok && c < nch; ++c) ok = fwrite(planes[c], sizeof(float), h.samples, f) == h.samples;
    if (ok && ray_weight) ok = fwrite(ray_weight, sizeof(float), h.samples, f) == h.samples;
    ok = (fclose(f) == 0) && ok;
    return ok ? RPF_OK : RPF_ERR_INVALID;
}

extern "C" int rpf_buffer_info_read(const char *path, rpf_buffer_info *info) {
    if (!path || !info) return RPF_ERR_INVALID;
    FILE *f = fopen(path, "rb");
    if (!f) return RPF_ERR_INVALID;
    Header h;
    const int rc = read_header(f, h);
    fclose(f);
    if (rc != RPF_OK) return rc;
    info->width = h.width; info->height = h.height; info->spp = h.spp;
    info->sample_x0 = h.sample_x0; info->sample_y0 = h.sample_y0;
    info->has_ray_weight = h.has_ray_weight;
    info->n_random = static_cast<int>(h.n_random);
    info->n_features = static_cast<int>(h.n_features);
    return RPF_OK;
}

extern "C" int rpf_buffer_read(const char *path, float *planes /* [19][samples] */, float *ray_weight /* [samples] or NULL */) {
    if (!path || !planes) return RPF_ERR_INVALID;
    FILE *f = fopen(path, "rb");
    if (!f) return RPF_ERR_INVALID;
    Header h;
    int rc = read_header(f, h);
    if (rc == RPF_OK) {
        const size_t n = static_cast<size_t>(h.samples) * h.channels;
        if (fread(planes, sizeof(float), n, f) != n) rc = RPF_ERR_INVALID;  // truncated file
    }
    if (rc == RPF_OK && ray_weight) {
        if (h.has_ray_weight) {
            if (fread(ray_weight, sizeof(float), h.samples, f) != h.samples) rc = RPF_ERR_INVALID;
        } else {
            for (uint64_t i = 0; i < h.samples; ++i) ray_weight[i] = 1.f;
        }
    }
    fclose(f);
    return rc;
}
