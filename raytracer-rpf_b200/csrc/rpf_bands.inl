// rpf_bands.inl -- row-band sharding of the RPF passes behind the C ABI; included by rpf_api.cu.
//
// The reference parallelises ApplyRPFFilter over 16x16 pixel tiles of one host (ParallelFor2D, rpf.cpp:542,
// core/parallel.cpp:247-299) and swaps the whole film after a pass (Jacobi: rpf.cpp:527,732). Here the frame is cut into
// row bands, one per GPU:
//   * n_gpus > 1 in rpf_config: ONE process, the parent context owns one child context (band + halo rows) per device;
//     colour halos move with cudaMemcpyPeerAsync over NVLink on a copy stream per device.
//   * rpf_band_attach: one PROCESS per band (torchrun); colour halos move with ncclSend / ncclRecv on the context's copy
//     stream. NCCL is dlopen'ed, the library has no link-time dependency on it.
// Either way a pass filters the rows next to a band edge first, hands them to the copy stream, and filters the interior
// rows while they travel; the only waits are stream-to-stream events. Features never change (rpf.cpp:715 writes colours
// only), so nothing but 3 colour planes x b_delta(next box) rows crosses a band edge per pass.
#pragma once
#include <dlfcn.h>
#include <nccl.h>  // types only; every entry point is resolved with dlsym

namespace {

inline int b_delta(int box) { return (box - 1) / 2; }  // rpf.cpp:561

// ---- equal-work plan -------------------------------------------------------------------------------------------------
int band_plan(int height, int n, const int *boxes, int n_iter, int *own0, const double *row_weight = nullptr) {
    if (height <= 0 || n < 1 || n > height || n_iter < 0 || (n_iter && !boxes) || !own0) return RPF_ERR_INVALID;
    // the exchange after a pass talks to the two adjacent bands only: a band must own at least the rows any LATER pass
    // reaches across its edge, or a pixel would need colours from two bands away
    int reach = 0;
    for (int i = 0; i < n_iter; ++i) {
        if (boxes[i] < 1) return RPF_ERR_INVALID;
        if (i > 0) reach = std::max(reach, b_delta(boxes[i]));
    }
    const int min_rows = std::max(1, reach);
    if (n > 1 && 1ll * n * min_rows > height) return RPF_ERR_INVALID;
    // weight of a row: sum over passes of box * (0.9 box + 0.1 rows of the box inside the frame). Pixels near the top and
    // bottom edge see cut boxes, but measured on B200 (profiles/r2/scaling/scale_n8_v29_calibrated.json) a cut box saves
    // little: most of a pixel's cost does not shrink with its candidates. bench.py re-cuts the bands from a measured step.
    std::vector<double> pre(height + 1, 0.0);
    for (int y = 0; y < height; ++y) {
        double w = 0.0;
        for (int i = 0; i < n_iter; ++i) {
            const int d = b_delta(boxes[i]);
            const int rows = std::min(height - 1, y + d) - std::max(0, y - d) + 1;
            w += static_cast<double>(boxes[i]) * (0.9 * boxes[i] + 0.1 * rows);
        }
        if (n_iter == 0) w = 1.0;
        if (row_weight) w = row_weight[y] > 0.0 ? row_weight[y] : 0.0;  // measured cost per row instead of the geometric model
        pre[y + 1] = pre[y] + w;
    }
    own0[0] = 0;
    own0[n] = height;
    int y = 0;
    for (int k = 1; k < n; ++k) {
        const double target = pre[height] * k / n;
        while (y < height && pre[y + 1] <= target) ++y;
        // y = last row boundary with prefix <= target; take the nearer of y and y + 1
        own0[k] = (y < height && target - pre[y] > pre[y + 1] - target) ? y + 1 : y;
    }
    for (int k = 1; k < n; ++k) own0[k] = std::max(own0[k], own0[k - 1] + min_rows);
    for (int k = n - 1; k >= 1; --k) own0[k] = std::min(own0[k], own0[k + 1] - min_rows);
    return RPF_OK;
}

// ---- NCCL, resolved at run time ---------------------------------------------------------------------------------------
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};

NcclApi *nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    // RTLD_NOLOAD first: a process that already carries an NCCL (torch's bundled one) must keep using that one
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { api.error = dlerror() ? dlerror() : "libnccl.so.2 not found"; return nullptr; }
    bool ok = true;
    auto sym = [&](const char *name) { void *p = dlsym(h, name); if (!p) { ok = false; api.error = std::string("libnccl lacks ") + name; } return p; };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    if (!ok) return nullptr;
    api.handle = h;
    return &api;
}

#define NC(ctx, call)                                                                                         \
    do {                                                                                                      \
        ncclResult_t r_ = (call);                                                                             \
        if (r_ != ncclSuccess)                                                                                \
            return fail(ctx, RPF_ERR_CUDA, "%s failed: %s", #call, nccl_api()->GetErrorString(r_));           \
    } while (0)

// ---- one pass over a band ------------------------------------------------------------------------------------------
// Rows [o0,o1) (context-local) are the band's own. d = rows the NEXT pass reaches across a band edge (0 after the last
// pass); up / down: there is a neighbouring band on that side. The rows a neighbour needs are filtered first, in slabs
// rounded up to the membership kernel's tile height so that no tile straddles the slab boundary; `hand_over` then enqueues
// the exchange on the copy stream; the interior follows on the compute stream.
constexpr int kEdgeQuantum = 8;

template <class F>
int band_pass(rpf_ctx *c, int box, double seed, int o0, int o1, int d, bool up, bool down, F &&hand_over) {
    int rc;
    if (d <= 0 || (!up && !down)) {
        if ((rc = launch_rows(c, box, seed, o0, o1)) != RPF_OK) return rc;
        return hand_over();
    }
    const int e = ((d + kEdgeQuantum - 1) / kEdgeQuantum) * kEdgeQuantum;
    const int top1 = up ? std::min(o0 + e, o1) : o0;
    const int bot0 = down ? std::max(o1 - e, top1) : o1;
    // both edge slabs in ONE set of launches (a chunk with a gap): two slabs of a few dozen rows each would leave most of
    // the SMs idle in the tile-granular kernels
    if (top1 > o0 && o1 > bot0) rc = launch_rows(c, box, seed, o0, top1, bot0, o1);
    else if (top1 > o0) rc = launch_rows(c, box, seed, o0, top1);
    else if (o1 > bot0) rc = launch_rows(c, box, seed, bot0, o1);
    else rc = RPF_OK;
    if (rc != RPF_OK) return rc;
    if ((rc = hand_over()) != RPF_OK) return rc;
    if (bot0 > top1 && (rc = launch_rows(c, box, seed, top1, bot0)) != RPF_OK) return rc;
    return RPF_OK;
}

int ensure_band_streams(rpf_ctx *c) {
    if (c->copy_stream) return RPF_OK;
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    CU(c, cudaEventCreateWithFlags(&c->ev_edge, cudaEventDisableTiming));
    CU(c, cudaEventCreateWithFlags(&c->ev_copy, cudaEventDisableTiming));
    CU(c, cudaEventCreateWithFlags(&c->ev_start, cudaEventDisableTiming));
    return RPF_OK;
}

// ---- one band per process: NCCL send / recv ---------------------------------------------------------------------------
int attached_filter(rpf_ctx *c, const int *boxes, const double *seeds, double seed_all, int n_iter) {
    NcclApi *nc = nccl_api();
    if (!nc || !c->nccl_comm) return fail(c, RPF_ERR_STATE, "band context without an NCCL communicator");
    ncclComm_t comm = static_cast<ncclComm_t>(c->nccl_comm);
    const rpf_band &b = c->band;
    const int o0 = b.own0 - b.buf0, o1 = b.own1 - b.buf0;
    const size_t row = static_cast<size_t>(c->cfg.width) * c->cfg.spp;
    for (int i = 0; i < n_iter; ++i) {
        const int d = i + 1 < n_iter ? b_delta(boxes[i + 1]) : 0;
        if (b_delta(boxes[i]) > b.own0 - b.buf0 && b.rank > 0)
            return fail(c, RPF_ERR_INVALID, "box %d reaches %d rows across the band edge, the context holds %d halo rows", boxes[i],
                        b_delta(boxes[i]), b.own0 - b.buf0);
        const bool up = b.rank > 0, down = b.rank + 1 < b.world;
        if ((up || down) && d > o1 - o0)
            return fail(c, RPF_ERR_INVALID, "box %d reaches %d rows across a band edge, this band owns %d: use fewer bands", boxes[i + 1], d, o1 - o0);
        int rc = band_pass(c, boxes[i], seeds ? seeds[i] : seed_all, o0, o1, d, up, down, [&]() -> int {
            if (d <= 0 || (!up && !down)) return RPF_OK;
            double *next = c->d_col[c->cur ^ 1];
            CU(c, cudaEventRecord(c->ev_edge, c->stream));
            CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_edge, 0));
            const int n_send = std::min(d, o1 - o0);
            NC(c, nc->GroupStart());
            for (int k = 0; k < 3; ++k) {
                double *plane = next + k * c->S;
                if (up) {
                    const int n_recv = std::min(d, o0);
                    NC(c, nc->Send(plane + o0 * row, n_send * row, ncclDouble, b.rank - 1, comm, c->copy_stream));
                    NC(c, nc->Recv(plane + (o0 - n_recv) * row, n_recv * row, ncclDouble, b.rank - 1, comm, c->copy_stream));
                }
                if (down) {
                    const int n_recv = std::min(d, c->cfg.height - o1);
                    NC(c, nc->Send(plane + (o1 - n_send) * row, n_send * row, ncclDouble, b.rank + 1, comm, c->copy_stream));
                    NC(c, nc->Recv(plane + o1 * row, n_recv * row, ncclDouble, b.rank + 1, comm, c->copy_stream));
                }
            }
            NC(c, nc->GroupEnd());
            CU(c, cudaEventRecord(c->ev_copy, c->copy_stream));
            c->halo_bytes += 3ull * n_send * row * sizeof(double) * ((up ? 1 : 0) + (down ? 1 : 0));
            return RPF_OK;
        });
        if (rc != RPF_OK) return rc;
        if (d > 0 && (up || down)) CU(c, cudaStreamWaitEvent(c->stream, c->ev_copy, 0));
        c->cur ^= 1;  // rpf.cpp:732
        c->col4_valid = false;
    }
    return RPF_OK;
}

// ---- one process, one band per device: peer copies ----------------------------------------------------------------
int sharded_filter(rpf_ctx *par, const int *boxes, const double *seeds, double seed_all, int n_iter) {
    const int G = static_cast<int>(par->shards.size());
    const size_t row = static_cast<size_t>(par->cfg.width) * par->cfg.spp;
    for (int i = 0; i < n_iter; ++i) {
        if (b_delta(boxes[i]) > par->halo_rows)
            return fail(par, RPF_ERR_INVALID, "box %d needs %d halo rows, the sharded context was created for %d (rpf_config.max_box)",
                        boxes[i], b_delta(boxes[i]), par->halo_rows);
        const int d = i + 1 < n_iter ? b_delta(boxes[i + 1]) : 0;
        for (int g = 0; g < G && G > 1; ++g)
            if (d > par->shard_own0[g + 1] - par->shard_own0[g])
                return fail(par, RPF_ERR_INVALID, "box %d reaches %d rows across a band edge, band %d owns %d: use fewer devices", boxes[i + 1], d, g,
                            par->shard_own0[g + 1] - par->shard_own0[g]);
        // every band marks "my previous pass has drained" before it starts this one: a neighbour may overwrite the halo
        // rows of this band's NEXT buffer (the buffer its previous pass read) only after that
        for (int g = 0; g < G; ++g) {
            rpf_ctx *c = par->shards[g];
            CU(par, cudaSetDevice(c->cfg.device));
            CU(par, cudaEventRecord(c->ev_start, c->stream));
        }
        for (int g = 0; g < G; ++g) {
            rpf_ctx *c = par->shards[g];
            const int o0 = par->shard_own0[g] - par->shard_buf0[g], o1 = par->shard_own0[g + 1] - par->shard_buf0[g];
            const bool up = g > 0, down = g + 1 < G;
            CU(par, cudaSetDevice(c->cfg.device));
            int rc = band_pass(c, boxes[i], seeds ? seeds[i] : seed_all, o0, o1, d, up, down, [&]() -> int {
                if (d <= 0 || (!up && !down)) return RPF_OK;
                CU(par, cudaEventRecord(c->ev_edge, c->stream));
                CU(par, cudaStreamWaitEvent(c->copy_stream, c->ev_edge, 0));
                const double *src = c->d_col[c->cur ^ 1];
                const int n_send = std::min(d, o1 - o0);
                for (int side = 0; side < 2; ++side) {
                    if (side == 0 ? !up : !down) continue;
                    const int h = side == 0 ? g - 1 : g + 1;
                    rpf_ctx *t = par->shards[h];
                    CU(par, cudaStreamWaitEvent(c->copy_stream, t->ev_start, 0));
                    double *dst = t->d_col[t->cur ^ 1];
                    // global rows [r0, r0 + n_send) of this band's fresh colours -> the same global rows in the neighbour
                    const int r0 = side == 0 ? par->shard_own0[g] : par->shard_own0[g + 1] - n_send;
                    const size_t so = static_cast<size_t>(r0 - par->shard_buf0[g]) * row, to = static_cast<size_t>(r0 - par->shard_buf0[h]) * row;
                    for (int k = 0; k < 3; ++k)
                        CU(par, cudaMemcpyPeerAsync(dst + k * t->S + to, t->cfg.device, src + k * c->S + so, c->cfg.device,
                                                    n_send * row * sizeof(double), c->copy_stream));
                    par->halo_bytes += 3ull * n_send * row * sizeof(double);
                }
                CU(par, cudaEventRecord(c->ev_copy, c->copy_stream));
                return RPF_OK;
            });
            if (rc != RPF_OK) { par->err = c->err; return rc; }
        }
        for (int g = 0; g < G; ++g) {
            rpf_ctx *c = par->shards[g];
            CU(par, cudaSetDevice(c->cfg.device));
            if (d > 0) {  // the next pass of this band reads what its neighbours copied in
                if (g > 0) CU(par, cudaStreamWaitEvent(c->stream, par->shards[g - 1]->ev_copy, 0));
                if (g + 1 < G) CU(par, cudaStreamWaitEvent(c->stream, par->shards[g + 1]->ev_copy, 0));
            }
            c->cur ^= 1;  // rpf.cpp:732
            c->col4_valid = false;
        }
    }
    return RPF_OK;
}

}  // namespace
