// smem_micro.cu -- design microbenchmarks for the large-neighbourhood RPF path (not product code).
// Measures on one B200: shared-memory atomicAdd throughput for histogram-like address streams, the
// lane-private read-modify-write alternative, and the latency of a dependent FP64 add chain (the
// reference's sequential sums, ops.h:111-144). Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// mode 0: random word in [0,words); 1: groups of 8 consecutive lanes share an address; 2: all lanes same bank,
// distinct addresses; 3: random, packed u16 halves (add 1 or 1<<16)
template <int MODE>
__global__ void atom_kernel(int words, int iters, unsigned *sink) {
    extern __shared__ unsigned h[];
    for (int i = threadIdx.x; i < words; i += blockDim.x) h[i] = 0;
    __syncthreads();
    uint32_t s = mix(blockIdx.x * 7919u + threadIdx.x);
    for (int it = 0; it < iters; ++it) {
        s = s * 1664525u + 1013904223u;
        uint32_t r = s >> 8;
        if (MODE == 1) r = mix((threadIdx.x >> 3) * 2654435761u + it);
        uint32_t w = r % (unsigned)words;
        if (MODE == 2) w = ((r % (unsigned)(words / 32)) * 32);
        unsigned inc = 1u;
        if (MODE == 3) inc = (r & 0x800000u) ? 0x10000u : 1u;
        atomicAdd(&h[w], inc);
    }
    __syncthreads();
    unsigned acc = 0;
    for (int i = threadIdx.x; i < words; i += blockDim.x) acc += h[i];
    if (acc == 0xdeadbeefu) sink[0] = acc;
}

// indices come from a byte stream in global memory (4 bins per u32 load for each of two channels), like the MI kernel
__global__ void atom_stream_kernel(const uint32_t *__restrict__ ba, const uint32_t *__restrict__ bb, int n4, int nb, int reps,
                                   unsigned *sink) {
    extern __shared__ unsigned h[];
    const int words = (nb * nb + 1) / 2;
    for (int i = threadIdx.x; i < words; i += blockDim.x) h[i] = 0;
    __syncthreads();
    for (int r = 0; r < reps; ++r)
        for (int i = threadIdx.x; i < n4; i += blockDim.x) {
            const uint32_t a = ba[i], b = bb[i];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned c = ((a >> (8 * k)) & 255u) * nb + ((b >> (8 * k)) & 255u);
                atomicAdd(&h[c >> 1], (c & 1) ? 0x10000u : 1u);
            }
        }
    __syncthreads();
    unsigned acc = 0;
    for (int i = threadIdx.x; i < words; i += blockDim.x) acc += h[i];
    if (acc == 0xdeadbeefu) sink[0] = acc;
}

// lane-private u16 read-modify-write, column layout [word][32 lanes] per warp
__global__ void private_kernel(int wpl, int iters, unsigned *sink) {
    extern __shared__ unsigned h[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned *col = h + warp * wpl * 32 + lane;
    for (int w = 0; w < wpl; ++w) col[w * 32] = 0;
    uint32_t s = mix(blockIdx.x * 7919u + threadIdx.x);
    unsigned short *c16 = reinterpret_cast<unsigned short *>(col);
    for (int it = 0; it < iters; ++it) {
        s = s * 1664525u + 1013904223u;
        const unsigned c = (s >> 8) % (unsigned)(wpl * 2);
        unsigned short *p = c16 + (c >> 1) * 64 + (c & 1);
        *p = *p + 1;
    }
    unsigned acc = 0;
    for (int w = 0; w < wpl; ++w) acc += col[w * 32];
    if (acc == 0xdeadbeefu) sink[0] = acc;
}

__global__ void dadd_chain_kernel(const double *__restrict__ v, int n, int active_lanes, double *out) {
    const int lane = threadIdx.x & 31;
    if (lane >= active_lanes) return;
    double acc = 0.0;
    for (int i = 0; i < n; ++i) acc = __dadd_rn(acc, v[i] * (lane + 1));
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
// same but the operand comes from shared memory (one value per step, broadcast), unrolled by the compiler
__global__ void dadd_chain_smem_kernel(const double *__restrict__ v, int n, double *out) {
    extern __shared__ double sv[];
    for (int i = threadIdx.x; i < n; i += blockDim.x) sv[i] = v[i];
    __syncthreads();
    double acc = 0.0, acc2 = 0.0;
    const double m = threadIdx.x + 1;
    for (int i = 0; i < n; ++i) {
        const double x = sv[i] * m;
        acc = __dadd_rn(acc, x);
        acc2 = __dadd_rn(acc2, __dmul_rn(x, x));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + acc2;
}

template <class F>
static float time_ms(F f, int reps = 3) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(a);
        f();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz / 1e6;
    printf("device %s, %d SMs, clock %.3f GHz (nominal max)\n", prop.name, sms, ghz);
    unsigned *sink; CK(cudaMalloc(&sink, 64));
    auto report = [&](const char *name, double updates, float ms) {
        printf("%-58s %8.3f ms  %7.2f updates/clk/SM  (%.2f Gupd/s chip)\n", name, ms, updates / (ms * 1e-3) / (ghz * 1e9) / sms,
               updates / (ms * 1e-3) / 1e9);
    };
    const int iters = 20000;
    for (int threads : {128, 256, 512, 1024}) {
        for (int words : {12288, 24576, 49152}) {
            const int smem = words * 4;
            cudaFuncSetAttribute(atom_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaFuncSetAttribute(atom_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaFuncSetAttribute(atom_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaFuncSetAttribute(atom_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            char nm[128];
            const double upd = 1.0 * sms * threads * iters;
            snprintf(nm, sizeof nm, "ATOMS random     %4d thr %3d KB", threads, smem / 1024);
            report(nm, upd, time_ms([&] { atom_kernel<0><<<sms, threads, smem>>>(words, iters, sink); }));
            if (words == 12288) {
                snprintf(nm, sizeof nm, "ATOMS 8-lane same addr %4d thr %3d KB", threads, smem / 1024);
                report(nm, upd, time_ms([&] { atom_kernel<1><<<sms, threads, smem>>>(words, iters, sink); }));
                snprintf(nm, sizeof nm, "ATOMS same bank  %4d thr %3d KB", threads, smem / 1024);
                report(nm, upd, time_ms([&] { atom_kernel<2><<<sms, threads, smem>>>(words, iters, sink); }));
                snprintf(nm, sizeof nm, "ATOMS random u16 packed %4d thr %3d KB", threads, smem / 1024);
                report(nm, upd, time_ms([&] { atom_kernel<3><<<sms, threads, smem>>>(words, iters, sink); }));
            }
        }
    }
    // 2 CTAs per SM
    {
        const int words = 12288, smem = words * 4, threads = 512;
        report("ATOMS random 2 CTA/SM 512 thr 48 KB", 2.0 * sms * threads * iters,
               time_ms([&] { atom_kernel<0><<<2 * sms, threads, smem>>>(words, iters, sink); }));
    }
    // streamed bins from global memory
    {
        const int N = 24200, nb = 155, n4 = N / 4;
        uint32_t *ba, *bb;
        CK(cudaMalloc(&ba, n4 * 4)); CK(cudaMalloc(&bb, n4 * 4));
        uint32_t *h = new uint32_t[n4];
        uint32_t s = 12345;
        auto fill = [&](uint32_t *d, bool corr) {
            for (int i = 0; i < n4; ++i) {
                uint32_t w = 0;
                for (int k = 0; k < 4; ++k) {
                    s = s * 1664525u + 1013904223u;
                    uint32_t b = corr ? ((i * 4 + k) / 8 * 7 % nb) : (s >> 8) % nb;
                    w |= b << (8 * k);
                }
                h[i] = w;
            }
            cudaMemcpy(d, h, n4 * 4, cudaMemcpyHostToDevice);
        };
        const int smem = ((nb * nb + 1) / 2) * 4;
        cudaFuncSetAttribute(atom_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        for (int threads : {256, 512, 1024}) {
            fill(ba, false); fill(bb, false);
            char nm[128];
            const int reps = 16;
            snprintf(nm, sizeof nm, "ATOMS streamed bins random N=24200 nb=155 %4d thr", threads);
            report(nm, 1.0 * sms * N * reps, time_ms([&] { atom_stream_kernel<<<sms, threads, smem>>>(ba, bb, n4, nb, reps, sink); }));
            fill(ba, true); fill(bb, true);
            snprintf(nm, sizeof nm, "ATOMS streamed bins 8-sample runs same cell %4d thr", threads);
            report(nm, 1.0 * sms * N * reps, time_ms([&] { atom_stream_kernel<<<sms, threads, smem>>>(ba, bb, n4, nb, reps, sink); }));
        }
    }
    // lane-private RMW
    for (int threads : {128, 256}) {
        const int wpl = 64;  // 128 u16 cells per lane
        const int smem = (threads / 32) * wpl * 32 * 4;
        cudaFuncSetAttribute(private_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        char nm[128];
        for (int cps : {1, 4}) {
            snprintf(nm, sizeof nm, "lane-private u16 RMW %4d thr x %d CTA/SM", threads, cps);
            report(nm, 1.0 * cps * sms * threads * iters, time_ms([&] { private_kernel<<<cps * sms, threads, smem>>>(wpl, iters, sink); }));
        }
    }
    // dependent DADD chain
    {
        const int n = 24200;
        double *v, *out;
        CK(cudaMalloc(&v, n * 8)); CK(cudaMalloc(&out, 8 * 1024 * sms * 8));
        double *hv = new double[n];
        for (int i = 0; i < n; ++i) hv[i] = 1.0 / (i + 1);
        cudaMemcpy(v, hv, n * 8, cudaMemcpyHostToDevice);
        float ms = time_ms([&] { dadd_chain_kernel<<<1, 32>>>(v, n, 1, out); });
        printf("DADD chain (global operand, 1 lane): %.1f cycles/step\n", ms * 1e-3 * ghz * 1e9 / n);
        ms = time_ms([&] { dadd_chain_kernel<<<1, 32>>>(v, n, 32, out); });
        printf("DADD chain (global operand, 32 lanes): %.1f cycles/step\n", ms * 1e-3 * ghz * 1e9 / n);
        cudaFuncSetAttribute(dadd_chain_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, n * 8);
        for (int threads : {32, 64, 128, 256, 512}) {
            ms = time_ms([&] { dadd_chain_smem_kernel<<<sms, threads, n * 8>>>(v, n, out); });
            printf("DADD two chains/thread (smem operand) %3d thr/SM: %.2f cycles/step  (%.2f warp-DADD/clk/SM)\n", threads,
                   ms * 1e-3 * ghz * 1e9 / n, 2.0 * (threads / 32) * n / (ms * 1e-3 * ghz * 1e9));
        }
    }
    return 0;
}
