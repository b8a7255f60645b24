// microbench.cu -- design measurements for the MAP kernel on the target GPU (not part of
// the library): throughput of the primitives the tokeniser/combiner is built from.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include "common.cuh"

using namespace locust;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t xorshift(uint32_t &s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }

// mode 0: uniform random slots; mode 1: 14% of lanes hit slot 7 (heavy hitter), rest random
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k_atoms(int iters, uint32_t *sink) {
    extern __shared__ uint32_t tab[];
    for (int i = threadIdx.x; i < 16384; i += 1024) tab[i] = 0;
    __syncthreads();
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 1u;
    for (int i = 0; i < iters; ++i) {
        uint32_t r = xorshift(s);
        uint32_t slot = r >> 18;
        if (MODE == 1 && (r & 127u) < 18u) slot = 7;
        atomicAdd(&tab[slot], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) sink[blockIdx.x] = tab[7];
}

// 64-bit key compare + atomic count: the combiner's hit path
__global__ void __launch_bounds__(1024, 1) k_probe_atoms(int iters, uint32_t *sink) {
    extern __shared__ __align__(16) uint8_t raw[];
    uint64_t *keys = (uint64_t *)raw;
    uint32_t *cnt = (uint32_t *)(raw + 16384 * 8);
    for (int i = threadIdx.x; i < 16384; i += 1024) { keys[i] = i * 0x9E3779B97F4A7C15ull; cnt[i] = 0; }
    __syncthreads();
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 1u;
    uint32_t miss = 0;
    for (int i = 0; i < iters; ++i) {
        uint32_t r = xorshift(s);
        uint32_t slot = r >> 18;
        if ((r & 127u) < 18u) slot = 7;
        uint64_t k = slot * 0x9E3779B97F4A7C15ull;
        if (keys[slot] == k) atomicAdd(&cnt[slot], 1u); else ++miss;
    }
    __syncthreads();
    if (threadIdx.x == 0) sink[blockIdx.x] = cnt[7] + miss;
}

// plain read-modify-write (racy; only to see what LDS+STS costs)
__global__ void __launch_bounds__(1024, 1) k_lds_sts(int iters, uint32_t *sink) {
    extern __shared__ uint32_t tab[];
    for (int i = threadIdx.x; i < 16384; i += 1024) tab[i] = 0;
    __syncthreads();
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 1u;
    for (int i = 0; i < iters; ++i) {
        uint32_t slot = xorshift(s) >> 18;
        tab[slot] = tab[slot] + 1u;
    }
    __syncthreads();
    if (threadIdx.x == 0) sink[blockIdx.x] = tab[7];
}

__global__ void __launch_bounds__(1024, 1) k_match(int iters, uint32_t *sink) {
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 1u;
    uint32_t acc = 0;
    for (int i = 0; i < iters; ++i) {
        uint32_t v = xorshift(s) & 63u;
        acc += __match_any_sync(0xFFFFFFFFu, v);
    }
    if (acc == 0x12345u) sink[0] = acc;
}

// global atomics on a table of `slots` 32-byte slots
__global__ void __launch_bounds__(1024, 1) k_red_global(int iters, unsigned long long *tab, uint32_t mask) {
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 1u;
    for (int i = 0; i < iters; ++i) {
        uint32_t slot = xorshift(s) & mask;
        atomicAdd(&tab[(uint64_t)slot * 4 + 2], 1ull);
    }
}

// streaming read + SWAR classify + popcount: the classifier's own ceiling
__global__ void __launch_bounds__(1024, 1) k_classify(const uint4 *in, uint64_t n16, unsigned long long *out, int classify) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (; i + stride < n16; i += 2 * stride) {
        const uint4 a = __ldg(in + i), b = __ldg(in + i + stride);
        if (classify) {
            acc += __popc(classify16(a.x, a.y, a.z, a.w)) + __popc(classify16(b.x, b.y, b.z, b.w));
        } else {
            acc += (a.x ^ a.y ^ a.z ^ a.w) + (b.x ^ b.y ^ b.z ^ b.w);
        }
    }
    acc = __reduce_add_sync(0xFFFFFFFFu, acc);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, (unsigned long long)acc);
}

template <class F>
static float time_ms(F f, int reps = 5) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(a));
        f();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount;
    printf("device %s, %d SMs, clock %d kHz\n", p.name, sms, p.clockRate);
    uint32_t *sink; CK(cudaMalloc(&sink, 4096 * 4));
    const int iters = 4096;
    const double ops = (double)sms * 1024 * iters;
    CK(cudaFuncSetAttribute(k_atoms<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CK(cudaFuncSetAttribute(k_atoms<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CK(cudaFuncSetAttribute(k_lds_sts, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CK(cudaFuncSetAttribute(k_probe_atoms, cudaFuncAttributeMaxDynamicSharedMemorySize, 196608));
    float ms;
    ms = time_ms([&] { k_atoms<0><<<sms, 1024, 65536>>>(iters, sink); });
    printf("ATOMS.ADD uniform 16K slots      : %8.3f ms  %7.1f Gop/s  %.2f lanes/clk/SM @1.9GHz\n", ms, ops / ms * 1e-6, ops / ms * 1e-6 / sms / 1.9);
    ms = time_ms([&] { k_atoms<1><<<sms, 1024, 65536>>>(iters, sink); });
    printf("ATOMS.ADD 14%% heavy hitter       : %8.3f ms  %7.1f Gop/s  %.2f lanes/clk/SM\n", ms, ops / ms * 1e-6, ops / ms * 1e-6 / sms / 1.9);
    ms = time_ms([&] { k_probe_atoms<<<sms, 1024, 196608>>>(iters, sink); });
    printf("LDS.64 probe + ATOMS (14%% heavy) : %8.3f ms  %7.1f Gop/s  %.2f lanes/clk/SM\n", ms, ops / ms * 1e-6, ops / ms * 1e-6 / sms / 1.9);
    ms = time_ms([&] { k_lds_sts<<<sms, 1024, 65536>>>(iters, sink); });
    printf("LDS+STS uniform                  : %8.3f ms  %7.1f Gop/s  %.2f lanes/clk/SM\n", ms, ops / ms * 1e-6, ops / ms * 1e-6 / sms / 1.9);
    ms = time_ms([&] { k_match<<<sms, 1024>>>(iters, sink); });
    printf("match.any (6-bit values)         : %8.3f ms  %7.1f Gop/s  %.2f lanes/clk/SM\n", ms, ops / ms * 1e-6, ops / ms * 1e-6 / sms / 1.9);
    for (int lg : {17, 20, 26}) {
        unsigned long long *tab; CK(cudaMalloc(&tab, (size_t)32 << lg)); CK(cudaMemset(tab, 0, (size_t)32 << lg));
        const int it2 = 512;
        ms = time_ms([&] { k_red_global<<<sms, 1024>>>(it2, tab, (1u << lg) - 1); });
        printf("RED.64 global, 2^%d slots x 32 B  : %8.3f ms  %7.1f Gop/s\n", lg, ms, (double)sms * 1024 * it2 / ms * 1e-6);
        CK(cudaFree(tab));
    }
    const uint64_t n = 1ull << 30;
    uint4 *buf; CK(cudaMalloc(&buf, n)); CK(cudaMemset(buf, 0x61, n));
    unsigned long long *out; CK(cudaMalloc(&out, 8)); CK(cudaMemset(out, 0, 8));
    for (int c = 0; c < 2; ++c) {
        ms = time_ms([&] { k_classify<<<sms, 1024>>>(buf, n / 16, out, c); });
        printf("stream 1 GiB %-20s: %8.3f ms  %7.1f GB/s\n", c ? "+ SWAR classify" : "(read only)", ms, (double)n / ms * 1e-6);
    }
    for (int c = 0; c < 2; ++c) {
        ms = time_ms([&] { k_classify<<<sms * 2, 1024>>>(buf, n / 16, out, c); });
        printf("stream 1 GiB x2 CTAs %-12s: %8.3f ms  %7.1f GB/s\n", c ? "+ classify" : "(read)", ms, (double)n / ms * 1e-6);
    }
    return 0;
}
