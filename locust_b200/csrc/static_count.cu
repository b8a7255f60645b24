// static_count.cu -- locust_count_static(): the MAP stage (tokenise + combine) for
// callers that must not allocate device memory.
//
// Who needs this.  liblocust's normal path (locust_create) sizes its combiner table
// with cudaMalloc and grows it when a job turns out to be larger.  Some hosts of this
// library cannot do that: a worker that shares its GPU with a serving process which
// has already claimed the device memory (cudaMalloc simply fails there), a plug-in
// running inside a framework whose pooling allocator owns the heap, or a latency
// path that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply.
// For them the combiner's scratch -- the job counters and the hash table -- lives in
// one fixed array that is part of the library image itself.  Every GPU that loads
// liblocust.so carries that array (about 1 MiB) until the process exits, whether or
// not this entry point is ever called.
//
// The price is a fixed ceiling: at most LOCUST_STATIC_MAX_KEYS distinct words of at most 16
// bytes each, and an input of at most LOCUST_STATIC_MAX_BYTES.  Anything larger is
// refused with LOCUST_E_CAPACITY / LOCUST_E_ARG and a message; nothing is grown.
#include <cstdio>
#include <cstring>
#include <mutex>

#include "kernels.cuh"
#include "../../include/locust.h"

using namespace locust;

namespace {

constexpr uint32_t kStaticSlots = LOCUST_STATIC_MAX_KEYS * 2;           // load factor <= 0.5
constexpr size_t kCountersBytes = 256;                                  // Counters, padded
constexpr size_t kScratchBytes = kCountersBytes + (size_t)kStaticSlots * sizeof(Slot);
static_assert(sizeof(Counters) <= kCountersBytes, "Counters must fit its slot");
static_assert((kStaticSlots & (kStaticSlots - 1)) == 0, "table size must be a power of two");

// The scratch itself.  In host code this name is only a handle for
// cudaGetSymbolAddress; it is not a device address and must never be passed to a
// kernel as a pointer.
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void *g_scratch_addr[kMaxDevices];      // per device: resolved device address, nullptr until set up

void say(char *errbuf, size_t errlen, const char *msg, cudaError_t e = cudaSuccess) {
    if (!errbuf || errlen == 0) return;
    if (e != cudaSuccess) snprintf(errbuf, errlen, "%s: %s", msg, cudaGetErrorString(e));
    else snprintf(errbuf, errlen, "%s", msg);
}

// copies the occupied slots out as 32-byte records {k0, k1, count, 0}
__global__ void __launch_bounds__(256) static_extract_kernel(const Slot *__restrict__ table, uint32_t slots,
                                                             uint64_t *__restrict__ out, uint64_t max_records,
                                                             Counters *ctr) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += gridDim.x * blockDim.x) {
        const ulonglong4 v = *reinterpret_cast<const ulonglong4 *>(table + i);
        if ((v.x | v.y) == 0ull) continue;
        const unsigned long long e = atomicAdd(&ctr->n_entries, 1ull);
        if (e < max_records) {
            out[4 * e + 0] = v.x; out[4 * e + 1] = v.y; out[4 * e + 2] = v.z; out[4 * e + 3] = 0ull;
        }
    }
}

}  // namespace

extern "C" int locust_count_static(const void *dev_data, uint64_t nbytes, void *dev_records, uint64_t max_records,
                                   uint64_t *n_records, void *stream, char *errbuf, size_t errlen) {
    if (errbuf && errlen) errbuf[0] = '\0';
    if (!n_records || (!dev_data && nbytes) || (!dev_records && max_records)) {
        say(errbuf, errlen, "locust_count_static: NULL argument");
        return LOCUST_E_ARG;
    }
    *n_records = 0;
    if (nbytes > LOCUST_STATIC_MAX_BYTES) {
        say(errbuf, errlen, "locust_count_static: input larger than LOCUST_STATIC_MAX_BYTES");
        return LOCUST_E_ARG;
    }
    if (max_records > LOCUST_STATIC_MAX_KEYS) max_records = LOCUST_STATIC_MAX_KEYS;   // a larger output buffer is harmless
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { say(errbuf, errlen, "cudaGetDevice", e); return LOCUST_E_CUDA; }
    if (dev < 0 || dev >= kMaxDevices) { say(errbuf, errlen, "locust_count_static: device ordinal out of range"); return LOCUST_E_ARG; }

    // Set-up, once per device, with that device current: turn the symbol into the
    // address kernels can use, and let the map kernel opt into its shared memory.
    unsigned char *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_mu);
        if (!g_scratch_addr[dev]) {
            void *ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, g_static_scratch);
            if (e != cudaSuccess || !ptr) { say(errbuf, errlen, "cudaGetSymbolAddress(g_static_scratch)", e); return LOCUST_E_CUDA; }
            e = map_kernel_setup();
            if (e != cudaSuccess) { say(errbuf, errlen, "map kernel set-up", e); return LOCUST_E_CUDA; }
            g_scratch_addr[dev] = ptr;
        }
        scratch = static_cast<unsigned char *>(g_scratch_addr[dev]);
    }
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) { say(errbuf, errlen, "cudaDeviceGetAttribute", e); return LOCUST_E_CUDA; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // The array still holds the previous call's table and counters: both must start at zero.
    e = cudaMemsetAsync(scratch, 0, kScratchBytes, st);
    if (e != cudaSuccess) { say(errbuf, errlen, "cudaMemsetAsync(scratch)", e); return LOCUST_E_CUDA; }

    Counters *ctr = reinterpret_cast<Counters *>(scratch);
    Slot *table = reinterpret_cast<Slot *>(scratch + kCountersBytes);
    if (nbytes) {
        MapParams p{};
        const uintptr_t a = reinterpret_cast<uintptr_t>(dev_data);
        const uintptr_t base = a & ~uintptr_t(15);
        p.base = reinterpret_cast<const uint8_t *>(base);
        p.lo = a - base;
        p.hi = p.lo + nbytes;
        p.table = table;
        p.table_mask = kStaticSlots - 1;
        p.long_list = nullptr;            // no room for words longer than 16 bytes: they are reported, not stored
        p.long_cap = 0;
        p.ctr = ctr;
        p.flags = 0;
        launch_tokenize_combine(p, sms, st);
    }
    static_extract_kernel<<<sms * 2, 256, 0, st>>>(table, kStaticSlots, static_cast<uint64_t *>(dev_records), max_records, ctr);
    e = cudaGetLastError();
    if (e != cudaSuccess) { say(errbuf, errlen, "kernel launch", e); return LOCUST_E_CUDA; }
    Counters h;
    e = cudaMemcpyAsync(&h, ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { say(errbuf, errlen, "reading the counters back", e); return LOCUST_E_CUDA; }
    if (h.status & (kStatusLongFull | kStatusFarFull)) {
        say(errbuf, errlen, "locust_count_static: the input holds a word longer than 16 bytes; use locust_create/locust_map");
        return LOCUST_E_CAPACITY;
    }
    if ((h.status & kStatusTableFull) || h.n_entries > LOCUST_STATIC_MAX_KEYS) {
        say(errbuf, errlen, "locust_count_static: more than LOCUST_STATIC_MAX_KEYS distinct words; use locust_create/locust_map");
        return LOCUST_E_CAPACITY;
    }
    if (h.n_entries > max_records) {
        *n_records = h.n_entries;
        say(errbuf, errlen, "locust_count_static: dev_records is too small (n_records holds the number needed)");
        return LOCUST_E_CAPACITY;
    }
    *n_records = h.n_entries;
    return LOCUST_OK;
}
