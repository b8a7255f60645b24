// shuffle_kernels.cu -- packing of a device's partial (word, count) records into one
// part per destination, destination = hash(word) % nparts.
//
// Stands in for writeKeyIntValues -> /tmp/out.txt (MapReduce/src/main.cu:116-124,
// :428-430): the reference hands partials to the reduce stage through a text file,
// here they are packed in device memory for an all-to-all over NVLink.  The merge
// side (loadIntermediateFile, main.cu:66-103) is shuffle_merge_kernel in
// map_kernel.cu, next to the table insert it shares with MAP.
#include "kernels.cuh"

namespace locust {

LC_D uint32_t part_of_short(uint64_t k0, uint64_t k1, uint32_t nparts) {
    return (uint32_t)(hash_key128(k0, k1) >> 32) % nparts;
}
LC_D uint32_t part_of_long(const uint8_t *ptr, uint64_t len, uint32_t nparts) {
    uint64_t h = 1469598103934665603ull;
    for (uint64_t i = 0; i < len; ++i) { h ^= ptr[i]; h *= 1099511628211ull; }
    return (uint32_t)(mix64(h) >> 32) % nparts;
}
LC_D uint64_t long_record_bytes(uint64_t len) { return 16ull + ((len + 15ull) & ~15ull); }

constexpr uint32_t kMaxParts = 256;

// Both kernels first add up per part inside the block (shared-memory counters) and touch the
// global per-part counters once per block and round: a million records would otherwise
// serialise on a handful of addresses.
__global__ void __launch_bounds__(256) shuffle_count_kernel(const Slot *__restrict__ table, uint64_t slots,
                                                            const LongOcc *__restrict__ long_list, uint64_t long_cap,
                                                            const Counters *ctr, uint32_t nparts,
                                                            unsigned long long *part_counts) {
    __shared__ unsigned long long sh[2 * kMaxParts];
    for (uint32_t i = threadIdx.x; i < 2 * nparts; i += blockDim.x) sh[i] = 0ull;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += stride) {
        const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(table + i);
        if ((k.x | k.y) != 0ull) atomicAdd(&sh[part_of_short(k.x, k.y, nparts)], 1ull);
    }
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nl; i += stride) {
        const LongOcc t = long_list[i];
        atomicAdd(&sh[nparts + part_of_long(t.ptr, t.len, nparts)], long_record_bytes(t.len));
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 2 * nparts; i += blockDim.x)
        if (sh[i]) atomicAdd(&part_counts[i], sh[i]);
}

__global__ void __launch_bounds__(256) shuffle_pack_kernel(const Slot *__restrict__ table, uint64_t slots,
                                                           const LongOcc *__restrict__ long_list, uint64_t long_cap,
                                                           const Counters *ctr, uint32_t nparts, uint8_t *out,
                                                           const unsigned long long *__restrict__ part_offsets,
                                                           const unsigned long long *__restrict__ part_counts,
                                                           unsigned long long *part_cursors) {
    __shared__ unsigned long long sh_cnt[kMaxParts], sh_base[kMaxParts];
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid < nparts) {       // headers
        unsigned long long *h = reinterpret_cast<unsigned long long *>(out + part_offsets[tid]);
        h[0] = part_counts[tid]; h[1] = part_counts[nparts + tid]; h[2] = 32ull + 32ull * part_counts[tid]; h[3] = 0ull;
    }
    // short records: every round the block reserves its records' places per part with one atomic
    const uint64_t rounds = (slots + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; ++r) {
        for (uint32_t i = threadIdx.x; i < nparts; i += blockDim.x) sh_cnt[i] = 0ull;
        __syncthreads();
        const uint64_t i = r * stride + tid;
        ulonglong4 v = make_ulonglong4(0ull, 0ull, 0ull, 0ull);
        if (i < slots) v = *reinterpret_cast<const ulonglong4 *>(table + i);
        const bool occ = (v.x | v.y) != 0ull;
        uint32_t part = 0;
        unsigned long long local = 0;
        if (occ) { part = part_of_short(v.x, v.y, nparts); local = atomicAdd(&sh_cnt[part], 1ull); }
        __syncthreads();
        for (uint32_t q = threadIdx.x; q < nparts; q += blockDim.x)
            sh_base[q] = sh_cnt[q] ? atomicAdd(&part_cursors[q], sh_cnt[q]) : 0ull;
        __syncthreads();
        if (occ) {
            ulonglong4 *dst = reinterpret_cast<ulonglong4 *>(out + part_offsets[part] + 32ull) + sh_base[part] + local;
            *dst = make_ulonglong4(v.x, v.y, v.z, 0ull);
        }
        __syncthreads();
    }
    // long records (rare): one atomic each
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = tid; i < nl; i += stride) {
        const LongOcc t = long_list[i];
        const uint32_t part = part_of_long(t.ptr, t.len, nparts);
        const unsigned long long at = atomicAdd(&part_cursors[nparts + part], long_record_bytes(t.len));
        uint8_t *dst = out + part_offsets[part] + 32ull + 32ull * part_counts[part] + at;
        reinterpret_cast<unsigned long long *>(dst)[0] = t.len;
        reinterpret_cast<unsigned long long *>(dst)[1] = t.count;
        const uint64_t padded = (t.len + 15ull) & ~15ull;
        for (uint64_t b = 0; b < padded; ++b) dst[16 + b] = b < t.len ? t.ptr[b] : (uint8_t)0;
    }
}

// ---- pack + send in one kernel ------------------------------------------------------------
// Each part is written straight into its destination: dst[p] is this rank's slot inside
// rank p's receive buffer, a peer-mapped address (NVLink) unless p is this rank.  No sizes are
// exchanged beforehand: a slot has room for `short_cap` short records at byte 32 and for
// `long_room` bytes of long records at byte 32 + 32 * short_cap, and its header (written last,
// by shuffle_headers_kernel) says how much of each arrived.  What does not fit is dropped and
// flagged in *xflag; short_cap = the sender's table size can never overflow.
__global__ void __launch_bounds__(256) shuffle_pack_peers_kernel(const Slot *__restrict__ table, uint64_t slots,
                                                                 const LongOcc *__restrict__ long_list,
                                                                 uint64_t long_cap, const Counters *ctr,
                                                                 uint32_t nparts, uint8_t *const *__restrict__ dst,
                                                                 uint64_t short_cap, uint64_t long_room,
                                                                 unsigned long long *part_cursors, uint32_t *xflag) {
    __shared__ unsigned long long sh_cnt[kMaxParts], sh_base[kMaxParts];
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t rounds = (slots + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; ++r) {
        for (uint32_t i = threadIdx.x; i < nparts; i += blockDim.x) sh_cnt[i] = 0ull;
        __syncthreads();
        const uint64_t i = r * stride + tid;
        ulonglong4 v = make_ulonglong4(0ull, 0ull, 0ull, 0ull);
        if (i < slots) v = *reinterpret_cast<const ulonglong4 *>(table + i);
        const bool occ = (v.x | v.y) != 0ull;
        uint32_t part = 0;
        unsigned long long local = 0;
        if (occ) { part = part_of_short(v.x, v.y, nparts); local = atomicAdd(&sh_cnt[part], 1ull); }
        __syncthreads();
        for (uint32_t q = threadIdx.x; q < nparts; q += blockDim.x)
            sh_base[q] = sh_cnt[q] ? atomicAdd(&part_cursors[q], sh_cnt[q]) : 0ull;
        __syncthreads();
        if (occ) {
            const unsigned long long at = sh_base[part] + local;
            if (at < short_cap) reinterpret_cast<ulonglong4 *>(dst[part] + 32)[at] = make_ulonglong4(v.x, v.y, v.z, 0ull);
            else atomicOr(xflag, 1u);
        }
        __syncthreads();
    }
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = tid; i < nl; i += stride) {
        const LongOcc t = long_list[i];
        const uint32_t part = part_of_long(t.ptr, t.len, nparts);
        const uint64_t need = long_record_bytes(t.len);
        const unsigned long long at = atomicAdd(&part_cursors[nparts + part], need);
        if (at + need > long_room) { atomicOr(xflag, 1u); continue; }
        uint8_t *d = dst[part] + 32ull + 32ull * short_cap + at;
        reinterpret_cast<unsigned long long *>(d)[0] = t.len;
        reinterpret_cast<unsigned long long *>(d)[1] = t.count;
        const uint64_t padded = (t.len + 15ull) & ~15ull;
        for (uint64_t b = 0; b < padded; ++b) d[16 + b] = b < t.len ? t.ptr[b] : (uint8_t)0;
    }
    __threadfence_system();          // the records have to be in the peer's memory before the header says so
}

// Long records that overflowed were dropped one by one, so the long area may have holes past
// the first drop: report only what is certainly contiguous (everything, when nothing overflowed;
// nothing, when something did -- the job fails on the flag anyway).
__global__ void shuffle_headers_kernel(uint32_t nparts, uint8_t *const *__restrict__ dst, uint64_t short_cap,
                                       uint64_t long_room, const unsigned long long *part_cursors,
                                       const Counters *ctr, uint32_t *xflag) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nparts) return;
    // the counters are about to be cleared for the merge: keep MAP's overflow bits where REDUCE looks
    if (p == 0 && (ctr->status & (kStatusTableFull | kStatusLongFull))) atomicOr(xflag, 2u);
    unsigned long long ns = part_cursors[p], lb = part_cursors[nparts + p];
    if (ns > short_cap) ns = short_cap;
    if (lb > long_room) lb = 0ull;
    unsigned long long *h = reinterpret_cast<unsigned long long *>(dst[p]);
    h[0] = ns; h[1] = lb; h[2] = 32ull + 32ull * short_cap; h[3] = 0ull;
    __threadfence_system();
}

static inline unsigned grid_for(uint64_t n, int num_sms) {
    uint64_t g = (n + 1023) / 1024;
    const uint64_t cap = (uint64_t)num_sms * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (unsigned)g;
}

int launch_shuffle_count(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                         const Counters *ctr, uint32_t nparts, unsigned long long *part_counts, int num_sms,
                         cudaStream_t st) {
    shuffle_count_kernel<<<grid_for(slots, num_sms), 256, 0, st>>>(table, slots, long_list, long_cap, ctr, nparts,
                                                                  part_counts);
    return 1;
}

int launch_shuffle_pack(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                        const Counters *ctr, uint32_t nparts, uint8_t *out, const unsigned long long *part_offsets,
                        const unsigned long long *part_counts, unsigned long long *part_cursors, int num_sms,
                        cudaStream_t st) {
    shuffle_pack_kernel<<<grid_for(slots, num_sms), 256, 0, st>>>(table, slots, long_list, long_cap, ctr, nparts, out,
                                                                 part_offsets, part_counts, part_cursors);
    return 1;
}

int launch_shuffle_pack_peers(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                              const Counters *ctr, uint32_t nparts, uint8_t *const *dev_dst, uint64_t short_cap,
                              uint64_t long_room, unsigned long long *part_cursors, uint32_t *xflag, int num_sms,
                              cudaStream_t st) {
    shuffle_pack_peers_kernel<<<grid_for(slots, num_sms), 256, 0, st>>>(table, slots, long_list, long_cap, ctr, nparts,
                                                                       dev_dst, short_cap, long_room, part_cursors, xflag);
    shuffle_headers_kernel<<<(nparts + 255) / 256, 256, 0, st>>>(nparts, dev_dst, short_cap, long_room, part_cursors, ctr, xflag);
    return 2;
}

}  // namespace locust
