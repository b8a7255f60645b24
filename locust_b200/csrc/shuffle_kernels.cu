// shuffle_kernels.cu -- packing of a device's partial (word, count) records into one
// part per destination.  Destination = hash(word) % nparts, or, for the ordered shuffle, the
// key range the word falls in (partmap: first two key bytes -> part, built from an all-reduced
// histogram so that ranges hold about the same number of keys and part p sorts before p + 1).
//
// Stands in for writeKeyIntValues -> /tmp/out.txt (MapReduce/src/main.cu:116-124,
// :428-430): the reference hands partials to the reduce stage through a text file,
// here they are packed in device memory for an all-to-all over NVLink.  The merge
// side (loadIntermediateFile, main.cu:66-103) is shuffle_merge_kernel in
// map_kernel.cu, next to the table insert it shares with MAP.
#include "kernels.cuh"

namespace locust {

// bin of a key = its first two bytes, big endian (a one-byte key has 0 as its second byte), so
// bins are ordered like the keys
LC_D uint32_t bin_of_short(uint64_t k0) { return (uint32_t)((k0 & 0xFFull) << 8) | (uint32_t)((k0 >> 8) & 0xFFull); }
LC_D uint32_t bin_of_long(const uint8_t *ptr) { return ((uint32_t)ptr[0] << 8) | (uint32_t)ptr[1]; }

LC_D uint32_t part_of_short(uint64_t k0, uint64_t k1, uint32_t nparts, const uint8_t *__restrict__ partmap) {
    if (partmap) return partmap[bin_of_short(k0)];
    return (uint32_t)(hash_key128(k0, k1) >> 32) % nparts;
}
LC_D uint32_t part_of_long(const uint8_t *ptr, uint64_t len, uint32_t nparts, const uint8_t *__restrict__ partmap) {
    if (partmap) return partmap[bin_of_long(ptr)];
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
                                                            const uint8_t *__restrict__ partmap,
                                                            unsigned long long *part_counts) {
    __shared__ unsigned long long sh[2 * kMaxParts];
    for (uint32_t i = threadIdx.x; i < 2 * nparts; i += blockDim.x) sh[i] = 0ull;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += stride) {
        const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(table + i);
        if ((k.x | k.y) != 0ull) atomicAdd(&sh[part_of_short(k.x, k.y, nparts, partmap)], 1ull);
    }
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nl; i += stride) {
        const LongOcc t = long_list[i];
        atomicAdd(&sh[nparts + part_of_long(t.ptr, t.len, nparts, partmap)], long_record_bytes(t.len));
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 2 * nparts; i += blockDim.x)
        if (sh[i]) atomicAdd(&part_counts[i], sh[i]);
}

__global__ void __launch_bounds__(256) shuffle_pack_kernel(const Slot *__restrict__ table, uint64_t slots,
                                                           const LongOcc *__restrict__ long_list, uint64_t long_cap,
                                                           const Counters *ctr, uint32_t nparts,
                                                           const uint8_t *__restrict__ partmap, uint8_t *out,
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
        if (i < slots) v = ld256(table + i);
        const bool occ = (v.x | v.y) != 0ull;
        uint32_t part = 0;
        unsigned long long local = 0;
        if (occ) { part = part_of_short(v.x, v.y, nparts, partmap); local = atomicAdd(&sh_cnt[part], 1ull); }
        __syncthreads();
        for (uint32_t q = threadIdx.x; q < nparts; q += blockDim.x)
            sh_base[q] = sh_cnt[q] ? atomicAdd(&part_cursors[q], sh_cnt[q]) : 0ull;
        __syncthreads();
        if (occ) {
            ulonglong4 *dst = reinterpret_cast<ulonglong4 *>(out + part_offsets[part] + 32ull) + sh_base[part] + local;
            st256(dst, make_ulonglong4(v.x, v.y, v.z, 0ull));
        }
        __syncthreads();
    }
    // long records (rare): one atomic each
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = tid; i < nl; i += stride) {
        const LongOcc t = long_list[i];
        const uint32_t part = part_of_long(t.ptr, t.len, nparts, partmap);
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
                                                                 uint32_t nparts, const uint8_t *__restrict__ partmap,
                                                                 uint8_t *const *__restrict__ dst,
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
        if (i < slots) v = ld256(table + i);
        const bool occ = (v.x | v.y) != 0ull;
        uint32_t part = 0;
        unsigned long long local = 0;
        if (occ) { part = part_of_short(v.x, v.y, nparts, partmap); local = atomicAdd(&sh_cnt[part], 1ull); }
        __syncthreads();
        for (uint32_t q = threadIdx.x; q < nparts; q += blockDim.x)
            sh_base[q] = sh_cnt[q] ? atomicAdd(&part_cursors[q], sh_cnt[q]) : 0ull;
        __syncthreads();
        if (occ) {
            const unsigned long long at = sh_base[part] + local;
            if (at < short_cap) st256(reinterpret_cast<ulonglong4 *>(dst[part] + 32) + at, make_ulonglong4(v.x, v.y, v.z, 0ull));
            else atomicOr(xflag, 1u);
        }
        __syncthreads();
    }
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = tid; i < nl; i += stride) {
        const LongOcc t = long_list[i];
        const uint32_t part = part_of_long(t.ptr, t.len, nparts, partmap);
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
                         const Counters *ctr, uint32_t nparts, const uint8_t *partmap, unsigned long long *part_counts,
                         int num_sms, cudaStream_t st) {
    shuffle_count_kernel<<<grid_for(slots, num_sms), 256, 0, st>>>(table, slots, long_list, long_cap, ctr, nparts,
                                                                  partmap, part_counts);
    return 1;
}

int launch_shuffle_pack(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                        const Counters *ctr, uint32_t nparts, const uint8_t *partmap, uint8_t *out,
                        const unsigned long long *part_offsets, const unsigned long long *part_counts,
                        unsigned long long *part_cursors, int num_sms, cudaStream_t st) {
    shuffle_pack_kernel<<<grid_for(slots, num_sms), 256, 0, st>>>(table, slots, long_list, long_cap, ctr, nparts, partmap,
                                                                 out, part_offsets, part_counts, part_cursors);
    return 1;
}

int launch_shuffle_pack_peers(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                              const Counters *ctr, uint32_t nparts, const uint8_t *partmap, uint8_t *const *dev_dst,
                              uint64_t short_cap, uint64_t long_room, unsigned long long *part_cursors, uint32_t *xflag,
                              int num_sms, cudaStream_t st) {
    shuffle_pack_peers_kernel<<<grid_for(slots, num_sms), 256, 0, st>>>(table, slots, long_list, long_cap, ctr, nparts,
                                                                       partmap, dev_dst, short_cap, long_room,
                                                                       part_cursors, xflag);
    shuffle_headers_kernel<<<(nparts + 255) / 256, 256, 0, st>>>(nparts, dev_dst, short_cap, long_room, part_cursors, ctr, xflag);
    return 2;
}


// ---- ordered shuffle: key histogram and range map ----------------------------------------------
// hist[bin] += (an estimate of the) number of this device's keys whose first two bytes are `bin` (65 536 bins).
__global__ void __launch_bounds__(256) partition_hist_kernel(const Slot *__restrict__ table, uint64_t slots,
                                                             const LongOcc *__restrict__ long_list, uint64_t long_cap,
                                                             const Counters *ctr, unsigned long long *hist) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    // The histogram only steers the balance of the ranges (any monotone map gives a correct result),
    // so a large table is sampled: every eighth slot, which the hash fills like any other.
    const uint64_t step = slots >= (1ull << 18) ? 8ull : 1ull;
    for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * step; i < slots; i += stride * step) {
        const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(table + i);
        if ((k.x | k.y) != 0ull) atomicAdd(&hist[bin_of_short(k.x)], step);
    }
    unsigned long long nl = ctr->n_long;
    if (nl > long_cap) nl = long_cap;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nl; i += stride)
        atomicAdd(&hist[bin_of_long(long_list[i].ptr)], 1ull);
}

// One block of 32 warps, 2048 consecutive bins per warp, read 32 at a time (coalesced): bin b
// goes to part floor(keys_before(b) * nparts / total), which is monotone in b, so part p holds
// a contiguous key range below part p + 1's, and the ranges hold about total / nparts keys each
// (as far as whole bins allow).  Every rank computes the same map from the same all-reduced
// histogram.
// nhists histograms that lie one after the other (one per rank, put there over peer memory) added up
__global__ void __launch_bounds__(256) hist_sum_kernel(const unsigned long long *__restrict__ hists, uint32_t nhists,
                                                       unsigned long long *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;      // 65 536 threads
    unsigned long long v = 0;
    for (uint32_t h = 0; h < nhists; ++h) v += hists[(uint64_t)h * 65536u + i];
    out[i] = v;
}
__global__ void __launch_bounds__(1024) partition_map_kernel(const unsigned long long *__restrict__ hist,
                                                             uint32_t nparts, uint8_t *partmap) {
    __shared__ unsigned long long warp_sum[32];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const unsigned long long *mine = hist + w * 2048u;
    unsigned long long acc = 0;
#pragma unroll 16
    for (uint32_t j = 0; j < 64; ++j) acc += mine[j * 32u + lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    if (lane == 0) warp_sum[w] = acc;
    __syncthreads();
    unsigned long long total = 0, carry = 0;
    for (uint32_t k = 0; k < 32; ++k) { if (k == w) carry = total; total += warp_sum[k]; }
#pragma unroll 4
    for (uint32_t j = 0; j < 64; ++j) {
        const unsigned long long v = mine[j * 32u + lane];
        unsigned long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= (uint32_t)o) incl += y;
        }
        const unsigned long long before = carry + incl - v;
        uint32_t part = 0;
        if (total) {
            // before < total <= 2^64 / 256 in any job that fits a GPU, so the product cannot overflow
            part = (uint32_t)(before * nparts / total);
            if (part >= nparts) part = nparts - 1;
        }
        partmap[w * 2048u + j * 32u + lane] = (uint8_t)part;
        carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
}

int launch_partition_hist(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                          const Counters *ctr, unsigned long long *hist, int num_sms, cudaStream_t st) {
    partition_hist_kernel<<<grid_for(slots >= (1ull << 18) ? slots / 8 : slots, num_sms), 256, 0, st>>>(table, slots, long_list,
                                                                                                       long_cap, ctr, hist);
    return 1;
}
int launch_partition_map(const unsigned long long *hist, uint32_t nhists, unsigned long long *sum_scratch,
                         uint32_t nparts, uint8_t *partmap, cudaStream_t st) {
    if (nhists > 1) {
        hist_sum_kernel<<<256, 256, 0, st>>>(hist, nhists, sum_scratch);
        hist = sum_scratch;
    }
    partition_map_kernel<<<1, 1024, 0, st>>>(hist, nparts, partmap);
    return nhists > 1 ? 2 : 1;
}

// ---- small transfers and a barrier over peer memory (one box, buffers mapped with locust_ipc_*) ------
// The same 16-byte words to every destination (a rank's histogram to all ranks, say).
__global__ void __launch_bounds__(256) peer_put_kernel(uint4 *const *__restrict__ dsts, uint32_t ndst,
                                                       const uint4 *__restrict__ src, uint64_t n16) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint4 v = src[i];
        for (uint32_t d = 0; d < ndst; ++d) dsts[d][i] = v;
    }
    __threadfence_system();
}
// Barrier between the ranks' streams: thread p tells rank p "rank `rank` has reached `epoch`"
// (flags[p][rank] = epoch, a store into p's memory) and then waits until rank p has told us the
// same (flags[rank][p] >= epoch, a load from our own memory).  Everything this stream wrote to
// peers before is visible to them once they pass.  Gives up after five seconds (a rank that
// died must not hang the others' GPUs) and says so in *xflag.
__global__ void peer_barrier_kernel(uint32_t *const *__restrict__ flags, uint32_t rank, uint32_t world, uint32_t epoch,
                                    uint32_t *xflag) {
    const uint32_t p = threadIdx.x;
    if (p >= world) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flags[p] + rank), "r"(epoch) : "memory");
    const uint32_t *mine = flags[rank] + p;
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    for (;;) {
        uint32_t seen;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
        if ((int32_t)(seen - epoch) >= 0) break;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (t - t0 > 5000000000ull) { atomicOr(xflag, 4u); break; }
    }
}
int launch_peer_put(uint4 *const *dev_dsts, uint32_t ndst, const uint4 *src, uint64_t n16, int num_sms, cudaStream_t st) {
    peer_put_kernel<<<grid_for(n16, num_sms), 256, 0, st>>>(dev_dsts, ndst, src, n16);
    return 1;
}
int launch_peer_barrier(uint32_t *const *dev_flags, uint32_t rank, uint32_t world, uint32_t epoch, uint32_t *xflag,
                        cudaStream_t st) {
    peer_barrier_kernel<<<1, 256, 0, st>>>(dev_flags, rank, world, epoch, xflag);
    return 1;
}

// ---- a rank's sorted result as one movable part, and the concatenation of such parts -----------
// Part: header {n, blob_bytes, total, part_bytes}, key_offsets[n + 1], counts[n], vals[n], key bytes.
__global__ void __launch_bounds__(256) result_pack_kernel(ReduceBufs red, const Counters *ctr, uint8_t *out) {
    const unsigned long long n = ctr->n_out, blob = ctr->blob_bytes;
    unsigned long long *h = reinterpret_cast<unsigned long long *>(out);
    unsigned long long *off = h + 4, *cnt = off + n + 1, *val = cnt + n;
    uint8_t *kb = reinterpret_cast<uint8_t *>(val + n);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid == 0) { h[0] = n; h[1] = blob; h[2] = ctr->total; h[3] = (32ull + 8ull * (n + 1) + 16ull * n + blob + 31ull) & ~31ull; }
    for (uint64_t i = tid; i <= n; i += stride) off[i] = red.key_offsets[i];
    for (uint64_t i = tid; i < n; i += stride) { cnt[i] = red.counts[i]; val[i] = red.vals[i]; }
    for (uint64_t i = tid; i < blob; i += stride) kb[i] = red.key_blob[i];
    __threadfence_system();          // `out` may be a peer's memory
}

// Appends one part behind `n_base` keys / `blob_base` key bytes / `total_base` occurrences.
__global__ void __launch_bounds__(256) result_append_kernel(ReduceBufs red, const uint8_t *__restrict__ part,
                                                            unsigned long long n_base, unsigned long long blob_base,
                                                            unsigned long long total_base) {
    const unsigned long long *h = reinterpret_cast<const unsigned long long *>(part);
    const unsigned long long n = h[0], blob = h[1];
    const unsigned long long *off = h + 4, *cnt = off + n + 1, *val = cnt + n;
    const uint8_t *kb = reinterpret_cast<const uint8_t *>(val + n);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (uint64_t i = tid; i <= n; i += stride) red.key_offsets[n_base + i] = off[i] + blob_base;
    for (uint64_t i = tid; i < n; i += stride) {
        red.counts[n_base + i] = cnt[i];
        red.vals[n_base + i] = val[i] + total_base;
    }
    for (uint64_t i = tid; i < blob; i += stride) red.key_blob[blob_base + i] = kb[i];
}

// A rank's four header words alone (for a slot too small to take the part: the reader then sees
// part_bytes > slot and refuses it).
__global__ void result_header_kernel(const Counters *ctr, unsigned long long *h) {
    const unsigned long long n = ctr->n_out, blob = ctr->blob_bytes;
    h[0] = n; h[1] = blob; h[2] = ctr->total; h[3] = (32ull + 8ull * (n + 1) + 16ull * n + blob + 31ull) & ~31ull;
    __threadfence_system();
}
int launch_result_header(const Counters *ctr, unsigned long long *dev_hdr, cudaStream_t st) {
    result_header_kernel<<<1, 1, 0, st>>>(ctr, dev_hdr);
    return 1;
}

int launch_result_pack(ReduceBufs red, const Counters *ctr, uint8_t *out, uint64_t n_hint, int num_sms, cudaStream_t st) {
    result_pack_kernel<<<grid_for(n_hint * 16 + 1024, num_sms), 256, 0, st>>>(red, ctr, out);
    return 1;
}
int launch_result_append(ReduceBufs red, const uint8_t *part, uint64_t n_base, uint64_t blob_base, uint64_t total_base,
                         uint64_t n_hint, int num_sms, cudaStream_t st) {
    result_append_kernel<<<grid_for(n_hint * 16 + 1024, num_sms), 256, 0, st>>>(red, part, n_base, blob_base, total_base);
    return 1;
}

}  // namespace locust
