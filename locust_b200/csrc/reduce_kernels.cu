// reduce_kernels.cu -- REDUCE stage and output formatting.
//
// Stands in for kernFindUniqBool -> thrust::partition -> kernGetCount
// (MapReduce/src/main.cu:161-238, host sequence :447-465): marks the head of every
// run of equal keys in the sorted records, merges the runs' counts, and lays the
// unique keys out as blob + offsets with `val` = exclusive prefix sum of counts
// (the index of a key's first occurrence in the reference's sorted token array,
// main.cu:132).  Also printKeyIntValues / printKeyValues (main.cu:126-134, :280-289)
// as device-side formatting.
#include "kernels.cuh"
#include "lookback.cuh"

namespace locust {

constexpr uint32_t kFull = 0xFFFFFFFFu;

// ---- device-wide exclusive scan (single pass, decoupled look-back) ----------------
__global__ void __launch_bounds__(256) scan_u64_kernel(const unsigned long long *in, unsigned long long *out,
                                                       const unsigned long long *n_ptr, uint64_t *lb_state,
                                                       uint32_t *ticket, uint32_t epoch, unsigned long long *total,
                                                       int append_total) {
    __shared__ unsigned long long warp_sum[8];
    __shared__ unsigned long long tile_prefix;
    __shared__ uint32_t tslot;
    const unsigned long long n = *n_ptr;
    const uint32_t ntiles = (uint32_t)((n + kScanTile - 1) / kScanTile);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, w = tid >> 5;
    if (n == 0) {
        if (blockIdx.x == 0 && tid == 0) { if (total) *total = 0ull; if (append_total) out[0] = 0ull; }
        return;
    }
    for (;;) {
        const uint32_t tile = take_ticket(ticket, &tslot);
        if (tile >= ntiles) break;
        const unsigned long long base = (unsigned long long)tile * kScanTile + tid * 8u;
        unsigned long long v[8];
        unsigned long long s = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { v[i] = (base + i < n) ? in[base + i] : 0ull; s += v[i]; }
        unsigned long long x = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(kFull, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31) warp_sum[w] = x;
        __syncthreads();
        unsigned long long woff = 0, agg = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { if ((uint32_t)k < w) woff += warp_sum[k]; agg += warp_sum[k]; }
        if (w == 0) {                          // the first warp chains this tile to its predecessors
            uint64_t *st = lb_state + tile;
            unsigned long long prefix = 0;
            if (tile == 0) {
                if (lane == 0) lb_store(st, lb_pack(epoch, kLbInclusive, agg));
            } else {
                if (lane == 0) lb_store(st, lb_pack(epoch, kLbAggregate, agg));
                prefix = lb_lookback_warp(lb_state, tile, epoch, lane);
                if (lane == 0) lb_store(st, lb_pack(epoch, kLbInclusive, prefix + agg));
            }
            if (lane == 0) {
                tile_prefix = prefix;
                if (tile == ntiles - 1) {
                    if (total) *total = prefix + agg;
                    if (append_total) out[n] = prefix + agg;
                }
            }
        }
        __syncthreads();
        unsigned long long run = tile_prefix + woff + (x - s);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (base + i < n) out[base + i] = run;
            run += v[i];
        }
        __syncthreads();
    }
    return_ticket(ticket);
}

int launch_exclusive_scan_u64(const unsigned long long *in, unsigned long long *out, const unsigned long long *n_ptr,
                              uint64_t n_max, uint64_t *lb_state, uint32_t *ticket, uint32_t epoch,
                              unsigned long long *total, bool append_total, int num_sms, cudaStream_t st) {
    uint64_t g = (n_max + kScanTile - 1) / kScanTile;
    const uint64_t cap = (uint64_t)num_sms * 4;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    scan_u64_kernel<<<(unsigned)g, 256, 0, st>>>(in, out, n_ptr, lb_state, ticket, epoch, total, append_total ? 1 : 0);
    return 1;
}

// ---- run heads -----------------------------------------------------------------------
LC_D bool same_key(const Records &r, const Rec &a, const Rec &b) {
    if (a.be0 != b.be0 || a.be1 != b.be1 || a.len != b.len) return false;
    if (a.len <= 16u) return true;
    const uint8_t *pa = r.longs[a.lidx].ptr, *pb = r.longs[b.lidx].ptr;
    for (uint32_t i = 16; i < a.len; ++i)
        if (pa[i] != pb[i]) return false;
    return true;
}

__global__ void __launch_bounds__(256) head_kernel(Records r, ReduceBufs o, const Counters *ctr) {
    const unsigned long long n = ctr->n_entries;
    const Rec *s = ctr->sort_parity != 0u ? r.b : r.a;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const Rec cur = load_rec(s + i);
        bool head = true;
        if (i > 0) {
            const Rec prev = load_rec(s + i - 1);
            if (prev.be0 == cur.be0 && prev.be1 == cur.be1) {
                if ((cur.be1 & 0xFFull) == 0ull) head = false;          // shorter than 16 bytes: prefix is the key
                else head = !same_key(r, prev, cur);
            }
        }
        o.head[i] = head ? 1 : 0;
        o.rank[i] = head ? 1ull : 0ull;
        o.csum[i] = cur.count;                      // counts in sorted order, prefix-summed next
    }
}

// one thread per run head: where the key sits and how long it is.  The merged count of a run
// is not summed here (a run can be millions of records long when every token is a record):
// it is the difference of two entries of the prefix sum over the sorted counts.
__global__ void __launch_bounds__(256) run_heads_kernel(Records r, ReduceBufs o, const Counters *ctr) {
    const unsigned long long n = ctr->n_entries;
    const Rec *s = ctr->sort_parity != 0u ? r.b : r.a;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        if (!o.head[i]) continue;
        const unsigned long long out = o.rank[i];
        o.key_offsets[out] = s[i].len;
        o.src_pos[out] = (uint32_t)i;
    }
}

// per unique key: val = records in front of its run, count = records in its run (both weighted
// by the records' counts), and the key bytes
__global__ void __launch_bounds__(256) blob_kernel(Records r, ReduceBufs o, Counters *ctr) {
    const unsigned long long n_out = ctr->n_out, n = ctr->n_entries;
    const Rec *s = ctr->sort_parity != 0u ? r.b : r.a;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < n_out;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t pos = o.src_pos[k];
        const unsigned long long before = o.csum[pos];
        const unsigned long long upto = k + 1 < n_out ? o.csum[o.src_pos[k + 1]] : o.csum[n];
        o.vals[k] = before;
        o.counts[k] = upto - before;
        const uint64_t off = o.key_offsets[k], len = o.key_offsets[k + 1] - off;
        if (off + len > o.blob_cap) { atomicOr(&ctr->status, kStatusBlobFull); continue; }
        uint8_t *dst = o.key_blob + off;
        const Rec rec = load_rec(s + pos);
        if (len <= 16) {
            for (uint32_t b = 0; b < (uint32_t)len; ++b)
                dst[b] = (uint8_t)(b < 8 ? rec.be0 >> (56 - 8 * b) : rec.be1 >> (56 - 8 * (b - 8)));
        } else {
            const uint8_t *src = r.longs[rec.lidx].ptr;
            for (uint64_t b = 0; b < len; ++b) dst[b] = src[b];
        }
    }
}

int launch_reduce(Records r, SortScratch s, ReduceBufs o, Counters *ctr, uint32_t epoch, int num_sms,
                  cudaStream_t st) {
    uint64_t g = (r.cap + 255) / 256;
    const uint64_t gcap = (uint64_t)num_sms * 8;
    if (g > gcap) g = gcap;
    if (g < 1) g = 1;
    head_kernel<<<(unsigned)g, 256, 0, st>>>(r, o, ctr);
    int k = 1;
    k += launch_exclusive_scan_u64(o.rank, o.rank, &ctr->n_entries, r.cap, s.lb_state, s.tickets + 32, epoch,
                                   &ctr->n_out, false, num_sms, st);
    // prefix sum over the sorted counts (entry n = sum of all counts = tokens)
    k += launch_exclusive_scan_u64(o.csum, o.csum, &ctr->n_entries, r.cap, s.lb_state, s.tickets + 36, epoch + 2,
                                   &ctr->total, true, num_sms, st);
    run_heads_kernel<<<(unsigned)g, 256, 0, st>>>(r, o, ctr);
    k += 1;
    // key lengths -> offsets (n_out + 1 entries)
    k += launch_exclusive_scan_u64(reinterpret_cast<unsigned long long *>(o.key_offsets),
                                   reinterpret_cast<unsigned long long *>(o.key_offsets), &ctr->n_out, r.cap,
                                   s.lb_state, s.tickets + 34, epoch + 1, &ctr->blob_bytes, true, num_sms, st);
    blob_kernel<<<(unsigned)g, 256, 0, st>>>(r, o, ctr);
    k += 1;
    return k;
}

// ---- formatting ---------------------------------------------------------------------------
LC_D uint32_t dec_digits(unsigned long long v) {
    uint32_t d = 1;
    while (v >= 10ull) { v /= 10ull; ++d; }
    return d;
}
LC_D char *put_dec(char *p, unsigned long long v) {
    const uint32_t d = dec_digits(v);
    for (uint32_t i = d; i-- > 0;) { p[i] = (char)('0' + (uint32_t)(v % 10ull)); v /= 10ull; }
    return p + d;
}
LC_D char *put_str(char *p, const char *s, int n) {
    for (int i = 0; i < n; ++i) p[i] = s[i];
    return p + n;
}

__global__ void __launch_bounds__(256) format_sizes_kernel(ReduceBufs o, const Counters *ctr, int cpu_format,
                                                           unsigned long long *line_off) {
    const unsigned long long n_out = ctr->n_out;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < n_out;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint64_t len = o.key_offsets[k + 1] - o.key_offsets[k];
        unsigned long long sz;
        if (cpu_format == 2) sz = len + 2 + dec_digits(o.counts[k]) + 1;
        else if (cpu_format) sz = 11 + len + 10 + dec_digits(o.counts[k]) + 1;
        else sz = 11 + len + 8 + dec_digits(o.vals[k]) + 10 + dec_digits(o.counts[k]) + 1;
        line_off[k] = sz;
    }
}

__global__ void __launch_bounds__(256) format_write_kernel(ReduceBufs o, Counters *ctr, int cpu_format,
                                                           const unsigned long long *line_off, char *text,
                                                           uint64_t text_cap) {
    const unsigned long long n_out = ctr->n_out;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < n_out;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        if (line_off[k + 1] > text_cap) { atomicOr(&ctr->status, kStatusTextFull); continue; }
        char *p = text + line_off[k];
        const uint64_t off = o.key_offsets[k], len = o.key_offsets[k + 1] - off;
        if (cpu_format != 2) p = put_str(p, "print key: ", 11);
        for (uint64_t b = 0; b < len; ++b) p[b] = (char)o.key_blob[off + b];
        p += len;
        if (cpu_format == 2) {                       // writeKeyIntValues, main.cu:121: "%s \t%d\n"
            p = put_str(p, " \t", 2);
            p = put_dec(p, o.counts[k]);
        } else if (cpu_format) {
            p = put_str(p, " \t value: ", 10);
            p = put_dec(p, o.counts[k]);
        } else {
            p = put_str(p, " \t val: ", 8);
            p = put_dec(p, o.vals[k]);
            p = put_str(p, " \t count: ", 10);
            p = put_dec(p, o.counts[k]);
        }
        *p = '\n';
    }
}

int launch_format_sizes(ReduceBufs o, Counters *ctr, int cpu_format, unsigned long long *line_off, SortScratch s,
                        uint32_t epoch, int num_sms, cudaStream_t st) {
    const unsigned g = (unsigned)num_sms * 4;
    format_sizes_kernel<<<g, 256, 0, st>>>(o, ctr, cpu_format, line_off);
    return 1 + launch_exclusive_scan_u64(line_off, line_off, &ctr->n_out, ~0ull >> 8, s.lb_state, s.tickets + 38,
                                         epoch, &ctr->text_bytes, true, num_sms, st);
}

int launch_format_write(ReduceBufs o, Counters *ctr, int cpu_format, const unsigned long long *line_off, char *text,
                        uint64_t text_cap, int num_sms, cudaStream_t st) {
    const unsigned g = (unsigned)num_sms * 4;
    format_write_kernel<<<g, 256, 0, st>>>(o, ctr, cpu_format, line_off, text, text_cap);
    return 1;
}

}  // namespace locust
