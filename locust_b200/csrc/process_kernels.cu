// process_kernels.cu -- PROCESS stage: compaction of the combined records out of
// the hash table and their lexicographic sort.
//
// Stands in for thrust::partition(KeyIntValueNotEmpty) + thrust::sort(KIVComparator)
// (MapReduce/src/main.cu:411-415; KeyValue.h:20-33, :79-84).  The sort is a
// hand-written onesweep LSD radix sort (8-bit digits, decoupled look-back, one
// kernel per digit that moves data) over the 16-byte big-endian key prefix; keys
// longer than 16 bytes are put in order afterwards by full comparison inside
// their run of equal prefixes.
#include "kernels.cuh"
#include "long_keys.cuh"
#include "lookback.cuh"

namespace locust {

constexpr uint32_t kFull = 0xFFFFFFFFu;

LC_D uint32_t key_len16(uint64_t k0, uint64_t k1) {
    // number of non-zero bytes of the zero padded little-endian key
    return k1 ? 16u - (uint32_t)(__clzll((long long)k1) >> 3) : 8u - (uint32_t)(__clzll((long long)k0) >> 3);
}

// ---- extract ------------------------------------------------------------------
// Compaction of the occupied slots.  A block takes 1024 consecutive slots per round (four per
// thread, all loads in flight together), counts its occupied ones in shared memory and reserves
// their places with ONE atomic on the global record counter: a warp-level reservation would put
// millions of atomics on that single address when the table is large (BASELINE config 4).
__global__ void __launch_bounds__(256) extract_kernel(const Slot *__restrict__ table, uint64_t slots, Records r,
                                                      Counters *ctr) {
    __shared__ uint32_t warp_cnt[8];
    __shared__ unsigned long long block_base;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint64_t stride = (uint64_t)gridDim.x * 1024u;
    const uint64_t rounds = (slots + stride - 1) / stride;
    for (uint64_t it = 0; it < rounds; ++it) {
        const uint64_t first = it * stride + (uint64_t)blockIdx.x * 1024u;
        ulonglong4 v[4];
        uint32_t mine = 0, m[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint64_t i = first + j * 256u + threadIdx.x;
            v[j] = i < slots ? ld256(table + i) : make_ulonglong4(0ull, 0ull, 0ull, 0ull);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            m[j] = __ballot_sync(kFull, (v[j].x | v[j].y) != 0ull);
            mine += (uint32_t)__popc(m[j]);                      // warp-uniform: this warp's occupied slots
        }
        if (lane == 0) warp_cnt[w] = mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t tot = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) tot += warp_cnt[k];
            block_base = tot ? atomicAdd(&ctr->n_entries, (unsigned long long)tot) : 0ull;
        }
        __syncthreads();
        unsigned long long e = block_base;
        for (uint32_t k = 0; k < w; ++k) e += warp_cnt[k];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if ((v[j].x | v[j].y) != 0ull) {
                const unsigned long long at = e + __popc(m[j] & ((1u << lane) - 1u));
                if (at < r.cap) {
                    Rec rec;
                    rec.be0 = bswap64(v[j].x); rec.be1 = bswap64(v[j].y); rec.count = v[j].z;
                    rec.len = key_len16(v[j].x, v[j].y); rec.lidx = 0xFFFFFFFFu;
                    store_rec(r.a + at, rec);
                } else {
                    atomicOr(&ctr->status, kStatusEntryFull);
                }
            }
            e += __popc(m[j]);
        }
        __syncthreads();
    }
}

// every occurrence of a long token becomes a record with count 1; REDUCE merges them
__global__ void __launch_bounds__(256) append_long_kernel(const LongOcc *__restrict__ list, uint64_t long_cap,
                                                          Records r, Counters *ctr) {
    unsigned long long n = ctr->n_long;
    if (n > long_cap) n = long_cap;
    const unsigned long long base = ctr->n_entries;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long e = base + i;
        if (e >= r.cap) { atomicOr(&ctr->status, kStatusEntryFull); continue; }
        const LongOcc t = list[i];
        uint64_t b0 = 0, b1 = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            b0 = (b0 << 8) | t.ptr[b];
            b1 = (b1 << 8) | t.ptr[8 + b];
        }
        Rec rec;
        rec.be0 = b0; rec.be1 = b1; rec.count = t.count;
        rec.len = t.len > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)t.len; rec.lidx = (uint32_t)i;
        store_rec(r.a + e, rec);
    }
}
__global__ void bump_entries_kernel(uint64_t long_cap, uint64_t cap, Counters *ctr) {
    unsigned long long n = ctr->n_long;
    if (n > long_cap) n = long_cap;
    unsigned long long e = ctr->n_entries + n;
    if (e > cap) e = cap;
    ctr->n_entries = e;
}

// ---- radix sort -----------------------------------------------------------------
LC_D uint32_t digit_of(uint64_t be0, uint64_t be1, int pass) {
    return pass < 8 ? (uint32_t)(be1 >> (8 * pass)) & 0xFFu : (uint32_t)(be0 >> (8 * (pass - 8))) & 0xFFu;
}

__global__ void __launch_bounds__(256) sort_hist_kernel(Records r, const Counters *ctr, uint32_t *hist) {
    __shared__ uint32_t sh[16 * 256];
    for (int i = threadIdx.x; i < 16 * 256; i += 256) sh[i] = 0u;
    __syncthreads();
    const unsigned long long n = ctr->n_entries;
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * 256) {
        const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(r.a + i);     // be0, be1
#pragma unroll
        for (int p = 0; p < 16; ++p) atomicAdd(&sh[p * 256 + digit_of(k.x, k.y, p)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 16 * 256; i += 256)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// one block: per digit position, bucket bases and whether the pass moves anything
__global__ void __launch_bounds__(256) sort_prep_kernel(const uint32_t *hist, uint32_t *bases, Counters *ctr) {
    __shared__ uint32_t warp_tot[2][8];
    const uint32_t d = threadIdx.x, lane = d & 31u, w = d >> 5;
    const unsigned long long n = ctr->n_entries;
    uint32_t parity = 0;
    uint32_t hv[16];
#pragma unroll
    for (int p = 0; p < 16; ++p) hv[p] = hist[p * 256 + d];        // all loads in flight together
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const uint32_t v = hv[p];
        // every key has the same digit here (also n == 0): the pass would move nothing
        const uint32_t uniform = __syncthreads_or((unsigned long long)v == n);
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(kFull, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31) warp_tot[p & 1][w] = x;
        __syncthreads();
        uint32_t off = 0;
        for (uint32_t k = 0; k < w; ++k) off += warp_tot[p & 1][k];
        bases[p * 256 + d] = off + x - v;
        const uint32_t active = uniform ? 0u : 1u;
        if (d == 0) { ctr->pass_active[p] = active; ctr->pass_src[p] = parity; }
        parity ^= active;
    }
    if (d == 0) ctr->sort_parity = parity;
}

// One LSD pass, onesweep style: a tile of 2048 records is ranked in registers
// (warp match + per-warp digit counters), its digit totals are chained to the
// previous tiles' by decoupled look-back, and the records are scattered to their
// final place for this digit.  Stable.
// kStatic: tiles are dealt out by block index (tile = block, block + grid, ...) instead of by
// ticket.  Only safe when every block of the grid is resident (cooperative launch): a tile waits
// for lower-numbered tiles only, and those then belong to blocks that are running or done.
LC_D void sort_one_pass(const bool kStatic, const Records &r, const uint32_t *__restrict__ bases, unsigned long long n,
                        bool from_b, int pass, uint64_t *lb_state, uint32_t *ticket, uint32_t epoch) {
    __shared__ uint32_t wcnt[8][256];
    __shared__ uint32_t tile_base[256];
    __shared__ uint32_t tslot;
    const uint32_t ntiles = (uint32_t)((n + kSortTile - 1) / kSortTile);
    const Rec *src = from_b ? r.b : r.a;
    Rec *dst = from_b ? r.a : r.b;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, w = tid >> 5;

    for (uint32_t round = 0;; ++round) {
        const uint32_t tile = kStatic ? blockIdx.x + round * gridDim.x : take_ticket(ticket, &tslot);
        if (tile >= ntiles) break;
        for (int i = tid; i < 8 * 256; i += 256) (&wcnt[0][0])[i] = 0u;
        __syncthreads();
        ulonglong4 rec[8];                 // {be0, be1, count, len | lidx << 32}
        uint32_t dg[8], rk[8];
        const unsigned long long tbase = (unsigned long long)tile * kSortTile + w * 256u + lane;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const unsigned long long e = tbase + i * 32u;
            const bool valid = e < n;
            rec[i] = valid ? ld256(src + e) : make_ulonglong4(0ull, 0ull, 0ull, 0ull);
            dg[i] = valid ? digit_of(rec[i].x, rec[i].y, pass) : 0x100u;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t peers = __match_any_sync(kFull, dg[i]);
            const uint32_t leader = (uint32_t)__ffs((int)peers) - 1u;
            uint32_t base = 0;
            if (lane == leader && dg[i] < 256u) {
                base = wcnt[w][dg[i]];
                wcnt[w][dg[i]] = base + (uint32_t)__popc(peers);
            }
            base = __shfl_sync(kFull, base, leader);
            rk[i] = base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
            __syncwarp();
        }
        __syncthreads();
        {   // thread = digit: offsets of each warp's records inside the tile, tile total
            uint32_t tot = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) { const uint32_t c = wcnt[k][tid]; wcnt[k][tid] = tot; tot += c; }
            uint64_t *st = lb_state + (uint64_t)tile * 256u + tid;
            uint64_t prefix = 0;
            if (tile == 0) {
                lb_store(st, lb_pack(epoch, kLbInclusive, tot));
            } else {
                lb_store(st, lb_pack(epoch, kLbAggregate, tot));
                prefix = lb_lookback(lb_state, tile, 256u, tid, epoch);
                lb_store(st, lb_pack(epoch, kLbInclusive, prefix + tot));
            }
            tile_base[tid] = bases[pass * 256 + tid] + (uint32_t)prefix;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (dg[i] < 256u) {
                const uint32_t pos = tile_base[dg[i]] + wcnt[w][dg[i]] + rk[i];
                st256(dst + pos, rec[i]);
            }
        }
        __syncthreads();
    }
    if (!kStatic) return_ticket(ticket);
}

// One launch per pass (fallback when a cooperative launch is not possible).
__global__ void __launch_bounds__(256) sort_pass_kernel(Records r, const uint32_t *__restrict__ bases,
                                                        const Counters *ctr, int pass, uint64_t *lb_state,
                                                        uint32_t *ticket, uint32_t epoch) {
    if (!ctr->pass_active[pass]) return;
    sort_one_pass(false, r, bases, ctr->n_entries, ctr->pass_src[pass] != 0u, pass, lb_state, ticket, epoch);
}

// All passes in one cooperative launch: between passes the whole grid meets at a barrier (every
// record of pass p must be in place before pass p + 1 reads it).  Passes that would move nothing
// cost nothing, and a pass no longer pays a kernel launch -- with 50 000 records that was most
// of its time.  *bar counts arrivals; the host zeroes it before the launch.
__global__ void __launch_bounds__(256) sort_all_kernel(Records r, const uint32_t *__restrict__ bases,
                                                       const Counters *ctr, uint64_t *lb_state, uint32_t *tickets,
                                                       uint32_t epoch, uint32_t *bar) {
    uint32_t generation = 0;
    const unsigned long long n = ctr->n_entries;
    uint32_t active = 0, from_b = 0;                                // the plan of sort_prep_kernel, read once
    for (int pass = 0; pass < 16; ++pass) {
        active |= (ctr->pass_active[pass] ? 1u : 0u) << pass;
        from_b |= (ctr->pass_src[pass] ? 1u : 0u) << pass;
    }
    // few tiles (one round): dealt by block index, no ticket traffic -- and blocks without a tile leave
    // at once, so the barrier is between the working blocks only; many tiles: tickets balance the load
    const bool by_index = n <= (unsigned long long)gridDim.x * kSortTile;
    const uint32_t members = by_index ? (uint32_t)((n + kSortTile - 1) / kSortTile) : gridDim.x;
    if (by_index && blockIdx.x >= members) return;
    for (int pass = 0; pass < 16; ++pass) {
        if (!((active >> pass) & 1u)) continue;                     // the same decision in every block
        sort_one_pass(by_index, r, bases, n, ((from_b >> pass) & 1u) != 0u, pass,
                      lb_state, tickets + 2 * pass, epoch + (uint32_t)pass);
        ++generation;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(bar, 1u);
            const uint32_t want = generation * members;
            uint32_t seen;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
                if (seen < want && members > 64u) __nanosleep(32);
            } while (seen < want);
            __threadfence();
        }
        __syncthreads();
    }
}

// ---- keys longer than 16 bytes ---------------------------------------------------
// order of two records that agree on their first 16 bytes
LC_D int cmp_tail(const Records &r, const Rec &a, const Rec &b) {
    const uint32_t la = a.len, lb = b.len;
    const uint32_t m = la < lb ? la : lb;
    if (m > 16u) {
        const uint8_t *pa = r.longs[a.lidx].ptr, *pb = r.longs[b.lidx].ptr;
        for (uint32_t i = 16; i < m; ++i) {
            const uint32_t x = pa[i], y = pb[i];
            if (x != y) return x < y ? -1 : 1;
        }
    }
    return la < lb ? -1 : (la > lb ? 1 : 0);
}

// Records that share their 16-byte prefix with a neighbour ("tied") are put in full-key order.
// The radix sort has left them next to each other in arbitrary order; since the long-key list
// holds distinct keys, a tied group has as many records as there are distinct keys with that
// prefix (times the number of senders after a multi-GPU merge), which can still be millions
// (URLs, paths, identifiers).  One cooperative launch:
//   1. every tied record's position is appended to two arrays K and Q (order irrelevant);
//   2. bitonic sort over both: K by full key (prefix, then the bytes behind it, then length),
//      Q by position; the grid meets at a barrier after every compare-exchange step;
//   3. the t-th smallest tied key belongs at the t-th smallest tied position (untied records do
//      not lie inside a group, and groups are in prefix order already): records are gathered in
//      key order into the spare record buffer and written back to the positions in Q.
// A job without long keys leaves at the first test.
LC_D int cmp_full(const Records &r, const Rec *recs, uint32_t a, uint32_t b) {
    if (a == b) return 0;
    if (a == 0xFFFFFFFFu) return 1;                  // padding sorts last
    if (b == 0xFFFFFFFFu) return -1;
    const Rec x = load_rec(recs + a), y = load_rec(recs + b);
    if (x.be0 != y.be0) return x.be0 < y.be0 ? -1 : 1;
    if (x.be1 != y.be1) return x.be1 < y.be1 ? -1 : 1;
    return cmp_tail(r, x, y);
}

constexpr uint32_t kTieLocal = 2048;

LC_D void grid_meet(uint32_t *bar, uint32_t &generation) {
    ++generation;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        const uint32_t want = generation * gridDim.x;
        uint32_t seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
            if (seen < want) __nanosleep(64);
        } while (seen < want);
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) tie_sort_kernel(Records r, SortScratch sc, Counters *ctr, uint32_t *bar) {
    if (*reinterpret_cast<volatile unsigned long long *>(&ctr->n_long) == 0ull) return;      // the same in every block
    const unsigned long long n = ctr->n_entries;
    Rec *recs = ctr->sort_parity != 0u ? r.b : r.a;
    Rec *spare = ctr->sort_parity != 0u ? r.a : r.b;
    uint32_t *K = sc.tie_k, *Q = sc.tie_q;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t generation = 0;
    // 1. tied records
    for (uint64_t i0 = tid - lane; i0 < n; i0 += nthreads) {
        const uint64_t i = i0 + lane;
        bool tied = false;
        if (i < n) {
            const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(recs + i);
            if ((k.y & 0xFFull) != 0ull) {                                       // 16 bytes or more
                if (i > 0) { const ulonglong2 q = *reinterpret_cast<const ulonglong2 *>(recs + i - 1); tied = q.x == k.x && q.y == k.y; }
                if (!tied && i + 1 < n) { const ulonglong2 q = *reinterpret_cast<const ulonglong2 *>(recs + i + 1); tied = q.x == k.x && q.y == k.y; }
            }
        }
        const uint32_t m = __ballot_sync(kFull, tied);
        if (m) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(&ctr->n_tied, (unsigned long long)__popc(m));
            base = __shfl_sync(kFull, base, 0);
            if (tied) {
                const unsigned long long at = base + __popc(m & ((1u << lane) - 1u));
                if (at < sc.tie_cap) { K[at] = (uint32_t)i; Q[at] = (uint32_t)i; }
                else atomicOr(&ctr->status, kStatusEntryFull | kStatusTieFull);
            }
        }
    }
    grid_meet(bar, generation);
    unsigned long long m;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(m) : "l"(&ctr->n_tied) : "memory");
    if (m < 2ull || m > sc.tie_cap) return;
    uint64_t M = 2;
    while (M < m) M <<= 1;
    if (M <= kTieLocal) {
        // a handful of ties (the usual case: two spellings of one long word, duplicates of a racing insert): one
        // block sorts them in shared memory, no grid barriers
        if (blockIdx.x != 0) return;
        __shared__ uint32_t sk[kTieLocal], sq[kTieLocal];
        for (uint32_t j = threadIdx.x; j < (uint32_t)M; j += blockDim.x) { sk[j] = j < m ? K[j] : 0xFFFFFFFFu; sq[j] = sk[j]; }
        __syncthreads();
        for (uint32_t k = 2; k <= (uint32_t)M; k <<= 1) {
            for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                for (uint32_t i = threadIdx.x; i < (uint32_t)M; i += blockDim.x) {
                    const uint32_t l = i ^ j;
                    if (l <= i) continue;
                    const bool asc = (i & k) == 0;
                    const uint32_t ka = sk[i], kb = sk[l];
                    if ((cmp_full(r, recs, ka, kb) > 0) == asc) { sk[i] = kb; sk[l] = ka; }
                    const uint32_t qa = sq[i], qb = sq[l];
                    if ((qa > qb) == asc) { sq[i] = qb; sq[l] = qa; }
                }
                __syncthreads();
            }
        }
        for (uint32_t j = threadIdx.x; j < (uint32_t)m; j += blockDim.x) st256(spare + j, ld256(recs + sk[j]));
        __syncthreads();
        __threadfence_block();
        for (uint32_t j = threadIdx.x; j < (uint32_t)m; j += blockDim.x) st256(recs + sq[j], ld256(spare + j));
        return;
    }
    for (uint64_t j = m + tid; j < M; j += nthreads) { K[j] = 0xFFFFFFFFu; Q[j] = 0xFFFFFFFFu; }
    grid_meet(bar, generation);
    // 2. bitonic sort of K (by key) and Q (by position)
    for (uint64_t k = 2; k <= M; k <<= 1) {
        for (uint64_t j = k >> 1; j > 0; j >>= 1) {
            for (uint64_t i = tid; i < M; i += nthreads) {
                const uint64_t l = i ^ j;
                if (l <= i) continue;
                const bool asc = (i & k) == 0;
                const uint32_t ka = K[i], kb = K[l];
                if ((cmp_full(r, recs, ka, kb) > 0) == asc) { K[i] = kb; K[l] = ka; }
                const uint32_t qa = Q[i], qb = Q[l];
                if ((qa > qb) == asc) { Q[i] = qb; Q[l] = qa; }
            }
            grid_meet(bar, generation);
        }
    }
    // 3. into place
    for (uint64_t j = tid; j < m; j += nthreads) st256(spare + j, ld256(recs + K[j]));
    grid_meet(bar, generation);
    for (uint64_t j = tid; j < m; j += nthreads) st256(recs + Q[j], ld256(spare + j));
}

// Fallback where a cooperative launch is not possible: the thread that owns the first record of a
// tied run puts the run in order by insertion (fine for the short runs distinct long keys give).
__global__ void __launch_bounds__(256) tie_fix_kernel(Records r, const Counters *ctr) {
    const unsigned long long n = ctr->n_entries;
    if (ctr->n_long == 0ull) return;
    Rec *s = ctr->sort_parity != 0u ? r.b : r.a;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const uint64_t a = s[i].be0, b = s[i].be1;
        if ((b & 0xFFull) == 0ull) continue;                          // key shorter than 16 bytes
        if (i > 0 && s[i - 1].be0 == a && s[i - 1].be1 == b) continue;      // not the first of its run
        if (s[i + 1].be0 != a || s[i + 1].be1 != b) continue;               // run of one
        unsigned long long j = i + 2;
        while (j < n && s[j].be0 == a && s[j].be1 == b) ++j;
        for (unsigned long long x = i + 1; x < j; ++x) {
            const Rec v = load_rec(s + x);
            unsigned long long y = x;
            while (y > i) {
                const Rec u = load_rec(s + y - 1);
                if (cmp_tail(r, u, v) <= 0) break;
                s[y] = u; --y;
            }
            s[y] = v;
        }
    }
}

// One thread per listed long token: end and hash in one word-wise pass, then the long-key list.  This kernel
// has the registers and the L1 the MAP kernels lack (they keep the SM's L1 as shared memory).  A token the list
// says is long may turn out to have <= 16 bytes only if the caller listed it wrongly; it is counted all the same.
__global__ void __launch_bounds__(256) long_tokens_kernel(const __grid_constant__ MapParams p) {
    unsigned long long n = p.ctr->n_far;
    if (n == 0ull) return;
    if (n > p.far_cap) n = p.far_cap;
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < n; i0 += stride) {
        const unsigned long long i = i0 + lane;
        const bool valid = i < n;
        const uint8_t *s = nullptr;
        uint64_t h = 0, L = 0;
        if (valid) {
            const ulonglong2 e = p.far_list[i];
            s = reinterpret_cast<const uint8_t *>(e.x);
            L = long_scan(s, e.y - e.x, h);
        }
        // equal keys of the warp are added up first: a frequent long word would otherwise put every one of its
        // occurrences on the same counter in L2
        const uint32_t peers = __match_any_sync(kFull, valid ? h : 0ull);
        const uint32_t leader = (uint32_t)__ffs((int)peers) - 1u;
        const unsigned long long ls = __shfl_sync(kFull, reinterpret_cast<unsigned long long>(s), leader);
        const unsigned long long ll = __shfl_sync(kFull, (unsigned long long)L, leader);
        const bool same = valid && ll == L && long_equal(reinterpret_cast<const uint8_t *>(ls), s, L);
        const uint32_t agg = __ballot_sync(kFull, same) & peers;
        if (valid) {
            if (!same) long_insert(p, s, L, h, 1ull);
            else if (lane == leader) long_insert(p, s, L, h, (unsigned long long)__popc(agg));
        }
    }
}
__global__ void far_reset_kernel(Counters *ctr) {
    if (ctr->n_far > ctr->far_peak) ctr->far_peak = ctr->n_far;
    ctr->n_far = 0ull;
}

__global__ void __launch_bounds__(256) long_index_clear_kernel(const Counters *ctr, uint32_t *index, uint64_t slots) {
    if (ctr->n_long == 0ull) return;
    uint4 *z = reinterpret_cast<uint4 *>(index);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slots / 4; i += (uint64_t)gridDim.x * blockDim.x)
        z[i] = make_uint4(0u, 0u, 0u, 0u);
}

// ---- host launchers -----------------------------------------------------------------
static inline unsigned grid_for(uint64_t n, unsigned per_block, int num_sms, int blocks_per_sm) {
    uint64_t g = (n + per_block - 1) / per_block;
    const uint64_t cap = (uint64_t)num_sms * blocks_per_sm;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (unsigned)g;
}

int launch_long_tokens(const MapParams &p, int num_sms, cudaStream_t st) {
    long_tokens_kernel<<<num_sms * 8, 256, 0, st>>>(p);
    far_reset_kernel<<<1, 1, 0, st>>>(p.ctr);
    return 2;
}

int launch_long_index_clear(const Counters *ctr, uint32_t *index, uint64_t slots, int num_sms, cudaStream_t st) {
    long_index_clear_kernel<<<num_sms, 256, 0, st>>>(ctr, index, slots);
    return 1;
}

int launch_extract(const Slot *table, uint64_t slots, Records r, Counters *ctr, int num_sms, cudaStream_t st) {
    extract_kernel<<<grid_for(slots, 1024, num_sms, 8), 256, 0, st>>>(table, slots, r, ctr);
    return 1;
}

int launch_append_long(const LongOcc *list, uint64_t long_cap, Records r, Counters *ctr, int num_sms,
                       cudaStream_t st) {
    append_long_kernel<<<grid_for(long_cap, 256, num_sms, 4), 256, 0, st>>>(list, long_cap, r, ctr);
    bump_entries_kernel<<<1, 1, 0, st>>>(long_cap, r.cap, ctr);
    return 2;
}

int launch_sort_records(Records r, SortScratch s, Counters *ctr, uint32_t epoch, int num_sms, cudaStream_t st) {
    // hist holds the 16 x 256 digit counters and, behind them, the grid barrier's arrival counter
    cudaMemsetAsync(s.hist, 0, (16 * 256 + 16) * sizeof(uint32_t), st);
    sort_hist_kernel<<<grid_for(r.cap, 256 * 8, num_sms, 4), 256, 0, st>>>(r, ctr, s.hist);
    sort_prep_kernel<<<1, 256, 0, st>>>(s.hist, s.bases, ctr);
    static int resident_per_sm = -1;        // blocks of sort_all_kernel that fit one SM (same for every device here)
    if (resident_per_sm < 0) {
        int nb = 0, coop = 0, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, sort_all_kernel, 256, 0) != cudaSuccess) nb = 0;
        resident_per_sm = coop ? nb : 0;
    }
    if (resident_per_sm > 0) {
        unsigned g = grid_for(r.cap, kSortTile, num_sms, resident_per_sm);
        uint32_t *bar = s.hist + 16 * 256;
        const uint32_t *bases = s.bases;
        const Counters *cctr = ctr;
        void *args[] = {&r, &bases, &cctr, &s.lb_state, &s.tickets, &epoch, &bar};
        if (cudaLaunchCooperativeKernel(reinterpret_cast<void *>(sort_all_kernel), dim3(g), dim3(256), args, 0, st) ==
            cudaSuccess)
            return 3;
        cudaGetLastError();                 // fall through to one launch per pass
    }
    const unsigned g = grid_for(r.cap, kSortTile, num_sms, 4);
    for (int p = 0; p < 16; ++p)
        sort_pass_kernel<<<g, 256, 0, st>>>(r, s.bases, ctr, p, s.lb_state, s.tickets + 2 * p, epoch + (uint32_t)p);
    return 18;
}

int launch_tie_fix(Records r, SortScratch sc, Counters *ctr, int num_sms, cudaStream_t st) {
    static int resident_per_sm = -1;
    if (resident_per_sm < 0) {
        int nb = 0, coop = 0, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, tie_sort_kernel, 256, 0) != cudaSuccess) nb = 0;
        resident_per_sm = coop ? nb : 0;
    }
    if (resident_per_sm > 0 && sc.tie_k && sc.tie_q) {
        unsigned g = (unsigned)num_sms * (unsigned)(resident_per_sm > 4 ? 4 : resident_per_sm);
        uint32_t *bar = sc.hist + 16 * 256 + 1;          // zeroed with the histograms by launch_sort_records
        void *args[] = {&r, &sc, &ctr, &bar};
        if (cudaLaunchCooperativeKernel(reinterpret_cast<void *>(tie_sort_kernel), dim3(g), dim3(256), args, 0, st) == cudaSuccess)
            return 1;
        cudaGetLastError();
    }
    tie_fix_kernel<<<grid_for(r.cap, 256, num_sms, 8), 256, 0, st>>>(r, ctr);
    return 1;
}

}  // namespace locust
