// map_kernel.cu -- MAP stage: streaming tokeniser fused with the combiner.
//
// Stands in for kernMap/map (MapReduce/src/main.cu:136-159) and, through the
// combiner, for most of the work of thrust::sort + kernFindUniqBool/kernGetCount
// (main.cu:415, :161-238): instead of emitting one 40-byte record per token and
// sorting them all, every token is counted into a hash table keyed by its bytes.
//
// Layout.  One persistent CTA of 32 warps per SM; warps work on their own.  A warp
// owns a contiguous range of 512-byte rows.  Per row:
//   A. each lane holds 16 bytes (one coalesced 128-bit load, two rows prefetched in
//      registers), classifies them with the SWAR delimiter classifier (common.cuh),
//      stages bytes and delimiter bits in the warp's shared-memory area and writes
//      the positions of its token starts into a dense list (warp prefix sum);
//   B. the warp walks that list 32 tokens at a time, one token per lane, so every
//      lane is busy whatever the token lengths: token length from the staged
//      delimiter bits, token bytes from the staged row (PRMT), tokens of <= 8 bytes
//      counted in a 2-way shared-memory table (exact 64-bit keys), misses and tokens
//      of 9..16 bytes queued and inserted into the global table (exact 128-bit keys,
//      128-bit CAS) 32 at a time; longer tokens are listed by address.
#include <cstdlib>

#include "kernels.cuh"
#include "long_keys.cuh"
#include "smem_ptx.cuh"
#include "../../include/locust.h"

namespace locust {

constexpr int kMapThreads = 1024;
constexpr int kMapWarps = kMapThreads / 32;
#ifndef LOCUST_INS
#define LOCUST_INS 0
#endif
// EXPERIMENT, off by default (LOCUST_INS = 0): of the 32 warps, kInsWarps only insert.  The tokenising warps hand
// them their queued keys through a ring in shared memory (kXSlots batches of 32 keys, an mbarrier pair per slot), so
// that the L2 round trips of the global table stall warps that have nothing else to do.  Correct, but slower on every
// text measured (1 GiB, one B200: C2 1.75 ms against 1.49 ms, C3 2.10 against 1.80 with two inserting warps taking one
// batch at a time; 1.76 / 2.17 taking up to four): two warps fewer tokenise, the hand-over costs the tokenisers about
// what the inline insert did, and eight warps per scheduler already hide most of that latency.
constexpr int kInsWarps = LOCUST_INS;
constexpr uint32_t kXSlots = 8;
constexpr uint32_t oXHead = 0, oXTail = 4, oXDone = 8, oXFull = 16, oXEmpty = oXFull + 8 * kXSlots, oXCount = oXEmpty + 8 * kXSlots, oXData = 256;
constexpr uint32_t kXBytes = oXData + kXSlots * 512;
constexpr uint32_t kPoison = 0xFFFFFFFFu;
constexpr uint32_t kBuckets = (232448u - kXBytes - (kMapThreads / 32) * 2160u) / 24u / 16u * 16u;       // shared table: buckets of two 64-bit keys
constexpr uint32_t kSmemSlots = 2 * kBuckets;
constexpr uint32_t kQueueSlots = 64;              // per warp, 16 bytes each
// per-warp area: staged row (512 + 32), delimiter bits (17 words, padded), token list, queue
constexpr uint32_t kRowOff = 0, kBitsOff = 544, kDescOff = 624, kQueueOff = 1136;
constexpr uint32_t kWarpArea = kQueueOff + kQueueSlots * 16;           // 2160
constexpr uint32_t kKeysBytes = kBuckets * 16, kCntBytes = kSmemSlots * 4;
constexpr uint32_t kXOff = kKeysBytes + kCntBytes;                      // hand-over ring
constexpr uint32_t kAreasOff = kXOff + kXBytes;
constexpr uint32_t kSmemBytes = kAreasOff + kMapWarps * kWarpArea;
static_assert(kXOff % 128 == 0, "alignment of the hand-over ring");
static_assert(kSmemBytes <= 232448, "shared memory budget of one CTA");
static_assert(kWarpArea % 16 == 0 && kQueueOff % 16 == 0, "alignment");
constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr uint32_t kMaxProbe = 1u << 14;
#ifndef LOCUST_IMAD
#define LOCUST_IMAD 4
#endif
#ifndef LOCUST_SCANSTEP
#define LOCUST_SCANSTEP 1
#endif
#ifndef LOCUST_LISTUNROLL
#define LOCUST_LISTUNROLL 5        // token list: this many predicated steps, then a loop for what is left (0: loop only)
#endif
#ifndef LOCUST_SLOT32
#define LOCUST_SLOT32 1
#endif
constexpr int kImadAdds = LOCUST_IMAD;                     // of the classifier's nine adds per word, issued on the FMA pipe
constexpr uint32_t kSegments = 1;                 // pieces of the input per warp (8 measured: -2.5% on C2, +4% on C5)

LC_D uint4 ldg_stream(const uint8_t *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

LC_D void cas128(Slot *s, uint64_t n0, uint64_t n1, uint64_t &o0, uint64_t &o1) {
    // compare value is the empty key (0, 0)
    asm volatile(
        "{\n\t"
        ".reg .b128 c, n, d;\n\t"
        "mov.b128 c, {%3, %3};\n\t"
        "mov.b128 n, {%4, %5};\n\t"
        "atom.global.relaxed.gpu.cas.b128 d, [%2], c, n;\n\t"
        "mov.b128 {%0, %1}, d;\n\t"
        "}\n"
        : "=l"(o0), "=l"(o1) : "l"(s), "l"(0ull), "l"(n0), "l"(n1) : "memory");
}

// Adds `inc` occurrences of key (k0, k1) to the global table.
// kPlain: the job runs with flags == 0, so no test switch has to be looked at.
// Out of line, with the table handed over by value: through a reference to the kernel's parameters the function
// would re-read table and mask with generic loads on every probe and use generic atomics (it cannot see that the
// pointers are global memory); the inlined wrapper reads them from the constant bank at the call.
template <bool kPlain>
__device__ __noinline__ void global_insert_at(Slot *table, uint64_t table_mask, Counters *ctr, uint32_t flags, uint64_t k0, uint64_t k1,
                                              unsigned long long inc) {
    __builtin_assume(__isGlobal(table));
    __builtin_assume(__isGlobal(ctr));
#if LOCUST_SLOT32
    uint64_t h = slot_hash32(k0, k1);
#else
    uint64_t h = hash_key128(k0, k1);
#endif
    if (!kPlain && (flags & LOCUST_F_WEAK_HASH)) h &= 0xFFull;
    uint64_t slot = h & table_mask;
    for (uint32_t probe = 0; probe < kMaxProbe; ++probe) {
        Slot *s = table + slot;
        const ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2 *>(s));
        if (cur.x == k0 && cur.y == k1) { atomicAdd(&s->count, inc); return; }
        if ((cur.x | cur.y) == 0ull) {
            uint64_t o0, o1;
            cas128(s, k0, k1, o0, o1);
            if ((o0 | o1) == 0ull || (o0 == k0 && o1 == k1)) { atomicAdd(&s->count, inc); return; }
        }
        slot = (slot + 1) & table_mask;
    }
    atomicOr(&ctr->status, kStatusTableFull);
}
template <bool kPlain>
LC_D void global_insert(const MapParams &p, uint64_t k0, uint64_t k1, unsigned long long inc) {
    global_insert_at<kPlain>(p.table, p.table_mask, p.ctr, p.flags, k0, k1, inc);
}

// Per-warp view of shared memory, kept as 32-bit offsets into the CTA's dynamic shared memory.
extern __shared__ __align__(16) uint8_t smem_raw[];
struct WarpCtx {
    uint32_t area;         // byte offset of this warp's area
    uint32_t qn;           // entries queued (warp-uniform)
    uint32_t lane;
    uint32_t sb;           // shared address of smem_raw
    bool use_smem;
};
#define WC_SKEYS reinterpret_cast<ulonglong2 *>(smem_raw)
#define WC_SCNT reinterpret_cast<uint32_t *>(smem_raw + kKeysBytes)
#define WC_ROW(wc) (smem_raw + (wc).area + kRowOff)
#define WC_BITS(wc) reinterpret_cast<uint32_t *>(smem_raw + (wc).area + kBitsOff)
#define WC_DESC(wc) reinterpret_cast<uint16_t *>(smem_raw + (wc).area + kDescOff)
#define WC_QUEUE(wc) reinterpret_cast<ulonglong2 *>(smem_raw + (wc).area + kQueueOff)

// The CTA's table counts tokens of <= 8 bytes (key = lo | hi << 32, never 0); a bucket is two
// keys side by side, so one 128-bit load reads both ways.
LC_D uint32_t smem_bucket(uint32_t lo, uint32_t hi) {
    return __umulhi(lo * 0x9E3779B1u + hi * 0x85EBCA77u, kBuckets);
}
// Fast probe used per token: counts the token if its key already sits in the bucket, without
// a divergent branch.  A miss is queued by the caller; the queue's drain claims free ways.
// Branch free, on 32-bit shared addresses (wc.sb = the CTA's shared window, kept in a register: the
// generic-pointer form recomputed it from a special register for every batch): a predicated 128-bit load of
// the bucket, two compares, one predicated atomic (a key sits in one way only).  `probe` says whether this lane has a key of <= 8 bytes.
#ifndef LOCUST_ASM_PROBE
#define LOCUST_ASM_PROBE 1
#endif
#if !LOCUST_ASM_PROBE
LC_D bool smem_hit(const WarpCtx &wc, uint32_t lo, uint32_t hi, bool probe) {
    if (!probe) return false;
    const uint32_t b = smem_bucket(lo, hi);
    const uint4 kb = reinterpret_cast<const uint4 *>(WC_SKEYS)[b];
    const uint32_t d0 = (kb.x ^ lo) | (kb.y ^ hi), d1 = (kb.z ^ lo) | (kb.w ^ hi);
    const bool hit = d0 == 0u || d1 == 0u;
    if (hit) atomicAdd(&WC_SCNT[2 * b + (d1 == 0u ? 1u : 0u)], 1u);
    return hit;
}
#else
LC_D bool smem_hit(const WarpCtx &wc, uint32_t lo, uint32_t hi, bool probe) {
    uint32_t hit;
    asm volatile(
        "{\n\t"
        ".reg .pred ps, p0, p1;\n\t"
        ".reg .u32 h, ad, k0, k1, k2, k3, x0, x1;\n\t"
        "setp.ne.u32 ps, %4, 0;\n\t"
        "mul.lo.u32 h, %1, 0x9E3779B1;\n\t"
        "mad.lo.u32 h, %2, 0x85EBCA77, h;\n\t"
        "mul.hi.u32 h, h, %5;\n\t"
        "mad.lo.u32 ad, h, 16, %3;\n\t"
        "mov.u32 k0, 0;\n\tmov.u32 k1, 0;\n\tmov.u32 k2, 0;\n\tmov.u32 k3, 0;\n\t"
        "@ps ld.shared.v4.u32 {k0, k1, k2, k3}, [ad];\n\t"
        "xor.b32 x0, k0, %1;\n\txor.b32 x1, k1, %2;\n\tor.b32 x0, x0, x1;\n\t"
        "setp.eq.and.u32 p0, x0, 0, ps;\n\t"
        "xor.b32 x0, k2, %1;\n\txor.b32 x1, k3, %2;\n\tor.b32 x0, x0, x1;\n\t"
        "setp.eq.and.u32 p1, x0, 0, ps;\n\t"
        "mad.lo.u32 ad, h, 8, %3;\n\t"
        "selp.u32 x0, 4, 0, p1;\n\t"
        "add.u32 ad, ad, x0;\n\t"
        "or.pred p0, p0, p1;\n\t"
        "@p0 red.shared.add.u32 [ad+%6], 1;\n\t"          // a key sits in one way only: one atomic serves both
        "selp.u32 %0, 1, 0, p0;\n\t"
        "}"
        : "=r"(hit) : "r"(lo), "r"(hi), "r"(wc.sb), "r"((uint32_t)probe), "n"(kBuckets), "n"(kKeysBytes) : "memory");
    return hit != 0u;
}
#endif
// Full insert: counts `inc` occurrences, claiming a free way if the key is not there yet.  False
// if the bucket is taken by two other keys (or a claim was lost): the key goes to the global table.
LC_D bool smem_insert(const WarpCtx &wc, uint32_t lo, uint32_t hi, uint32_t inc) {
    const uint32_t b = smem_bucket(lo, hi);
    const unsigned long long k = (unsigned long long)lo | ((unsigned long long)hi << 32);
    const ulonglong2 kb = WC_SKEYS[b];
    if (kb.x == k) { atomicAdd(&WC_SCNT[2 * b], inc); return true; }
    if (kb.y == k) { atomicAdd(&WC_SCNT[2 * b + 1], inc); return true; }
    if (kb.x == 0ull || kb.y == 0ull) {
        const uint32_t way = kb.x == 0ull ? 0u : 1u;
        unsigned long long *slot = reinterpret_cast<unsigned long long *>(WC_SKEYS) + 2 * b + way;
        const unsigned long long old = atomicCAS(slot, 0ull, k);
        if (old == 0ull || old == k) { atomicAdd(&WC_SCNT[2 * b + way], inc); return true; }
    }
    return false;
}

// Byte-serial path for one token starting at base[pos]: tokens longer than 16 bytes
// and the LOCUST_F_GENERIC_TOKENISER test switch.
template <bool kPlain>
__device__ __noinline__ void token_generic(const MapParams &p, const WarpCtx &wc, uint64_t pos) {
    const uint8_t *s = p.base + pos;
    const uint64_t maxlen = p.hi - pos;
    uint64_t k0 = 0, k1 = 0, L = 0;
    while (L < 16 && L < maxlen) {
        const uint32_t c = s[L];
        if (is_delim(c)) break;
        if (L < 8) k0 |= (uint64_t)c << (8 * L); else k1 |= (uint64_t)c << (8 * (L - 8));
        ++L;
    }
    if (L == 16 && L < maxlen && !is_delim(s[16])) {
        long_token_at(p, s, maxlen);
        return;
    }
    if (L <= 8 && (kPlain || wc.use_smem)) {
        if (smem_insert(wc, (uint32_t)k0, (uint32_t)(k0 >> 32), 1u)) return;
    }
    global_insert<kPlain>(p, k0, k1, 1ull);
}

// Chunk at byte offset `off` that straddles lo or hi: bytes outside [lo, hi) read as '\n'.
__device__ __noinline__ uint4 mask_chunk(const MapParams &p, uint64_t off, uint4 v) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int b = 0; b < 16; ++b) {
        const uint64_t q = off + b;
        if (q < p.lo || q >= p.hi)
            w[b >> 2] = (w[b >> 2] & ~(0xFFu << (8 * (b & 3)))) | (0x0Au << (8 * (b & 3)));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// 16 bytes of row `row` for this lane; bytes outside [lo, hi) read as '\n'.
LC_D uint4 load_chunk(const MapParams &p, uint64_t row, uint32_t lane) {
    const uint64_t off = row * 512ull + lane * 16u;
    uint4 v = make_uint4(0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au);
    if (off < p.hi) {
        v = ldg_stream(p.base + off);
        if (off < p.lo || off + 16 > p.hi) v = mask_chunk(p, off, v);
    }
    return v;
}

template <bool kPlain>
LC_D void drain_queue(const MapParams &p, WarpCtx &wc, uint32_t keep_below) {
    // Pops batches of up to 32 keys until fewer than keep_below remain.  Equal keys of a
    // batch are added up first and inserted once: a heavy-hitter word of 9..16 bytes would
    // otherwise hammer one global counter from every SM (same-address atomics serialise
    // in L2; BASELINE config 5 is built to show exactly that).
    while (wc.qn >= keep_below && wc.qn > 0u) {
        __syncwarp();
        const uint32_t take = wc.qn < 32u ? wc.qn : 32u;
        const bool valid = wc.lane < take;
        ulonglong2 k = make_ulonglong2(0ull, 0ull);
        if (valid) k = WC_QUEUE(wc)[wc.qn - take + wc.lane];
        const uint32_t h = (uint32_t)k.x * 0x9E3779B1u + (uint32_t)(k.x >> 32) * 0x85EBCA77u +
                           (uint32_t)k.y * 0xC2B2AE3Du + (uint32_t)(k.y >> 32) * 0x27D4EB2Fu;
        const uint32_t peers = __match_any_sync(kFull, h);
        const uint32_t leader = (uint32_t)__ffs((int)peers) - 1u;
        const unsigned long long lx = __shfl_sync(kFull, k.x, leader), ly = __shfl_sync(kFull, k.y, leader);
        const bool same = k.x == lx && k.y == ly;                  // equal hash, equal key (else: insert alone)
        const uint32_t agg = __ballot_sync(kFull, same) & peers;
        // one lane per distinct key carries the sum; short keys get a chance to claim a free way
        // of the CTA's table first (the only place ways are claimed on the fast path)
        const uint32_t inc = !same ? 1u : (wc.lane == leader ? (uint32_t)__popc(agg) : 0u);
        if (valid && inc) {
            const bool to_smem = k.y == 0ull && (kPlain || wc.use_smem) &&
                                 smem_insert(wc, (uint32_t)k.x, (uint32_t)(k.x >> 32), inc);
            if (!to_smem) global_insert<kPlain>(p, k.x, k.y, (unsigned long long)inc);
        }
        wc.qn -= take;
        __syncwarp();
    }
}

// kDepth batches of 32 queued keys at once (key j of this lane in k[j], zero = none), for the inserting warps:
// the same steps as drain_queue, but the home slots of everything bound for the global table are read together,
// so that the L2 round trips of up to kDepth batches overlap.
template <int kDepth>
LC_D void insert_keys(const MapParams &p, const WarpCtx &wc, const ulonglong2 (&k)[kDepth]) {
    uint32_t inc[kDepth];
    Slot *home[kDepth];
    ulonglong2 cur[kDepth];
#pragma unroll
    for (int j = 0; j < kDepth; ++j) {
        const uint32_t h = (uint32_t)k[j].x * 0x9E3779B1u + (uint32_t)(k[j].x >> 32) * 0x85EBCA77u +
                           (uint32_t)k[j].y * 0xC2B2AE3Du + (uint32_t)(k[j].y >> 32) * 0x27D4EB2Fu;
        const uint32_t peers = __match_any_sync(kFull, h);
        const uint32_t leader = (uint32_t)__ffs((int)peers) - 1u;
        const unsigned long long lx = __shfl_sync(kFull, k[j].x, leader), ly = __shfl_sync(kFull, k[j].y, leader);
        const bool same = k[j].x == lx && k[j].y == ly;
        const uint32_t agg = __ballot_sync(kFull, same) & peers;
        inc[j] = !same ? 1u : (wc.lane == leader ? (uint32_t)__popc(agg) : 0u);
        if ((k[j].x | k[j].y) == 0ull) inc[j] = 0u;
        if (inc[j] && k[j].y == 0ull && smem_insert(wc, (uint32_t)k[j].x, (uint32_t)(k[j].x >> 32), inc[j])) inc[j] = 0u;
        home[j] = p.table + ((uint64_t)slot_hash32(k[j].x, k[j].y) & p.table_mask);
        cur[j] = make_ulonglong2(0ull, 0ull);
        if (inc[j]) cur[j] = __ldcg(reinterpret_cast<const ulonglong2 *>(home[j]));
    }
#pragma unroll
    for (int j = 0; j < kDepth; ++j) {
        if (!inc[j]) continue;
        uint64_t o0 = cur[j].x, o1 = cur[j].y;
        if ((o0 | o1) == 0ull) {
            cas128(home[j], k[j].x, k[j].y, o0, o1);
            if ((o0 | o1) == 0ull) { o0 = k[j].x; o1 = k[j].y; }
        }
        if (o0 == k[j].x && o1 == k[j].y) atomicAdd(&home[j]->count, (unsigned long long)inc[j]);
        else global_insert<true>(p, k[j].x, k[j].y, (unsigned long long)inc[j]);
    }
}

// The top 32 queued keys of a tokenising warp go to an inserting warp; when the ring is full the warp inserts
// them itself.  A place in the ring is taken by compare-and-swap on the head while fewer than kXSlots batches are
// outstanding (the barriers' phase parity cannot tell use k of a slot from use k + 2).
LC_D void hand_over(const MapParams &p, WarpCtx &wc) {
    const uint32_t xb = wc.sb + kXOff;
    while (wc.qn >= 32u) {
        __syncwarp();
        uint32_t idx = 0, got = 0;
        if (wc.lane == 0) {
            idx = lds32(xb + oXHead);
            const uint32_t tail = lds32(xb + oXTail);
            if (idx - tail < kXSlots) got = atoms_cas32(xb + oXHead, idx, idx + 1u) == idx ? 1u : 0u;
        }
        got = __shfl_sync(kFull, got, 0);
        idx = __shfl_sync(kFull, idx, 0);
        if (!got) { drain_queue<true>(p, wc, wc.qn - 31u); continue; }          // one batch, here
        const uint32_t s = idx % kXSlots, round = idx / kXSlots;
        mbar_wait(xb + oXEmpty + 8 * s, (round & 1u) ^ 1u);
        const uint4 kk = lds128(wc.sb + wc.area + kQueueOff + 16u * (wc.qn - 32u + wc.lane));
        sts128(xb + oXData + s * 512u + wc.lane * 16u, kk.x, kk.y, kk.z, kk.w);
        if (wc.lane == 0) sts32(xb + oXCount + 4 * s, 32u);
        __syncwarp();
        if (wc.lane == 0) mbar_arrive(xb + oXFull + 8 * s);
        wc.qn -= 32u;
    }
}

// An inserting warp: batches idx = w, w + kInsWarps, ... in order, up to four that are ready at a time.
LC_D void inserter_loop(const MapParams &p, const WarpCtx &wc, uint32_t w) {
    const uint32_t xb = wc.sb + kXOff;
#ifndef LOCUST_INS_DEPTH
#define LOCUST_INS_DEPTH 4
#endif
    constexpr int kDepth = LOCUST_INS_DEPTH;
    bool stop = false;
    for (uint32_t idx = w; !stop;) {
        ulonglong2 k[kDepth];
        uint32_t taken = 0;
#pragma unroll
        for (int j = 0; j < kDepth; ++j) {
            k[j] = make_ulonglong2(0ull, 0ull);
            if (stop || (j > 0 && taken != (uint32_t)j)) continue;              // batches are taken in order, without gaps
            const uint32_t s = idx % kXSlots, round = idx / kXSlots;
            if (j == 0) mbar_wait(xb + oXFull + 8 * s, round & 1u);
            else if (!__shfl_sync(kFull, (uint32_t)mbar_test(xb + oXFull + 8 * s, round & 1u), 0)) continue;   // not handed over yet: next time
            if (lds32(xb + oXCount + 4 * s) == kPoison) { stop = true; continue; }
            k[j] = lds128q(xb + oXData + s * 512u + wc.lane * 16u);
            __syncwarp();
            if (wc.lane == 0) { mbar_arrive(xb + oXEmpty + 8 * s); atoms_add(xb + oXTail, 1u); }
            idx += (uint32_t)kInsWarps;
            ++taken;
        }
        if (taken) insert_keys<kDepth>(p, wc, k);
    }
}

// Tokens of more than 16 bytes, out of line: their addresses and the list's parameters are only formed here.
// Called by the lanes that hold one (converged among themselves); they take their places in the list of far
// tokens with one atomic between them.  Inside a function that is not inlined the compiler no longer knows that
// the counters and the list are global memory, and would use generic atomics and stores: it is told.
template <bool kPlain>
__device__ __noinline__ void far_token(const MapParams &p, const WarpCtx &wc, uint64_t row_byte, uint32_t pos) {
    if (kPlain) {
        Counters *ctr = p.ctr;
        ulonglong2 *list = p.far_list;
        __builtin_assume(__isGlobal(ctr));
        __builtin_assume(__isGlobal(list));
        const uint32_t m = __activemask();
        const uint32_t leader = (uint32_t)__ffs((int)m) - 1u;
        unsigned long long first = 0;
        if (wc.lane == leader) first = atomicAdd(&ctr->n_far, (unsigned long long)__popc(m));
        first = __shfl_sync(m, first, leader);
        const unsigned long long idx = first + __popc(m & ((1u << wc.lane) - 1u));
        if (idx < p.far_cap)           // walked by long_tokens_kernel behind this launch
            list[idx] = make_ulonglong2(reinterpret_cast<unsigned long long>(p.base + row_byte + pos), reinterpret_cast<unsigned long long>(p.base + p.hi));
        else atomicOr(&ctr->status, kStatusFarFull);
    } else {
        token_generic<kPlain>(p, wc, row_byte + pos);
    }
}

// Phase B for one batch: lane `wc.lane` takes token number t of the row (if active).
// All lanes call this together.  The row is number r0 + i counted from p.base (only looked at for a token that runs on).
template <bool kPlain, bool kAscii>
LC_D void process_batch(const MapParams &p, WarpCtx &wc, bool active, uint32_t t, uint64_t r0, uint32_t i) {
    const uint32_t *row32 = reinterpret_cast<const uint32_t *>(WC_ROW(wc));
    const uint32_t pos = active ? WC_DESC(wc)[t] : 0u;
    // token length: first delimiter at or after pos, within 17 bits
    const uint32_t wi = pos >> 5;
    const uint32_t win = __funnelshift_r(WC_BITS(wc)[wi], WC_BITS(wc)[wi + 1], pos & 31u) & 0x1FFFFu;
    const uint32_t L = (uint32_t)__ffs((int)win) - 1u;            // win == 0 -> 0xFFFFFFFF
    const bool has = active && win != 0u;
    const bool far = active && win == 0u;
    // token bytes: five aligned words starting at the word that holds pos
    const uint32_t bi = pos >> 2, o = pos & 3u;
    const uint32_t w0 = row32[bi], w1 = row32[bi + 1], w2 = row32[bi + 2], w3 = row32[bi + 3], w4 = row32[bi + 4];
    uint32_t a, b, c, d;
    if (kAscii) extract_key16_ascii(w0, w1, w2, w3, w4, o, L, a, b, c, d);
    else extract_key16_any(w0, w1, w2, w3, w4, o, L, a, b, c, d);
    if (!kPlain && (p.flags & LOCUST_F_SORT_PATH)) {          // alternative path: one record per token, no combiner
        const uint32_t m = __ballot_sync(kFull, has);
        unsigned long long base = 0;
        if (wc.lane == 0 && m) base = atomicAdd(&p.ctr->n_entries, (unsigned long long)__popc(m));
        base = __shfl_sync(kFull, base, 0);
        if (has) {
            const unsigned long long e = base + __popc(m & ((1u << wc.lane) - 1u));
            if (e < p.rec.cap) {
                Rec rec;
                rec.be0 = bswap64((uint64_t)a | ((uint64_t)b << 32));
                rec.be1 = bswap64((uint64_t)c | ((uint64_t)d << 32));
                rec.count = 1ull; rec.len = L; rec.lidx = 0xFFFFFFFFu;
                store_rec(p.rec.a + e, rec);
            } else {
                atomicOr(&p.ctr->status, kStatusEntryFull);
            }
        }
        if (__any_sync(kFull, far)) {
            if (far) far_token<kPlain>(p, wc, (r0 + i) * 512ull, pos);
            __syncwarp();
        }
        return;
    }
    const bool hit = smem_hit(wc, a, b, has && L <= 8u && (kPlain || wc.use_smem));
    const bool defer = has && !hit;
    const uint32_t m = __ballot_sync(kFull, defer);
    if (m) {
        if (defer)
            WC_QUEUE(wc)[wc.qn + __popc(m & ((1u << wc.lane) - 1u))] =
                make_ulonglong2((uint64_t)a | ((uint64_t)b << 32), (uint64_t)c | ((uint64_t)d << 32));
        wc.qn += __popc(m);
        if (wc.qn >= 32u) {
            if (kPlain && kInsWarps > 0) hand_over(p, wc);
            else drain_queue<kPlain>(p, wc, 32u);
        }
    }
    if (far) far_token<kPlain>(p, wc, (r0 + i) * 512ull, pos);       // rare; nothing the other lanes have to wait for
}

template <bool kPlain>
__global__ void __launch_bounds__(kMapThreads, 1) tokenize_combine_kernel(const __grid_constant__ MapParams p) {
    {
        uint4 *z = reinterpret_cast<uint4 *>(smem_raw);
        for (uint32_t i = threadIdx.x; i < (kKeysBytes + kCntBytes + oXData) / 16; i += kMapThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    constexpr bool kSplit = kPlain && kInsWarps > 0;
    constexpr uint32_t kTokWarps = kSplit ? kMapWarps - kInsWarps : kMapWarps;
    if (kSplit) {
        if (threadIdx.x == 0) {
            const uint32_t xb = (uint32_t)__cvta_generic_to_shared(smem_raw) + kXOff;
            for (uint32_t s = 0; s < kXSlots; ++s) { mbar_init(xb + oXFull + 8 * s, 1); mbar_init(xb + oXEmpty + 8 * s, 1); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }

    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint64_t nwarps = (uint64_t)gridDim.x * kTokWarps;
    const uint64_t gw = (uint64_t)blockIdx.x * kTokWarps + warp;
    // rows of 512 bytes counted from base; only rows that overlap [lo, hi) are handed out
    const uint64_t row_lo = p.lo >> 9, nrows = ((p.hi + 511ull) >> 9) - row_lo;
    // Every warp gets kSegments contiguous pieces spread evenly over the input instead of one:
    // where token density varies along the file (BASELINE config 5) each warp still sees an
    // average mix, and pieces stay long enough that their start-up cost does not matter.
    const uint64_t seg_len = (nrows + nwarps * kSegments - 1) / (nwarps * kSegments);
    const bool use_fast = kPlain || !(p.flags & LOCUST_F_GENERIC_TOKENISER);
    WarpCtx wc;
    wc.area = kAreasOff + warp * kWarpArea;
    wc.qn = 0u;
    wc.lane = lane;
    wc.use_smem = kPlain || !(p.flags & LOCUST_F_NO_SMEM_COMBINER);
    wc.sb = (uint32_t)__cvta_generic_to_shared(smem_raw);
    asm volatile("mov.u32 %0, %0;" : "+r"(wc.sb));            // kept in a register
    const uint32_t one = opaque_one();
    uint16_t *bits16 = reinterpret_cast<uint16_t *>(WC_BITS(wc));
    if (kSplit && warp >= kTokWarps) {
        inserter_loop(p, wc, warp - kTokWarps);
    } else {

    for (uint32_t seg = 0; seg < kSegments; ++seg) {
        const uint64_t r0 = row_lo + (seg * nwarps + gw) * seg_len;
        uint64_t r1 = r0 + seg_len;
        if (r1 > row_lo + nrows) r1 = row_lo + nrows;
        if (r0 >= r1) break;
        const uint32_t n_rows = (uint32_t)(r1 - r0);
        // Rows counted from r0 from here on.  Rows [f0, f0 + flen) lie entirely inside [lo, hi) and
        // are loaded without any bounds arithmetic; the few others (the first row of an unaligned
        // range, the rows around hi) go through load_chunk.
        uint32_t f0, flen;
        {
            const uint64_t full_lo = (p.lo + 511ull) >> 9, full_hi = p.hi >> 9, lim = (uint64_t)n_rows + 2ull;
            const uint64_t a = full_lo > r0 ? (full_lo - r0 < lim ? full_lo - r0 : lim) : 0ull;
            const uint64_t b = full_hi > r0 ? (full_hi - r0 < lim ? full_hi - r0 : lim) : 0ull;
            f0 = (uint32_t)a;
            flen = b > a ? (uint32_t)(b - a) : 0u;
        }
        const uint8_t *q = p.base + r0 * 512ull + lane * 16u;              // this lane's bytes of row r0
        auto load_row = [&](uint32_t i) -> uint4 {
            if (i - f0 < flen) return ldg_stream(q + (uint64_t)i * 512ull);
            return load_chunk(p, r0 + i, lane);
        };
        uint4 c = load_row(0u);
        uint4 n = load_row(1u);
        // inside the loop only rows 2.. are loaded, and the one row that can start before lo is row 0 or 1 (f0 <= 1):
        // one bound is enough there
        const uint32_t f_end = f0 + flen;
        // dc carries two rows' masks: this row's in the low half, the previous row's above it (only the top bit of
        // lane 31's is ever looked at: the byte in front of the row) -- one register instead of two
        uint32_t dc = classify16_bal<kImadAdds>(c.x, c.y, c.z, c.w, one);
        uint32_t dn = classify16_bal<kImadAdds>(n.x, n.y, n.z, n.w, one);
        bool asc_c = !__any_sync(kFull, (c.x | c.y | c.z | c.w) & 0x80808080u);
        bool asc_n = !__any_sync(kFull, (n.x | n.y | n.z | n.w) & 0x80808080u);
        {
            const uint64_t b = r0 * 512ull;
            dc |= (b <= p.lo || is_delim(p.base[b - 1])) ? 0x80000000u : 0u;
        }
        for (uint32_t i = 0; i < n_rows; ++i) {
            // in flight while this row is tokenised
            const uint4 nn = i + 2u < f_end ? ldg_stream(q + (uint64_t)(i + 2u) * 512ull) : load_chunk(p, r0 + (uint64_t)(i + 2u), lane);
            // mask of the chunk in front (the source lane is taken modulo 32: lane 0 hears from lane 31)
            const uint32_t d_dn = __shfl_sync(kFull, lane == 31 ? dc >> 16 : dc, lane - 1u);
            uint32_t starts = ~dc & ((dc << 1) | ((d_dn >> 15) & 1u)) & 0xFFFFu;
            if (use_fast) {
                // A: stage bytes and delimiter bits (with the next row's first chunks as look-ahead)
                reinterpret_cast<uint4 *>(WC_ROW(wc))[lane] = c;
                if (lane < 2) reinterpret_cast<uint4 *>(WC_ROW(wc))[32 + lane] = n;
                bits16[lane] = (uint16_t)dc;
                if (lane < 4) bits16[32 + lane] = (uint16_t)dn;
                // dense list of token starts: warp prefix sum of the per-lane counts
                const uint32_t cnt = __popc(starts);
                uint32_t incl = cnt;
#if LOCUST_SCANSTEP
                incl = scan_step(incl, 1); incl = scan_step(incl, 2); incl = scan_step(incl, 4);
                incl = scan_step(incl, 8); incl = scan_step(incl, 16);
#else
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(kFull, incl, o);
                    if (lane >= (uint32_t)o) incl += y;
                }
#endif
                const uint32_t total = __shfl_sync(kFull, incl, 31);
                uint32_t at = incl;                                         // filled from the back: order within a row is free
#if LOCUST_LISTUNROLL > 0
                {
                    // predicated steps instead of a loop whose trip count differs from lane to lane
                    const uint32_t dl = (uint32_t)__cvta_generic_to_shared(WC_DESC(wc)) + 2u * incl;
                    const uint32_t l16 = lane * 16u;
#pragma unroll
                    for (int k = 0; k < LOCUST_LISTUNROLL; ++k) {
                        uint32_t top;
                        asm("bfind.u32 %0, %1;" : "=r"(top) : "r"(starts));
                        if (starts) sts16(dl - 2u * (uint32_t)(k + 1), l16 + top);
                        starts ^= shl_clamp(1u, top);              // top = 0xFFFFFFFF when nothing is left: no bit
                    }
                    at -= LOCUST_LISTUNROLL;
                }
                if (__any_sync(kFull, starts != 0u))
#endif
                while (starts) {
                    const uint32_t top = 31u - (uint32_t)__clz((int)starts);
                    WC_DESC(wc)[--at] = (uint16_t)(lane * 16u + top);
                    starts ^= 1u << top;
                }
                __syncwarp();
                // B: one token per lane, 32 at a time; the text is pure ASCII or not for a whole row: two loops
                if (asc_c && asc_n) {
#pragma unroll 1
                    for (uint32_t t0 = 0; t0 < total; t0 += 32) process_batch<kPlain, true>(p, wc, t0 + lane < total, t0 + lane, r0, i);
                } else {
#pragma unroll 1
                    for (uint32_t t0 = 0; t0 < total; t0 += 32) process_batch<kPlain, false>(p, wc, t0 + lane < total, t0 + lane, r0, i);
                }
                __syncwarp();
            } else {
                while (starts) {
                    const uint32_t pp = (uint32_t)__ffs((int)starts) - 1u;
                    starts &= starts - 1u;
                    token_generic<kPlain>(p, wc, (r0 + i) * 512ull + lane * 16u + pp);
                }
                __syncwarp();
            }
            c = n; dc = (dc << 16) | dn; asc_c = asc_n;
            n = nn;
            dn = classify16_bal<kImadAdds>(n.x, n.y, n.z, n.w, one);
            asc_n = !__any_sync(kFull, (n.x | n.y | n.z | n.w) & 0x80808080u);
        }
    }
    drain_queue<kPlain>(p, wc, 1u);
    if (kSplit) {
        // the last tokenising warp to finish tells every inserting warp to stop
        const uint32_t xb = wc.sb + kXOff;
        uint32_t d = 0;
        if (lane == 0) d = atoms_add_ret(xb + oXDone, 1u);
        d = __shfl_sync(kFull, d, 0);
        if (d == kTokWarps - 1u) {
            for (int j = 0; j < kInsWarps; ++j) {
                uint32_t idx = 0;
                if (lane == 0) idx = atoms_add_ret(xb + oXHead, 1u);
                idx = __shfl_sync(kFull, idx, 0);
                const uint32_t s = idx % kXSlots, round = idx / kXSlots;
                mbar_wait(xb + oXEmpty + 8 * s, (round & 1u) ^ 1u);
                if (lane == 0) { sts32(xb + oXCount + 4 * s, kPoison); mbar_arrive(xb + oXFull + 8 * s); }
                __syncwarp();
            }
        }
    }
    }   // tokenising warp
    __syncthreads();
    // Flush the CTA's combiner into the global table.  All 148 CTAs arrive here at about the same
    // time with mostly the same keys, and a thread has a dozen slots to flush: done one after the
    // other, every slot costs two or three L2 round trips (probe, claim, add).  So four slots go
    // together: their home slots are read at once, the empty ones are claimed at once, and only a
    // slot that turns out to be taken by another key takes the ordinary probing path.
    const unsigned long long *keys = reinterpret_cast<const unsigned long long *>(WC_SKEYS);
    constexpr uint32_t kFlush = 4;
    for (uint32_t base = threadIdx.x; base < kSmemSlots; base += kMapThreads * kFlush) {
        unsigned long long k[kFlush];
        uint64_t o0[kFlush], o1[kFlush];
        Slot *home[kFlush];
        ulonglong2 cur[kFlush];
#pragma unroll
        for (uint32_t j = 0; j < kFlush; ++j) {
            const uint32_t i = base + j * kMapThreads;
            k[j] = i < kSmemSlots ? keys[i] : 0ull;
#if LOCUST_SLOT32
            uint64_t h = slot_hash32(k[j], 0ull);
#else
            uint64_t h = hash_key128(k[j], 0ull);
#endif
            if (!kPlain && (p.flags & LOCUST_F_WEAK_HASH)) h &= 0xFFull;
            home[j] = p.table + (h & p.table_mask);
            cur[j] = make_ulonglong2(0ull, 0ull);
            if (k[j]) cur[j] = __ldcg(reinterpret_cast<const ulonglong2 *>(home[j]));
        }
#pragma unroll
        for (uint32_t j = 0; j < kFlush; ++j) {
            o0[j] = cur[j].x; o1[j] = cur[j].y;
            if (k[j] && (cur[j].x | cur[j].y) == 0ull) cas128(home[j], k[j], 0ull, o0[j], o1[j]);     // returns what was there
        }
#pragma unroll
        for (uint32_t j = 0; j < kFlush; ++j) {
            if (!k[j]) continue;
            const unsigned long long cnt = (unsigned long long)WC_SCNT[base + j * kMapThreads];
            const bool claimed = (cur[j].x | cur[j].y) == 0ull && (o0[j] | o1[j]) == 0ull;        // our CAS went through
            if (claimed || (o0[j] == k[j] && o1[j] == 0ull)) atomicAdd(&home[j]->count, cnt);
            else global_insert<kPlain>(p, k[j], 0ull, cnt);
        }
    }
}

// ---- merge side of the shuffle (loadIntermediateFile's role, main.cu:66-103) ---------
// One received part: header {n_short, long_bytes, long_off, 0}, short records {k0, k1,
// count, 0} from byte 32, long records {len, count, bytes padded to 16} from byte long_off.  Short records go through the same
// table insert as MAP; long records are appended to the long-token list, pointing
// into the received buffer.
__global__ void __launch_bounds__(256) shuffle_merge_kernel(const uint8_t *__restrict__ part, uint64_t nbytes,
                                                            const __grid_constant__ MapParams p) {
    const unsigned long long *hdr = reinterpret_cast<const unsigned long long *>(part);
    const unsigned long long n_short = hdr[0], long_bytes = hdr[1], long_off = hdr[2];
    if (32ull + 32ull * n_short > long_off || long_off + long_bytes > nbytes) {     // malformed part: refuse loudly
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(&p.ctr->status, kStatusEntryFull | kStatusBadPart);
        return;
    }
    const ulonglong4 *rec = reinterpret_cast<const ulonglong4 *>(part + 32);
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_short;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const ulonglong4 v = ld256(rec + i);
        global_insert<false>(p, v.x, v.y, v.z);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {                      // long records are rare: one thread walks them
        const uint8_t *q = part + long_off;
        unsigned long long done = 0;
        while (done < long_bytes) {
            const unsigned long long len = reinterpret_cast<const unsigned long long *>(q + done)[0];
            const unsigned long long cnt = reinterpret_cast<const unsigned long long *>(q + done)[1];
            long_insert(p, q + done + 16, len, long_hash(q + done + 16, len), cnt);             // the same key from several senders becomes one entry
            done += 16ull + ((len + 15ull) & ~15ull);
        }
    }
}

int launch_shuffle_merge(const uint8_t *part, uint64_t nbytes, MapParams p, int num_sms, cudaStream_t st) {
    if (nbytes < 32) return 0;
    uint64_t g = (nbytes / 32 + 255) / 256;
    const uint64_t cap = (uint64_t)num_sms * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    shuffle_merge_kernel<<<(unsigned)g, 256, 0, st>>>(part, nbytes, p);
    return 1;
}

cudaError_t map_kernel_setup() {
    cudaError_t e = cudaFuncSetAttribute(tokenize_combine_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tokenize_combine_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return e;
    return map_stream_setup();
}

// LOCUST_MAP_VARIANT (development switch): -1 (default) = the register-staged kernel of this file;
// 0.. = a configuration of the streaming kernel (map_stream_kernel.cu: bulk-async row ring, inserter
// warps), which is correct on every test but measured slower.
static int map_variant() {
    static const int v = [] { const char *s = getenv("LOCUST_MAP_VARIANT"); return s ? atoi(s) : -1; }();
    return v;
}

int launch_tokenize_combine(const MapParams &p, int num_sms, cudaStream_t st) {
    if (p.hi <= p.lo) return 0;
    if (p.flags == 0u && map_variant() >= 0) return launch_tokenize_stream(p, map_variant(), num_sms, st);
    const uint64_t nrows = ((p.hi + 511ull) >> 9) - (p.lo >> 9);
    const uint64_t per_cta = p.flags == 0u ? (uint64_t)(kMapWarps - kInsWarps) : (uint64_t)kMapWarps;
    uint64_t ctas = (nrows + per_cta - 1) / per_cta;
    if (ctas > (uint64_t)num_sms) ctas = (uint64_t)num_sms;
    if (p.flags == 0u) tokenize_combine_kernel<true><<<(unsigned)ctas, kMapThreads, kSmemBytes, st>>>(p);
    else tokenize_combine_kernel<false><<<(unsigned)ctas, kMapThreads, kSmemBytes, st>>>(p);
    return 1;
}
}  // namespace locust
