// map_stream_kernel.cu -- MAP stage, production path (flags == 0): a streaming tokeniser fed by
// bulk asynchronous copies, fused with the combiner, with the global-table traffic moved to
// dedicated warps.
//
// Stands in for kernMap/map (MapReduce/src/main.cu:136-159), my_strtok_r (util.cu:54-89) and,
// through the combiner, for most of thrust::sort + kernFindUniqBool/kernGetCount (main.cu:415,
// :161-238).  map_kernel.cu keeps the older register-staged kernel for the test switches
// (LOCUST_F_*); this file is what a job with flags == 0 runs.
//
// One persistent CTA per SM, two kinds of warps:
//   * kTok tokeniser warps.  Each owns a contiguous range of 1024-byte rows and a ring of three
//     row slots in shared memory.  Lane 0 fetches rows with cp.async.bulk (one 1056-byte bulk
//     copy per row: the row plus the first 32 bytes of the next, so that a token crossing the
//     row end is contiguous in the slot); an mbarrier per slot says when the bytes have landed.
//     Per row:
//       A. every lane reads two 16-byte chunks of the *next* row (chunk l and chunk 32 + l, both
//          conflict free), classifies them with the SWAR classifier (common.cuh; part of the
//          adds issued as IMAD so the FMA pipe shares the load), and keeps the 2 x 16 delimiter
//          bits in a register;
//          for the *current* row the lane derives its token starts, looks the token length up in
//          a 32-bit delimiter window (own bits + the next chunk's), and writes one 16-bit entry
//          {position, length} per token into a dense, position-ordered list (packed warp scan);
//       B. the list is walked 64 tokens at a time, two tokens per lane (two independent chains):
//          key bytes cut out of the row slot with PRMT, keys of <= 8 bytes counted in the CTA's
//          2-way shared-memory table, misses and keys of 9..16 bytes pushed to a per-warp queue.
//     When 32 keys are queued they are handed to an inserter warp through a ring in shared
//     memory (mbarrier full/empty pairs); when the ring is full the tokeniser inserts them itself.
//   * kIns inserter warps.  They take batches of queued keys, add up equal keys, let short keys
//     claim a free way of the shared table, and put the rest into the global table (128-bit CAS,
//     64-bit RED).  The L2 round trips of those probes stall only these warps.
// At the end the CTA's table is flushed into the global table.
#include <type_traits>

#include "kernels.cuh"
#include "long_keys.cuh"
#include "smem_ptx.cuh"
#include "../../include/locust.h"

namespace locust {
namespace {

constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr uint32_t kRow = 1024;               // bytes per row
constexpr uint32_t kSlot = kRow + 32;         // row + look-ahead
constexpr uint32_t kRing = 3;                 // slots per tokeniser warp: tokenised / classified / in flight
constexpr uint32_t kListBytes = 1088;         // <= 512 token starts in 1024 bytes, 2 bytes each, + 32 of padding
constexpr uint32_t kQueue = 96;               // queued keys per warp, 16 bytes each (31 left over + 64 new)
constexpr uint32_t oRing = 0, oBar = kRing * kSlot, oList = oBar + 32, oQueue = oList + kListBytes;
constexpr uint32_t kWarpArea = oQueue + kQueue * 16;
static_assert(kWarpArea % 16 == 0 && oBar % 8 == 0 && oQueue % 16 == 0, "alignment");
// hand-over ring (CTA): control words, full/empty barriers, entry counts, key batches
constexpr uint32_t kXSlots = 8;
constexpr uint32_t oXHead = 0, oXTail = 4, oXDone = 8, oXFull = 16, oXEmpty = oXFull + 8 * kXSlots,
                   oXCount = oXEmpty + 8 * kXSlots, oXData = 256;
constexpr uint32_t kXBytes = oXData + kXSlots * 512;
static_assert(oXCount + 4 * kXSlots <= oXData, "control block");
constexpr uint32_t kSmemBudget = 232448;
constexpr uint32_t kMaxProbe = 1u << 14;
constexpr uint32_t kPoison = 0xFFFFFFFFu;

template <int kTok>
struct Layout {
    static constexpr uint32_t kBuckets = (kSmemBudget - kXBytes - kTok * kWarpArea) / 24u / 16u * 16u;
    static constexpr uint32_t kSlots = 2 * kBuckets;
    static constexpr uint32_t kKeysBytes = kBuckets * 16, kCntBytes = kSlots * 4;
    static constexpr uint32_t oKeys = 0, oCnt = kKeysBytes, oX = oCnt + kCntBytes, oWarps = oX + kXBytes;
    static constexpr uint32_t kBytes = oWarps + kTok * kWarpArea;
    static_assert(oX % 128 == 0 && oWarps % 128 == 0, "alignment");
    static_assert(kBytes <= kSmemBudget, "shared memory budget of one CTA");
};

LC_D void cas128(Slot *s, uint64_t n0, uint64_t n1, uint64_t &o0, uint64_t &o1) {
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

// ---- the CTA's combiner table (keys of <= 8 bytes): buckets of two 64-bit keys side by side ----
template <int kTok>
struct Table {
    using L = Layout<kTok>;
    uint32_t keys, cnt;        // shared addresses
    LC_D uint32_t bucket(uint32_t lo, uint32_t hi) const { return __umulhi(lo * 0x9E3779B1u + hi * 0x85EBCA77u, L::kBuckets); }
    // counts `inc` occurrences, claiming a free way if the key is new; false if both ways belong to other keys
    LC_D bool insert(uint32_t lo, uint32_t hi, uint32_t inc) const {
        const uint32_t b = bucket(lo, hi);
        const unsigned long long k = (unsigned long long)lo | ((unsigned long long)hi << 32);
        const ulonglong2 kb = lds128q(keys + b * 16u);
        if (kb.x == k) { atoms_add(cnt + b * 8u, inc); return true; }
        if (kb.y == k) { atoms_add(cnt + b * 8u + 4u, inc); return true; }
        if (kb.x == 0ull || kb.y == 0ull) {
            const uint32_t way = kb.x == 0ull ? 0u : 1u;
            const unsigned long long old = atoms_cas64(keys + b * 16u + way * 8u, 0ull, k);
            if (old == 0ull || old == k) { atoms_add(cnt + b * 8u + way * 4u, inc); return true; }
        }
        return false;
    }
};

// Adds `inc` occurrences of key (k0, k1) to the global table, starting at slot `slot`.
__device__ __noinline__ void global_probe(const MapParams &p, uint64_t slot, uint64_t k0, uint64_t k1, unsigned long long inc) {
    const uint64_t lim = p.table_mask + 1ull < (uint64_t)kMaxProbe ? p.table_mask + 1ull : (uint64_t)kMaxProbe;
    for (uint64_t probe = 0; probe < lim; ++probe) {
        Slot *s = p.table + slot;
        const ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2 *>(s));
        if (cur.x == k0 && cur.y == k1) { atomicAdd(&s->count, inc); return; }
        if ((cur.x | cur.y) == 0ull) {
            uint64_t o0, o1;
            cas128(s, k0, k1, o0, o1);
            if ((o0 | o1) == 0ull || (o0 == k0 && o1 == k1)) { atomicAdd(&s->count, inc); return; }
        }
        slot = (slot + 1) & p.table_mask;
        // a table that is already known to be full: the job will be redone with a larger one
        if ((probe & 63u) == 63u && (*reinterpret_cast<volatile uint32_t *>(&p.ctr->status) & kStatusTableFull)) return;
    }
    atomicOr(&p.ctr->status, kStatusTableFull);
}

// Up to kDepth batches of 32 queued keys at once (key j of this lane in k[j], zero = none): equal
// keys of a batch are added up first (a heavy-hitter word of 9..16 bytes would otherwise hammer
// one L2 counter from every SM -- BASELINE config 5), short keys try to claim a way of the CTA's
// table, and the home slots of everything else are read together so that the L2 round trips
// overlap; only keys whose home slot belongs to another key take the probing path.
template <int kTok, int kDepth>
LC_D void insert_batches(const MapParams &p, const Table<kTok> &tab, const ulonglong2 (&k)[kDepth], uint32_t lane) {
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
        const bool same = k[j].x == lx && k[j].y == ly;             // equal hash, equal key (else: insert alone)
        const uint32_t agg = __ballot_sync(kFull, same) & peers;
        inc[j] = !same ? 1u : (lane == leader ? (uint32_t)__popc(agg) : 0u);
        if ((k[j].x | k[j].y) == 0ull) inc[j] = 0u;
        if (inc[j] && k[j].y == 0ull && tab.insert((uint32_t)k[j].x, (uint32_t)(k[j].x >> 32), inc[j])) inc[j] = 0u;
        home[j] = p.table + ((uint64_t)slot_hash32(k[j].x, k[j].y) & p.table_mask);
        cur[j] = make_ulonglong2(0ull, 0ull);
        if (inc[j]) cur[j] = __ldcg(reinterpret_cast<const ulonglong2 *>(home[j]));
    }
#pragma unroll
    for (int j = 0; j < kDepth; ++j) {
        if (!inc[j]) continue;
        uint64_t o0 = cur[j].x, o1 = cur[j].y;
        if ((o0 | o1) == 0ull) {
            cas128(home[j], k[j].x, k[j].y, o0, o1);               // returns what was there
            if ((o0 | o1) == 0ull) { o0 = k[j].x; o1 = k[j].y; }
        }
        if (o0 == k[j].x && o1 == k[j].y) atomicAdd(&home[j]->count, (unsigned long long)inc[j]);
        else global_probe(p, (uint64_t)((home[j] - p.table) + 1) & p.table_mask, k[j].x, k[j].y, (unsigned long long)inc[j]);
    }
}

// 16 bytes at byte offset `off` from base that straddle lo or hi: bytes outside [lo, hi) read as '\n'
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

// One token of the list: entry e = position in the row | length << 10 (length 1..16; 17..32: far; an
// entry of 0 is padding).  Cuts the key out of the row slot (five aligned words around the position,
// PRMT) and, for keys of <= 8 bytes, counts it in the CTA's table if it is there.  Returns 1 when
// the key has to be queued (a miss, or 9..16 bytes).  The probe is branch free: a predicated 128-bit
// load of the bucket, two compares, two predicated atomics.
template <int kTok, bool kAscii>
LC_D uint32_t take_token(const Table<kTok> &tab, uint32_t e, uint32_t slot, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    const uint32_t wa = slot + (e & 0x3FCu), o = e & 3u, L = e >> 10;
    const uint32_t w0 = lds32(wa), w1 = lds32(wa + 4), w2 = lds32(wa + 8), w3 = lds32(wa + 12), w4 = lds32(wa + 16);
    if (kAscii) extract_key16_ascii(w0, w1, w2, w3, w4, o, L, a, b, c, d);
    else extract_key16_any(w0, w1, w2, w3, w4, o, L, a, b, c, d);
    uint32_t defer;
    asm volatile(
        "{\n\t"
        ".reg .pred ps, ph, p0, p1;\n\t"
        ".reg .u32 t, h, ad, k0, k1, k2, k3, x0, x1;\n\t"
        "sub.u32 t, %1, 0x400;\n\t"
        "setp.lt.u32 ps, t, 0x2000;\n\t"                 // length 1..8: the CTA's table
        "setp.lt.u32 ph, t, 0x4000;\n\t"                 // length 1..16: a key at all
        "mul.lo.u32 h, %2, 0x9E3779B1;\n\t"
        "mad.lo.u32 h, %3, 0x85EBCA77, h;\n\t"
        "mul.hi.u32 h, h, %6;\n\t"                       // bucket
        "mad.lo.u32 ad, h, 16, %4;\n\t"
        "mov.u32 k0, 0;\n\tmov.u32 k1, 0;\n\tmov.u32 k2, 0;\n\tmov.u32 k3, 0;\n\t"
        "@ps ld.shared.v4.u32 {k0, k1, k2, k3}, [ad];\n\t"
        "xor.b32 x0, k0, %2;\n\txor.b32 x1, k1, %3;\n\tor.b32 x0, x0, x1;\n\t"
        "setp.eq.and.u32 p0, x0, 0, ps;\n\t"
        "xor.b32 x0, k2, %2;\n\txor.b32 x1, k3, %3;\n\tor.b32 x0, x0, x1;\n\t"
        "setp.eq.and.u32 p1, x0, 0, ps;\n\t"
        "mad.lo.u32 ad, h, 8, %5;\n\t"
        "@p0 red.shared.add.u32 [ad], 1;\n\t"
        "@p1 red.shared.add.u32 [ad+4], 1;\n\t"
        "or.pred p0, p0, p1;\n\t"
        "and.pred p0, ph, !p0;\n\t"
        "selp.u32 %0, 1, 0, p0;\n\t"
        "}"
        : "=r"(defer) : "r"(e), "r"(a), "r"(b), "r"(tab.keys), "r"(tab.cnt), "n"(Layout<kTok>::kBuckets) : "memory");
    return defer;
}

template <int kTok, int kIns, int kImad>
__global__ void __launch_bounds__((kTok + kIns) * 32, 1) tokenize_stream_kernel(const __grid_constant__ MapParams p) {
    using L = Layout<kTok>;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem_raw);
    asm volatile("mov.u32 %0, %0;" : "+r"(sbase));            // kept in a register: ptxas otherwise recomputes it per use
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    constexpr uint32_t kThreads = (kTok + kIns) * 32;
    Table<kTok> tab;
    tab.keys = sbase + L::oKeys;
    tab.cnt = sbase + L::oCnt;
    const uint32_t xb = sbase + L::oX;
    {
        uint4 *z = reinterpret_cast<uint4 *>(smem_raw);
        for (uint32_t i = threadIdx.x; i < (L::kKeysBytes + L::kCntBytes + oXData) / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < kXSlots; ++s) { mbar_init(xb + oXFull + 8 * s, 1); mbar_init(xb + oXEmpty + 8 * s, 1); }
    }
    if (warp < kTok && lane == 0) {
        for (uint32_t s = 0; s < kRing; ++s) mbar_init(sbase + L::oWarps + warp * kWarpArea + oBar + 8 * s, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
    __syncthreads();

    if (warp < kTok) {
        // =========================== tokeniser warp ===========================================
        const uint32_t wa = sbase + L::oWarps + warp * kWarpArea;
        const uint32_t one = p.flags + 1u;                    // == 1 on this path; opaque to the compiler
        const uint64_t nwarps = (uint64_t)gridDim.x * kTok;
        const uint64_t gw = (uint64_t)blockIdx.x * kTok + warp;
        const uint64_t row_lo = p.lo >> 10, nrows = ((p.hi + 1023ull) >> 10) - row_lo;
        const uint64_t seg_len = (nrows + nwarps - 1) / nwarps;
        const uint64_t r0 = row_lo + gw * seg_len;
        uint64_t r1 = r0 + seg_len;
        if (r1 > row_lo + nrows) r1 = row_lo + nrows;
        uint32_t qn = 0;                                      // queued keys (warp-uniform)
        const uint32_t lt = (1u << lane) - 1u;
        if (r0 < r1) {
            const uint32_t n_rows = (uint32_t)(r1 - r0);
            const uint64_t lo16 = p.lo & ~15ull, hi16 = (p.hi + 15ull) & ~15ull;
            // fetch of row r0 + i into slot i % 3 (rows 0..n_rows: the last one only as look-ahead)
            auto fetch = [&](uint32_t i) {
                const uint64_t b0 = (r0 + i) << 10;
                const uint64_t s = b0 > lo16 ? b0 : lo16;
                const uint64_t e = b0 + kSlot < hi16 ? b0 + kSlot : hi16;
                if (s < e && lane == 0 && i <= n_rows) {
                    const uint32_t sl = i % kRing;
                    const uint32_t bar = wa + oBar + 8 * sl;
                    fence_proxy_async();
                    mbar_expect_tx(bar, (uint32_t)(e - s));
                    bulk_g2s(wa + oRing + sl * kSlot + (uint32_t)(s - b0), p.base + s, (uint32_t)(e - s), bar);
                }
            };
            auto fetched = [&](uint32_t i) -> bool {       // i <= n_rows always
                const uint64_t b0 = (r0 + i) << 10;
                const uint64_t s = b0 > lo16 ? b0 : lo16;
                const uint64_t e = b0 + kSlot < hi16 ? b0 + kSlot : hi16;
                return s < e;
            };
            uint32_t phase = 0;                               // bit s: parity the next wait on slot s uses
            // delimiter bits of row r0 + i, chunk l in the low half, chunk 32 + l in the high half; sets asc_n
            bool asc_n = true;
            auto classify_row = [&](uint32_t i) -> uint32_t {
                if (!fetched(i)) { asc_n = true; return kFull; }
                const uint32_t sl = i % kRing;
                mbar_wait(wa + oBar + 8 * sl, (phase >> sl) & 1u);
                phase ^= 1u << sl;
                const uint32_t ra = wa + oRing + sl * kSlot + lane * 16u;
                uint4 c0 = lds128(ra), c1 = lds128(ra + 512u);
                const uint64_t b0 = (r0 + i) << 10;
                if (b0 < p.lo || b0 + kRow > p.hi) {          // first or last row of the input: bytes outside read as '\n'
                    c0 = mask_chunk(p, b0 + lane * 16u, c0);
                    c1 = mask_chunk(p, b0 + 512u + lane * 16u, c1);
                }
                asc_n = !__any_sync(kFull, (c0.x | c0.y | c0.z | c0.w | c1.x | c1.y | c1.z | c1.w) & 0x80808080u);
                return classify16_bal<kImad>(c0.x, c0.y, c0.z, c0.w, one) | (classify16_bal<kImad>(c1.x, c1.y, c1.z, c1.w, one) << 16);
            };
            fetch(0); fetch(1);
            uint32_t dc = classify_row(0);
            bool asc_c = asc_n;
            // lane 0 carries the delimiter bits of the chunk in front of the row in `pd` (bit 31 = last byte)
            uint32_t pd;
            {
                const uint64_t b = r0 << 10;
                pd = (b <= p.lo || is_delim(p.base[b - 1])) ? 0x80000000u : 0u;
            }
            for (uint32_t i = 0; i < n_rows; ++i) {
                fetch(i + 2u);
                const uint32_t dn = classify_row(i + 1u);                 // next row: its bits are this row's look-ahead
                const uint32_t slot = wa + oRing + (i % kRing) * kSlot;
                // --- A: token starts, lengths, dense ordered list ---------------------------------
                const uint32_t xr = __shfl_sync(kFull, dc, (lane + 31u) & 31u);   // chunk in front of each of this lane's chunks
                uint32_t x = xr;
                if (lane == 0) x = (xr << 16) | ((pd >> 16) & 0x8000u);
                pd = xr;                                                   // lane 0: the row's last chunk
                const uint32_t prev = ((dc << 1) & 0xFFFEFFFEu) | ((x >> 15) & 0x00010001u);
                const uint32_t starts = ~dc & prev;
                uint32_t y = __shfl_sync(kFull, dc, (lane + 1u) & 31u);    // chunk behind each of this lane's chunks
                const uint32_t dn0 = __shfl_sync(kFull, dn, 0);
                if (lane == 31) y = prmt(y, dn0, 0x5432u);
                // delimiter windows (own 16 bits + the next 16), shifted so that bit b belongs to byte b + 1, with a
                // sentinel on top (a token that starts in the chunk and has no delimiter in the window is "far"),
                // bit-reversed: the distance from a start bit b to the next delimiter is clz(reversed << b)
                const uint32_t wrA = __brev((prmt(dc, y, 0x5410u) >> 1) | 0x80000000u);
                const uint32_t wrB = __brev((prmt(dc, y, 0x7632u) >> 1) | 0x80000000u);
                uint32_t sA = starts & 0xFFFFu, sB = starts >> 16;
                uint32_t v = (uint32_t)__popc(sA) | ((uint32_t)__popc(sB) << 16);
                v = scan_step(v, 1); v = scan_step(v, 2); v = scan_step(v, 4); v = scan_step(v, 8); v = scan_step(v, 16);
                const uint32_t tot = __shfl_sync(kFull, v, 31);
                const uint32_t totA = tot & 0xFFFFu, total = totA + (tot >> 16);
                uint32_t ia = wa + oList + 2u * (v & 0xFFFFu), ib = wa + oList + 2u * (totA + (v >> 16));
                // entry = position | length << 10; 0x400: length = distance + 1
                const uint32_t pA = lane * 16u + 0x400u, pB = pA + 512u;
                if (sA | sB) {
                    // both chunks' starts side by side, highest first, branch free inside the loop
                    asm volatile(
                        "{\n\t"
                        ".reg .pred pa, pb, pl;\n\t"
                        ".reg .u32 ba, bb, ta, tb, ca, cb, ea, eb, ma, mb, u;\n\t"
                        "LIST_LOOP:\n\t"
                        "setp.ne.u32 pa, %0, 0;\n\t"
                        "setp.ne.u32 pb, %1, 0;\n\t"
                        "bfind.u32 ba, %0;\n\t"
                        "bfind.u32 bb, %1;\n\t"
                        "shl.b32 ta, %4, ba;\n\t"
                        "shl.b32 tb, %5, bb;\n\t"
                        "clz.b32 ca, ta;\n\t"
                        "clz.b32 cb, tb;\n\t"
                        "add.u32 ea, %6, ba;\n\t"
                        "add.u32 eb, %7, bb;\n\t"
                        "mad.lo.u32 ea, ca, 1024, ea;\n\t"
                        "mad.lo.u32 eb, cb, 1024, eb;\n\t"
                        "sub.u32 %2, %2, 2;\n\t"
                        "sub.u32 %3, %3, 2;\n\t"
                        "@pa st.shared.u16 [%2], ea;\n\t"
                        "@pb st.shared.u16 [%3], eb;\n\t"
                        "mov.u32 u, 1;\n\t"
                        "shl.b32 ma, u, ba;\n\t"
                        "shl.b32 mb, u, bb;\n\t"
                        "xor.b32 %0, %0, ma;\n\t"
                        "xor.b32 %1, %1, mb;\n\t"
                        "or.b32 u, %0, %1;\n\t"
                        "setp.ne.u32 pl, u, 0;\n\t"
                        "@pl bra LIST_LOOP;\n\t"
                        "}"
                        : "+r"(sA), "+r"(sB), "+r"(ia), "+r"(ib) : "r"(wrA), "r"(wrB), "r"(pA), "r"(pB) : "memory");
                }
                sts16(wa + oList + 2u * (total + lane), 0u);                // padding: the walk reads whole batches
                __syncwarp();
                // --- B: two tokens per lane ------------------------------------------------------
                const uint64_t row_pos = (r0 + i) << 10;
                auto hand_over = [&]() {
                    // the top 32 queued keys go to an inserter warp, or are inserted here when the ring is full
                    while (qn >= 32u) {
                        __syncwarp();
                        const uint32_t src = wa + oQueue + 16u * (qn - 32u) + 16u * lane;
                        // a place in the ring is taken only while fewer than kXSlots batches are outstanding (compare-and-swap on
                        // the head: the barriers' phase parity cannot tell use k of a slot from use k + 2)
                        uint32_t idx = 0, got = 0;
                        if (kIns > 0) {
                            if (lane == 0) {
                                idx = lds32(xb + oXHead);
                                const uint32_t tail = lds32(xb + oXTail);
                                if (idx - tail < kXSlots) got = atoms_cas32(xb + oXHead, idx, idx + 1u) == idx ? 1u : 0u;
                            }
                            got = __shfl_sync(kFull, got, 0);
                            idx = __shfl_sync(kFull, idx, 0);
                        }
                        if (got) {
                            const uint32_t s = idx % kXSlots, round = idx / kXSlots;
                            mbar_wait(xb + oXEmpty + 8 * s, (round & 1u) ^ 1u);
                            const uint4 kk = lds128(src);
                            sts128(xb + oXData + s * 512u + lane * 16u, kk.x, kk.y, kk.z, kk.w);
                            if (lane == 0) sts32(xb + oXCount + 4 * s, 32u);
                            __syncwarp();
                            if (lane == 0) mbar_arrive(xb + oXFull + 8 * s);
                        } else {
                            ulonglong2 kk[1];
                            kk[0] = lds128q(src);
                            insert_batches<kTok, 1>(p, tab, kk, lane);
                        }
                        qn -= 32u;
                    }
                };
                auto walk = [&](auto ascii_tag) {
                    constexpr bool kAscii = decltype(ascii_tag)::value;
                    uint32_t la = wa + oList + 2u * lane;
                    uint32_t t0 = 0;
#pragma unroll 1
                    for (; t0 + 32u < total; t0 += 64u, la += 128u) {
                        const uint32_t e0 = lds16(la), e1 = lds16(la + 64u);
                        uint32_t a0, b0, c0, d0, a1, b1, c1, d1;
                        const uint32_t f0 = take_token<kTok, kAscii>(tab, e0, slot, a0, b0, c0, d0);
                        const uint32_t f1 = take_token<kTok, kAscii>(tab, e1, slot, a1, b1, c1, d1);
                        const uint32_t m0 = __ballot_sync(kFull, f0 != 0u), m1 = __ballot_sync(kFull, f1 != 0u);
                        const uint32_t q = wa + oQueue + 16u * qn, n0 = (uint32_t)__popc(m0);
                        if (f0) sts128(q + 16u * (uint32_t)__popc(m0 & lt), a0, b0, c0, d0);
                        if (f1) sts128(q + 16u * (n0 + (uint32_t)__popc(m1 & lt)), a1, b1, c1, d1);
                        qn += n0 + (uint32_t)__popc(m1);
                        if (qn >= 32u) hand_over();
                        if (__any_sync(kFull, (e0 | e1) >= 0x4400u)) {
                            if (e0 >= 0x4400u) far_push(p, row_pos + (e0 & 0x3FFu));
                            if (e1 >= 0x4400u) far_push(p, row_pos + (e1 & 0x3FFu));
                            __syncwarp();
                        }
                    }
                    if (t0 < total) {
                        const uint32_t e0 = lds16(la);
                        uint32_t a0, b0, c0, d0;
                        const uint32_t f0 = take_token<kTok, kAscii>(tab, e0, slot, a0, b0, c0, d0);
                        const uint32_t m0 = __ballot_sync(kFull, f0 != 0u);
                        if (f0) sts128(wa + oQueue + 16u * (qn + (uint32_t)__popc(m0 & lt)), a0, b0, c0, d0);
                        qn += (uint32_t)__popc(m0);
                        if (qn >= 32u) hand_over();
                        if (__any_sync(kFull, e0 >= 0x4400u)) {
                            if (e0 >= 0x4400u) far_push(p, row_pos + (e0 & 0x3FFu));
                            __syncwarp();
                        }
                    }
                };
                if (asc_c && asc_n) walk(std::true_type{}); else walk(std::false_type{});
                __syncwarp();
                dc = dn; asc_c = asc_n;
            }
        }
        // what is left in the queue
        __syncwarp();
        if (qn) {
            ulonglong2 kk[1];
            kk[0] = lane < qn ? lds128q(wa + oQueue + 16u * lane) : make_ulonglong2(0ull, 0ull);
            insert_batches<kTok, 1>(p, tab, kk, lane);
        }
        // the last tokeniser warp to finish tells every inserter warp to stop
        if (kIns > 0) {
            uint32_t d = 0;
            if (lane == 0) d = atoms_add_ret(xb + oXDone, 1u);
            d = __shfl_sync(kFull, d, 0);
            if (d == kTok - 1) {
                for (int j = 0; j < kIns; ++j) {
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
    } else if (kIns > 0) {
        // =========================== inserter warp ============================================
        const uint32_t w = warp - kTok;
        for (uint32_t idx = w;; idx += (uint32_t)(kIns > 0 ? kIns : 1)) {
            const uint32_t s = idx % kXSlots, round = idx / kXSlots;
            mbar_wait(xb + oXFull + 8 * s, round & 1u);
            const uint32_t n = lds32(xb + oXCount + 4 * s);
            if (n == kPoison) break;
            ulonglong2 kk[1];
            kk[0] = lds128q(xb + oXData + s * 512u + lane * 16u);
            __syncwarp();
            if (lane == 0) { mbar_arrive(xb + oXEmpty + 8 * s); atoms_add(xb + oXTail, 1u); }
            insert_batches<kTok, 1>(p, tab, kk, lane);
        }
    }
    __syncthreads();
    // Flush the CTA's combiner into the global table, four slots at a time: home slots read together,
    // empty ones claimed together; only a slot taken by another key takes the probing path.
    constexpr uint32_t kFlush = 4;
    for (uint32_t base = threadIdx.x; base < L::kSlots; base += kThreads * kFlush) {
        unsigned long long k[kFlush];
        uint64_t o0[kFlush], o1[kFlush];
        Slot *home[kFlush];
        ulonglong2 cur[kFlush];
#pragma unroll
        for (uint32_t j = 0; j < kFlush; ++j) {
            const uint32_t i = base + j * kThreads;
            k[j] = i < L::kSlots ? lds64(tab.keys + 8u * i) : 0ull;
            home[j] = p.table + ((uint64_t)slot_hash32(k[j], 0ull) & p.table_mask);
            cur[j] = make_ulonglong2(0ull, 0ull);
            if (k[j]) cur[j] = __ldcg(reinterpret_cast<const ulonglong2 *>(home[j]));
        }
#pragma unroll
        for (uint32_t j = 0; j < kFlush; ++j) {
            o0[j] = cur[j].x; o1[j] = cur[j].y;
            if (k[j] && (cur[j].x | cur[j].y) == 0ull) cas128(home[j], k[j], 0ull, o0[j], o1[j]);
        }
#pragma unroll
        for (uint32_t j = 0; j < kFlush; ++j) {
            if (!k[j]) continue;
            const unsigned long long cnt = (unsigned long long)lds32(tab.cnt + 4u * (base + j * kThreads));
            const bool claimed = (cur[j].x | cur[j].y) == 0ull && (o0[j] | o1[j]) == 0ull;
            if (claimed || (o0[j] == k[j] && o1[j] == 0ull)) atomicAdd(&home[j]->count, cnt);
            else global_probe(p, (uint64_t)((home[j] - p.table) + 1) & p.table_mask, k[j], 0ull, cnt);
        }
    }
}

template <int kTok, int kIns, int kImad>
cudaError_t setup_one() {
    return cudaFuncSetAttribute(tokenize_stream_kernel<kTok, kIns, kImad>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)Layout<kTok>::kBytes);
}
template <int kTok, int kIns, int kImad>
void launch_one(const MapParams &p, int num_sms, cudaStream_t st) {
    const uint64_t nrows = ((p.hi + 1023ull) >> 10) - (p.lo >> 10);
    uint64_t ctas = (nrows + kTok - 1) / kTok;
    if (ctas > (uint64_t)num_sms) ctas = (uint64_t)num_sms;
    tokenize_stream_kernel<kTok, kIns, kImad><<<(unsigned)ctas, (kTok + kIns) * 32, Layout<kTok>::kBytes, st>>>(p);
}

}  // namespace

cudaError_t map_stream_setup() {
    cudaError_t e = setup_one<14, 2, 4>();
    if (e == cudaSuccess) e = setup_one<12, 4, 4>();
    if (e == cudaSuccess) e = setup_one<14, 2, 0>();
    if (e == cudaSuccess) e = setup_one<16, 0, 4>();
    return e;
}

int launch_tokenize_stream(const MapParams &p, int variant, int num_sms, cudaStream_t st) {
    if (p.hi <= p.lo) return 0;
    switch (variant) {
    case 1: launch_one<12, 4, 4>(p, num_sms, st); break;
    case 2: launch_one<14, 2, 0>(p, num_sms, st); break;
    case 3: launch_one<16, 0, 4>(p, num_sms, st); break;
    default: launch_one<14, 2, 4>(p, num_sms, st); break;
    }
    return 1;
}

}  // namespace locust
