// map_kernel.cu -- MAP stage: streaming tokeniser fused with the combiner.
//
// Stands in for kernMap/map (MapReduce/src/main.cu:136-159) and, through the
// combiner, for most of the work of thrust::sort + kernFindUniqBool/kernGetCount
// (main.cu:415, :161-238): instead of emitting one 40-byte record per token and
// sorting them all, every token is counted into a hash table keyed by its bytes.
//
// Layout.  One persistent CTA of 32 warps per SM.  A warp owns a contiguous range
// of 512-byte rows; each lane holds 16 bytes of the row in registers (one
// coalesced 128-bit load) and the next two rows are already in flight or
// classified.  Word boundaries come from a SWAR delimiter classifier
// (common.cuh) and neighbour-lane shuffles; token bytes are cut out of the
// registers with PRMT, never re-read from memory.  Tokens of up to 8 bytes are
// counted in a 16 K-slot shared-memory table (exact 64-bit keys), everything else
// in the global table (exact 128-bit keys, 128-bit CAS).  Tokens longer than 16
// bytes are listed by address for the process stage.
#include "kernels.cuh"
#include "../../include/locust.h"

namespace locust {

constexpr int kLogSmemSlots = 14;
constexpr uint32_t kSmemSlots = 1u << kLogSmemSlots;
constexpr uint32_t kQueueSlots = 64;             // per warp, 16 bytes each
constexpr int kMapThreads = 1024;
constexpr uint32_t kSmemBytes = kSmemSlots * (8 + 4) + (kMapThreads / 32) * kQueueSlots * 16;
constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr uint32_t kMaxProbe = 1u << 14;

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
__device__ __noinline__ void global_insert(const MapParams &p, uint64_t k0, uint64_t k1, unsigned long long inc) {
    uint64_t h = hash_key128(k0, k1);
    if (p.flags & LOCUST_F_WEAK_HASH) h &= 0xFFull;
    uint64_t slot = h & p.table_mask;
    for (uint32_t probe = 0; probe < kMaxProbe; ++probe) {
        Slot *s = p.table + slot;
        const ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2 *>(s));
        if (cur.x == k0 && cur.y == k1) { atomicAdd(&s->count, inc); return; }
        if ((cur.x | cur.y) == 0ull) {
            uint64_t o0, o1;
            cas128(s, k0, k1, o0, o1);
            if ((o0 | o1) == 0ull || (o0 == k0 && o1 == k1)) { atomicAdd(&s->count, inc); return; }
        }
        slot = (slot + 1) & p.table_mask;
    }
    atomicOr(&p.ctr->status, kStatusTableFull);
}

// Counts a token of <= 8 bytes in the CTA's shared table; false if both probed
// slots belong to other keys.
LC_D bool smem_insert(uint64_t *skeys, uint32_t *scnt, uint32_t lo, uint32_t hi) {
    const uint32_t h = lo * 0x9E3779B1u + hi * 0x85EBCA77u;
    uint32_t slot = h >> (32 - kLogSmemSlots);
    const uint64_t k = (uint64_t)lo | ((uint64_t)hi << 32);
#pragma unroll
    for (int probe = 0; probe < 2; ++probe) {
        const uint64_t cur = skeys[slot];
        if (cur == k) { atomicAdd(&scnt[slot], 1u); return true; }
        if (cur == 0ull) {
            const unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long *>(&skeys[slot]), 0ull, k);
            if (old == 0ull || old == k) { atomicAdd(&scnt[slot], 1u); return true; }
        }
        slot = (slot + 1) & (kSmemSlots - 1);
    }
    return false;
}

// Byte-serial path for one token starting at base[pos]: tokens that leave the
// register window, second tokens of a word, rows with bytes >= 0x80.
__device__ __noinline__ void token_generic(const MapParams &p, uint64_t *skeys, uint32_t *scnt, uint64_t pos) {
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
        L = 17;
        while (L < maxlen && !is_delim(s[L])) ++L;
        const unsigned long long idx = atomicAdd(&p.ctr->n_long, 1ull);
        if (idx < p.long_cap) { p.long_list[idx].ptr = s; p.long_list[idx].len = L; }
        else atomicOr(&p.ctr->status, kStatusLongFull);
        return;
    }
    if (L <= 8 && !(p.flags & LOCUST_F_NO_SMEM_COMBINER)) {
        if (smem_insert(skeys, scnt, (uint32_t)k0, (uint32_t)(k0 >> 32))) return;
    }
    global_insert(p, k0, k1, 1ull);
}

// 16 bytes of row `row` for this lane; bytes outside [lo, hi) read as '\n'.
LC_D uint4 load_chunk(const MapParams &p, uint64_t row, uint32_t lane) {
    const uint64_t off = row * 512ull + lane * 16u;
    uint4 v = make_uint4(0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au);
    if (off < p.hi) {
        v = ldg_stream(p.base + off);
        if (off < p.lo || off + 16 > p.hi) {
            uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int b = 0; b < 16; ++b) {
                const uint64_t q = off + b;
                if (q < p.lo || q >= p.hi)
                    w[b >> 2] = (w[b >> 2] & ~(0xFFu << (8 * (b & 3)))) | (0x0Au << (8 * (b & 3)));
            }
            v = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
    return v;
}

// Per-warp state of the combiner: the CTA's shared table and this warp's queue of
// keys that have to go to the global table.  Deferring them costs a few
// instructions per step; inserting them one batch of 32 at a time keeps every lane
// busy in the long global-insert code instead of one or two.
struct WarpCtx {
    uint64_t *skeys;
    uint32_t *scnt;
    ulonglong2 *queue;     // kQueueSlots entries, this warp's
    uint32_t qn;           // entries queued (warp-uniform)
    uint32_t lane;
    bool use_smem;
};

LC_D void drain_queue(const MapParams &p, WarpCtx &wc, uint32_t keep_below) {
    // pops batches of up to 32 keys until fewer than keep_below remain
    while (wc.qn >= keep_below && wc.qn > 0u) {
        __syncwarp();
        const uint32_t take = wc.qn < 32u ? wc.qn : 32u;
        if (wc.lane < take) {
            const ulonglong2 k = wc.queue[wc.qn - take + wc.lane];
            global_insert(p, k.x, k.y, 1ull);
        }
        wc.qn -= take;
        __syncwarp();
    }
}

// All lanes call this together.  `has` marks the lanes that hold a token of L <= 16
// bytes whose bytes start at byte `o` of w0.  Short tokens are counted in the shared
// table; misses and tokens of 9..16 bytes are queued for the global table.
template <bool ASCII>
LC_D void handle_token(const MapParams &p, WarpCtx &wc, bool has, uint32_t o, uint32_t L, uint32_t w0, uint32_t w1,
                       uint32_t w2, uint32_t w3, uint32_t w4) {
    uint32_t a = 0, b = 0, c = 0, d = 0;
    bool defer = false;
    if (has) {
        if (ASCII) extract_key16_ascii(w0, w1, w2, w3, w4, o, L, a, b, c, d);
        else extract_key16_any(w0, w1, w2, w3, w4, o, L, a, b, c, d);
        defer = !(L <= 8u && wc.use_smem && smem_insert(wc.skeys, wc.scnt, a, b));
    }
    const uint32_t m = __ballot_sync(kFull, defer);
    if (m) {
        if (defer)
            wc.queue[wc.qn + __popc(m & ((1u << wc.lane) - 1u))] =
                make_ulonglong2((uint64_t)a | ((uint64_t)b << 32), (uint64_t)c | ((uint64_t)d << 32));
        wc.qn += __popc(m);
        if (wc.qn >= 32u) drain_queue(p, wc, 32u);
    }
}

template <bool ASCII>
LC_D void tokenise_row(const MapParams &p, WarpCtx &wc, const uint32_t (&w)[8], uint32_t starts, uint32_t dd,
                       uint64_t chunk_pos) {
    uint32_t rem = starts;
    // first token of each 4-byte word: register indices are static
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t sj = (starts >> (4 * j)) & 0xFu;
        const uint32_t o = low_bit4(sj);
        const uint32_t pp = 4 * j + o;
        const uint32_t L = (uint32_t)__ffs((int)(dd >> pp)) - 1u;            // no delimiter in the window -> huge
        const bool has = sj != 0u && L <= 16u;
        if (has) rem &= ~(1u << pp);
        handle_token<ASCII>(p, wc, has, o, L, w[j], w[j + 1], w[j + 2], w[j + 3], w[j + 4]);
    }
    // what is left: second tokens of a word (dynamic word index) and tokens that leave the window
    while (__any_sync(kFull, rem != 0u)) {
        const uint32_t pp = rem ? (uint32_t)__ffs((int)rem) - 1u : 0u;
        const uint32_t L = (uint32_t)__ffs((int)(dd >> pp)) - 1u;
        const bool has = rem != 0u && L <= 16u;
        const bool far = rem != 0u && L > 16u;
        rem &= rem - 1u;
        const uint32_t j = pp >> 2;
        const uint32_t t0 = (j & 1u) ? w[1] : w[0], t1 = (j & 1u) ? w[2] : w[1], t2 = (j & 1u) ? w[3] : w[2],
                       t3 = (j & 1u) ? w[4] : w[3], t4 = (j & 1u) ? w[5] : w[4], t5 = (j & 1u) ? w[6] : w[5],
                       t6 = (j & 1u) ? w[7] : w[6];
        const uint32_t v0 = (j & 2u) ? t2 : t0, v1 = (j & 2u) ? t3 : t1, v2 = (j & 2u) ? t4 : t2,
                       v3 = (j & 2u) ? t5 : t3, v4 = (j & 2u) ? t6 : t4;
        handle_token<ASCII>(p, wc, has, pp & 3u, L, v0, v1, v2, v3, v4);
        if (far) token_generic(p, wc.skeys, wc.scnt, chunk_pos + pp);
        __syncwarp();
    }
}

__global__ void __launch_bounds__(kMapThreads, 1) tokenize_combine_kernel(const __grid_constant__ MapParams p) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t *skeys = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *scnt = reinterpret_cast<uint32_t *>(smem_raw + kSmemSlots * 8);
    for (uint32_t i = threadIdx.x; i < kSmemSlots; i += kMapThreads) { skeys[i] = 0ull; scnt[i] = 0u; }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t nwarps = (uint64_t)gridDim.x * (kMapThreads / 32);
    const uint64_t gw = (uint64_t)blockIdx.x * (kMapThreads / 32) + (threadIdx.x >> 5);
    const uint64_t nrows = (p.hi + 511ull) >> 9;
    const uint64_t r0 = nrows * gw / nwarps, r1 = nrows * (gw + 1) / nwarps;
    const bool use_fast = !(p.flags & LOCUST_F_GENERIC_TOKENISER);
    WarpCtx wc;
    wc.skeys = skeys;
    wc.scnt = scnt;
    wc.queue = reinterpret_cast<ulonglong2 *>(smem_raw + kSmemSlots * 12) + (threadIdx.x >> 5) * kQueueSlots;
    wc.qn = 0u;
    wc.lane = lane;
    wc.use_smem = !(p.flags & LOCUST_F_NO_SMEM_COMBINER);

    if (r0 < r1) {
        uint4 c = load_chunk(p, r0, lane);
        uint4 n = load_chunk(p, r0 + 1, lane);
        uint32_t dc = classify16(c.x, c.y, c.z, c.w);
        uint32_t dn = classify16(n.x, n.y, n.z, n.w);
        bool asc_c = !__any_sync(kFull, (c.x | c.y | c.z | c.w) & 0x80808080u);
        bool asc_n = !__any_sync(kFull, (n.x | n.y | n.z | n.w) & 0x80808080u);
        // mask of the chunk in front of lane 0; only its top bit (last byte) is used
        uint32_t pd;
        {
            const uint64_t b = r0 * 512ull;
            pd = (b <= p.lo || is_delim(p.base[b - 1])) ? 0x8000u : 0u;
        }
        for (uint64_t row = r0; row < r1; ++row) {
            const uint4 nn = load_chunk(p, row + 2, lane);                 // in flight while this row is tokenised
            const uint32_t d_up = __shfl_sync(kFull, lane == 0 ? dn : dc, (lane + 1) & 31);
            const uint32_t d_dn = __shfl_sync(kFull, lane == 31 ? pd : dc, (lane + 31) & 31);
            const uint32_t starts = ~dc & ((dc << 1) | (d_dn >> 15)) & 0xFFFFu;
            const uint32_t dd = dc | (d_up << 16);
            const uint64_t chunk_pos = row * 512ull + lane * 16u;
            if (use_fast) {
                // 32-byte window: this lane's chunk and the next lane's (lane 31: next row's first chunk)
                uint32_t w[8];
                w[0] = c.x; w[1] = c.y; w[2] = c.z; w[3] = c.w;
                w[4] = __shfl_sync(kFull, lane == 0 ? n.x : c.x, (lane + 1) & 31);
                w[5] = __shfl_sync(kFull, lane == 0 ? n.y : c.y, (lane + 1) & 31);
                w[6] = __shfl_sync(kFull, lane == 0 ? n.z : c.z, (lane + 1) & 31);
                w[7] = __shfl_sync(kFull, lane == 0 ? n.w : c.w, (lane + 1) & 31);
                if (asc_c && asc_n) tokenise_row<true>(p, wc, w, starts, dd, chunk_pos);
                else tokenise_row<false>(p, wc, w, starts, dd, chunk_pos);
            } else {
                uint32_t rem = starts;
                while (rem) {
                    const uint32_t pp = (uint32_t)__ffs((int)rem) - 1u;
                    rem &= rem - 1u;
                    token_generic(p, skeys, scnt, chunk_pos + pp);
                }
                __syncwarp();
            }
            pd = dc; c = n; dc = dn; asc_c = asc_n;
            n = nn;
            dn = classify16(n.x, n.y, n.z, n.w);
            asc_n = !__any_sync(kFull, (n.x | n.y | n.z | n.w) & 0x80808080u);
        }
        drain_queue(p, wc, 1u);
    }
    __syncthreads();
    // flush the CTA's combiner into the global table
    for (uint32_t i = threadIdx.x; i < kSmemSlots; i += kMapThreads) {
        const uint64_t k = skeys[i];
        if (k) global_insert(p, k, 0ull, (unsigned long long)scnt[i]);
    }
}

cudaError_t map_kernel_setup() {
    return cudaFuncSetAttribute(tokenize_combine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
}

int launch_tokenize_combine(const MapParams &p, int num_sms, cudaStream_t st) {
    if (p.hi <= p.lo) return 0;
    const uint64_t nrows = (p.hi + 511ull) >> 9;
    uint64_t ctas = (nrows + (kMapThreads / 32) - 1) / (kMapThreads / 32);
    if (ctas > (uint64_t)num_sms) ctas = (uint64_t)num_sms;
    tokenize_combine_kernel<<<(unsigned)ctas, kMapThreads, kSmemBytes, st>>>(p);
    return 1;
}

}  // namespace locust
