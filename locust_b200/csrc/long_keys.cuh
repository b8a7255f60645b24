// long_keys.cuh -- combiner for keys longer than 16 bytes.
//
// The reference treats every key alike up to its record width (KeyValue.h:6-18, main.cu:146);
// here keys of <= 16 bytes live in the 128-bit-key table and longer ones in a list of distinct
// long keys {ptr, len, count, hash} (LongOcc) with an open-addressed index over it: a slot of the
// index holds list position + 1.  A key is looked up by the 64-bit hash of all its bytes and
// matched by hash, length and a full byte compare against the bytes the list entry points at, so
// every further occurrence only adds to the entry's count.
//
// Claiming is two-step (slot -> kBusy, entry written, slot -> position).  A thread that meets a slot in the
// busy state waits a bounded moment for the position; if it still cannot compare it moves on and may
// create a second entry for the same key.  That is rare and harmless: REDUCE adds up records with equal keys.
#pragma once
#include "kernels.cuh"

namespace locust {

constexpr uint32_t kLongBusy = 0xFFFFFFFFu;

// The MAP kernels keep almost all of the SM's L1 as shared memory, so a byte-at-a-time walk over a token
// costs one L2 round trip per byte.  Long keys are therefore read eight bytes at a time: chunk k of a key is
// its bytes [8k, 8k + 8), put together from the one or two aligned 64-bit words that hold them (aligned words
// never straddle a page, and a word is only touched if it holds a byte of the key), zero padded past the end.
LC_D uint64_t long_chunk(const uint8_t *key, uint64_t k, uint64_t len) {
    const uint64_t a = reinterpret_cast<uint64_t>(key) + 8ull * k, rem = len - 8ull * k;      // rem > 0
    const uint64_t *w = reinterpret_cast<const uint64_t *>(a & ~7ull);
    const uint32_t off = (uint32_t)(a & 7ull), sh = 8u * off;
    uint64_t c = *w >> sh;                              // input text: read-only for the whole job
    if (off != 0u && rem > 8ull - off) c |= w[1] << (64u - sh);
    if (rem < 8ull) c &= (1ull << (8u * (uint32_t)rem)) - 1ull;
    return c;
}
LC_D uint64_t long_mix(uint64_t h, uint64_t chunk) {
    h = (h ^ chunk) * 0x9E3779B97F4A7C15ull;
    return h ^ (h >> 32);
}
LC_D uint64_t long_finish(uint64_t h, uint64_t len) {
    h = mix64(h ^ (len * 0xD6E8FEB86659FD93ull));
    return h ? h : 1ull;
}
constexpr uint64_t kLongSeed = 0x243F6A8885A308D3ull;

LC_D uint64_t long_hash(const uint8_t *s, uint64_t len) {
    uint64_t h = kLongSeed;
    for (uint64_t k = 0; 8ull * k < len; ++k) h = long_mix(h, long_chunk(s, k, len));
    return long_finish(h, len);
}

LC_D bool long_equal(const uint8_t *a, const uint8_t *b, uint64_t len) {
    if (a == b) return true;
    for (uint64_t k = 0; 8ull * k < len; ++k)
        if (long_chunk(a, k, len) != long_chunk(b, k, len)) return false;
    return true;
}

// Length and hash of the token that starts at s in one pass: at most maxlen bytes belong to the input (the
// token ends at the first delimiter or there).  Eight bytes at a time, delimiters found with the SWAR classifier.
LC_D uint64_t long_scan(const uint8_t *s, uint64_t maxlen, uint64_t &hash) {
    uint64_t h = kLongSeed, L = 0;
    for (uint64_t k = 0; 8ull * k < maxlen; ++k) {
        uint64_t c = long_chunk(s, k, maxlen);
        const uint64_t rem = maxlen - 8ull * k;
        if (rem < 8ull) c |= 0x0A0A0A0A0A0A0A0Aull << (8u * (uint32_t)rem);           // beyond the input: a delimiter
        const uint32_t f0 = classify4((uint32_t)c), f1 = classify4((uint32_t)(c >> 32));
        if (f0 | f1) {
            const uint32_t nb = f0 ? ((uint32_t)__ffs((int)f0) - 1u) >> 3 : 4u + (((uint32_t)__ffs((int)f1) - 1u) >> 3);
            if (nb) h = long_mix(h, c & ((1ull << (8u * nb)) - 1ull));
            L += nb;
            hash = long_finish(h, L);
            return L;
        }
        h = long_mix(h, c);
        L += 8ull;
    }
    hash = long_finish(h, L);
    return L;
}

// Adds `inc` occurrences of the key s[0..len) (len > 16, h = its long_hash) to the long-key list.  `s` must stay valid
// until the job's result has been produced (it points into the input, or into a received part).
__device__ __noinline__ static void long_insert(const MapParams &p, const uint8_t *s, uint64_t len, uint64_t h, unsigned long long inc) {
    // The parameters are copied out once and the compiler is told what they point at: read through the reference
    // inside the probe loop they would be fetched again with generic loads on every step, and the atomics would be
    // generic ones (the function is not inlined, so it cannot see that these are global-memory pointers).
    Counters *const ctr = p.ctr;
    uint32_t *const index = p.long_index;
    LongOcc *const list = p.long_list;
    const uint64_t mask = p.long_index_mask, cap = p.long_cap;
    __builtin_assume(__isGlobal(ctr));
    __builtin_assume(__isGlobal(index));
    __builtin_assume(__isGlobal(list));
    if (cap == 0) {                                  // no room at all (locust_count_static): reported, not stored
        atomicAdd(&ctr->n_long, 1ull);
        atomicOr(&ctr->status, kStatusLongFull);
        return;
    }
    // a list that is already known to be too small: the host grows it and redoes the job
    if (*reinterpret_cast<volatile uint32_t *>(&ctr->status) & kStatusLongFull) return;
    uint64_t slot = h & mask;
    for (uint64_t probe = 0; probe <= mask; ++probe, slot = (slot + 1) & mask) {
        // read through L2 (other SMs write these words), without an acquire: that would invalidate this SM's L1
        // on every long token.  The entry's fields are read through L2 too, at an address that depends on the
        // value read here, and the writer fences between writing the entry and publishing its position.
        uint32_t v = __ldcg(index + slot);
        if (v == 0u) {
            v = atomicCAS(index + slot, 0u, kLongBusy);
            if (v == 0u) {                           // ours: make the entry, then publish its position
                const unsigned long long idx = atomicAdd(&ctr->n_long, 1ull);
                if (idx >= cap) {                    // list full: the host grows it and redoes the job
                    atomicOr(&ctr->status, kStatusLongFull);
                    return;
                }
                LongOcc *e = list + idx;
                e->ptr = s; e->len = len; e->count = inc; e->hash = h;
                __threadfence();
                __stcg(index + slot, (uint32_t)idx + 1u);
                return;
            }
        }
        for (int spin = 0; v == kLongBusy && spin < 256; ++spin) {          // being written by its claimer: a moment
            __nanosleep(32);
            v = __ldcg(index + slot);
        }
        if (v == kLongBusy) continue;                // still not published: cannot compare, move on (a duplicate entry at worst)
        const ulonglong2 pl = __ldcg(reinterpret_cast<const ulonglong2 *>(list + (v - 1u)));          // ptr, len
        const uint64_t eh = __ldcg(&list[v - 1u].hash);
        if (eh == h && pl.y == len && long_equal(reinterpret_cast<const uint8_t *>(pl.x), s, len)) {
            atomicAdd(&list[v - 1u].count, inc);
            return;
        }
    }
    atomicOr(&ctr->status, kStatusLongFull);
}

// MAP's fast kernels do not walk long tokens themselves (that needs registers and an L1 they do not have):
// they list where each one starts, and long_tokens_kernel (process_kernels.cu) takes it from there.
LC_D void far_push(const MapParams &p, uint64_t pos) {
    const unsigned long long idx = atomicAdd(&p.ctr->n_far, 1ull);
    if (idx < p.far_cap) p.far_list[idx] = make_ulonglong2(reinterpret_cast<unsigned long long>(p.base + pos), reinterpret_cast<unsigned long long>(p.base + p.hi));
    else atomicOr(&p.ctr->status, kStatusFarFull);
}

// The token that starts at s is longer than 16 bytes: find its end, hash it, count it.  Out of line: the MAP
// kernels call it from their rare path and must not pay its registers in their main loop.
__device__ __noinline__ static void long_token_at(const MapParams &p, const uint8_t *s, uint64_t maxlen) {
    uint64_t h;
    const uint64_t L = long_scan(s, maxlen, h);
    long_insert(p, s, L, h, 1ull);
}

}  // namespace locust
