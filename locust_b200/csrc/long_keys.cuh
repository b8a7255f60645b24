// long_keys.cuh -- combiner for keys longer than 16 bytes.
//
// The reference treats every key alike up to its record width (KeyValue.h:6-18, main.cu:146);
// here keys of <= 16 bytes live in the 128-bit-key table and longer ones in a list of distinct
// long keys {ptr, len, count, hash} (LongOcc) with an open-addressed index over it: a slot of the
// index holds list position + 1.  A key is looked up by the 64-bit hash of all its bytes and
// matched by hash, length and a full byte compare against the bytes the list entry points at, so
// every further occurrence only adds to the entry's count.
//
// Claiming is two-step (slot -> kBusy, entry written, slot -> position) without waiting: a thread
// that meets a slot in the busy state cannot compare, so it moves on and may create a second entry
// for the same key.  That is rare and harmless: REDUCE adds up records with equal keys.
#pragma once
#include "kernels.cuh"

namespace locust {

constexpr uint32_t kLongBusy = 0xFFFFFFFFu;

LC_D uint64_t long_hash(const uint8_t *s, uint64_t len) {
    uint64_t h = 0x9E3779B97F4A7C15ull ^ len;
    for (uint64_t i = 0; i < len; ++i) h = (h ^ s[i]) * 0x100000001B3ull;
    h = mix64(h);
    return h ? h : 1ull;
}

LC_D bool long_equal(const uint8_t *a, const uint8_t *b, uint64_t len) {
    if (a == b) return true;
    for (uint64_t i = 0; i < len; ++i)
        if (a[i] != b[i]) return false;
    return true;
}

// Adds `inc` occurrences of the key s[0..len) (len > 16) to the long-key list.  `s` must stay valid
// until the job's result has been produced (it points into the input, or into a received part).
LC_D void long_insert(const MapParams &p, const uint8_t *s, uint64_t len, unsigned long long inc) {
    if (p.long_cap == 0) {                           // no room at all (locust_count_static): reported, not stored
        atomicAdd(&p.ctr->n_long, 1ull);
        atomicOr(&p.ctr->status, kStatusLongFull);
        return;
    }
    // a list that is already known to be too small: the host grows it and redoes the job
    if (*reinterpret_cast<volatile uint32_t *>(&p.ctr->status) & kStatusLongFull) return;
    const uint64_t h = long_hash(s, len);
    uint64_t slot = h & p.long_index_mask;
    for (uint64_t probe = 0; probe <= p.long_index_mask; ++probe, slot = (slot + 1) & p.long_index_mask) {
        uint32_t v;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p.long_index + slot) : "memory");
        if (v == 0u) {
            v = atomicCAS(p.long_index + slot, 0u, kLongBusy);
            if (v == 0u) {                           // ours: make the entry, then publish its position
                const unsigned long long idx = atomicAdd(&p.ctr->n_long, 1ull);
                if (idx >= p.long_cap) {             // list full: the host grows it and redoes the job
                    atomicOr(&p.ctr->status, kStatusLongFull);
                    return;
                }
                LongOcc *e = p.long_list + idx;
                e->ptr = s; e->len = len; e->count = inc; e->hash = h;
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.long_index + slot), "r"((uint32_t)idx + 1u) : "memory");
                return;
            }
        }
        if (v == kLongBusy) continue;                // being written: cannot compare, move on
        const LongOcc *e = p.long_list + (v - 1u);
        if (e->hash == h && e->len == len && long_equal(e->ptr, s, len)) {
            atomicAdd(&p.long_list[v - 1u].count, inc);
            return;
        }
    }
    atomicOr(&p.ctr->status, kStatusLongFull);
}

}  // namespace locust
