// common.cuh -- shared device/host helpers of the word-count path (sm_100a).
//
// Byte classification follows the reference's delimiter literal
// " ,.-;:'()\"\t" (MapReduce/src/main.cu:138,:150) plus '\n' (getline, main.cu:43);
// key order follows KIVComparator (KeyValue.h:20-33): unsigned bytes, first
// difference decides, shorter prefix first.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

#if defined(__CUDACC__)
#define LC_HD __host__ __device__ __forceinline__
#define LC_D __device__ __forceinline__
#else
#define LC_HD inline
#define LC_D inline
#endif

namespace locust {

// One bit per byte value < 64: '\t' '\n' ' ' '"' '\'' '(' ')' ',' '-' '.' ':' ';'
constexpr uint64_t kDelimMask64 = 0x0C00738500000600ull;

LC_HD bool is_delim(uint32_t c) { return c < 64u && ((kDelimMask64 >> c) & 1ull); }

// SWAR classifier: for the four bytes of x, returns 0x80 in every byte that is a
// delimiter, 0 elsewhere.  Thresholds t are applied to the low seven bits
// (x7 + (0x80 - t) sets the top bit iff x7 >= t); the delimiter set is a union of
// the ranges [09,0A] [27,29] [2C,2E] [3A,3B], whose membership is the parity of
// the thresholds passed, plus {20,22}, tested as ((x7 & 7D) ^ 20) == 0.  Bytes
// with the top bit set are never delimiters.
LC_HD uint32_t classify4(uint32_t x) {
    const uint32_t x7 = x & 0x7F7F7F7Fu;
    const uint32_t r1 = x7 + 0x77777777u;   // >= 09
    const uint32_t r2 = x7 + 0x75757575u;   // >= 0B
    const uint32_t r3 = x7 + 0x59595959u;   // >= 27
    const uint32_t r4 = x7 + 0x56565656u;   // >= 2A
    const uint32_t r5 = x7 + 0x54545454u;   // >= 2C
    const uint32_t r6 = x7 + 0x51515151u;   // >= 2F
    const uint32_t r7 = x7 + 0x46464646u;   // >= 3A
    const uint32_t r8 = x7 + 0x44444444u;   // >= 3C
    const uint32_t z  = (x7 & 0x7D7D7D7Du) ^ 0x20202020u;
    const uint32_t nz = z + 0x7F7F7F7Fu;    // top bit iff z != 0
    const uint32_t a = r1 ^ r2 ^ r3;
    const uint32_t b = r4 ^ r5 ^ r6;
    const uint32_t c = r7 ^ r8 ^ ~nz;
    return (a ^ b ^ c) & ~x & 0x80808080u;
}

// 16-bit delimiter mask of a 16-byte chunk (bit i = byte i is a delimiter): the four 0x80
// flags of each word are gathered into a nibble by one multiply (0x00204081 moves bits
// 7, 15, 23, 31 to bits 28..31 without carries).
LC_HD uint32_t classify16(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
    const uint32_t p0 = classify4(w0) * 0x00204081u;
    const uint32_t p1 = classify4(w1) * 0x00204081u;
    const uint32_t p2 = classify4(w2) * 0x00204081u;
    const uint32_t p3 = classify4(w3) * 0x00204081u;
    return (p0 >> 28) | ((p1 >> 24) & 0xF0u) | ((p2 >> 20) & 0xF00u) | ((p3 >> 16) & 0xF000u);
}

#if defined(__CUDACC__)
// The same classifier with `nimad` of its nine byte-wise adds issued as multiply-adds (x * one + c, `one` a
// register holding 1 that the compiler cannot see through): IMAD runs on the FMA pipe, the rest of the
// classifier on the ALU pipe, and the two pipes issue side by side.
template <int nimad>
__device__ __forceinline__ uint32_t classify4_bal(uint32_t x, uint32_t one) {
    const uint32_t x7 = x & 0x7F7F7F7Fu;
#define LC_ADD(i, c) ((i) < nimad ? x7 * one + (c) : x7 + (c))
    const uint32_t r1 = LC_ADD(0, 0x77777777u);
    const uint32_t r2 = LC_ADD(1, 0x75757575u);
    const uint32_t r3 = LC_ADD(2, 0x59595959u);
    const uint32_t r4 = LC_ADD(3, 0x56565656u);
    const uint32_t r5 = LC_ADD(4, 0x54545454u);
    const uint32_t r6 = LC_ADD(5, 0x51515151u);
    const uint32_t r7 = LC_ADD(6, 0x46464646u);
    const uint32_t r8 = LC_ADD(7, 0x44444444u);
#undef LC_ADD
    const uint32_t z = (x7 & 0x7D7D7D7Du) ^ 0x20202020u;
    const uint32_t nz = (8 < nimad) ? z * one + 0x7F7F7F7Fu : z + 0x7F7F7F7Fu;
    const uint32_t a = r1 ^ r2 ^ r3;
    const uint32_t b = r4 ^ r5 ^ r6;
    const uint32_t c = r7 ^ r8 ^ ~nz;
    return (a ^ b ^ c) & ~x & 0x80808080u;
}
template <int nimad>
__device__ __forceinline__ uint32_t classify16_bal(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t one) {
    const uint32_t p0 = classify4_bal<nimad>(w0, one) * 0x00204081u;
    const uint32_t p1 = classify4_bal<nimad>(w1, one) * 0x00204081u;
    const uint32_t p2 = classify4_bal<nimad>(w2, one) * 0x00204081u;
    const uint32_t p3 = classify4_bal<nimad>(w3, one) * 0x00204081u;
    return (p0 >> 28) | ((p1 >> 24) & 0xF0u) | ((p2 >> 20) & 0xF00u) | ((p3 >> 16) & 0xF000u);
}
__device__ __forceinline__ uint32_t opaque_one() { uint32_t v; asm volatile("mov.u32 %0, 1;" : "=r"(v)); return v; }
// v += (value of v `delta` lanes below), for lanes that have such a lane: the shuffle's own predicate says
// whether the source lane exists, which saves the compare + select of the intrinsic form
__device__ __forceinline__ uint32_t scan_step(uint32_t v, uint32_t delta) {
    asm("{\n\t.reg .pred p;\n\t.reg .u32 t;\n\t"
        "shfl.sync.up.b32 t|p, %0, %1, 0, 0xffffffff;\n\t"
        "@p add.u32 %0, %0, t;\n\t}" : "+r"(v) : "r"(delta));
    return v;
}
#endif

LC_HD uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) {
#if defined(__CUDA_ARCH__)
    // raw PRMT: the __byte_perm intrinsic only honours three bits per selector nibble
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s));
    return r;
#else
    // PTX prmt, default mode: nibble i of s picks one of the 8 bytes {a,b}; bit 3
    // of the nibble replicates that byte's sign bit instead.
    uint64_t t = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) {
        uint32_t n = (s >> (4 * i)) & 0xF;
        uint32_t v = (uint32_t)(t >> (8 * (n & 7))) & 0xFF;
        if (n & 8) v = (v & 0x80) ? 0xFF : 0x00;
        r |= v << (8 * i);
    }
    return r;
#endif
}

// Key of a token of L <= 16 bytes starting at byte `o` (0..3) of word w0, with
// w1..w4 the following words: four little-endian words, zero padded past L.
// ASCII variant (every byte of the window < 0x80): bytes past L are zeroed inside
// the PRMT -- bit 3 of a selector nibble replicates the selected byte's sign bit,
// which is 0x00 for any byte below 0x80.
LC_HD uint32_t shl_clamp(uint32_t v, uint32_t n) {      // v << n, 0 for n >= 32
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));   // PTX shl clamps the shift amount at 32
    return r;
#else
    return n >= 32u ? 0u : v << n;
#endif
}
// The 64-bit value 0x8888888888888888 << 4L (zlo = low word, zhi = high word) has nibble 8
// exactly at the key bytes at or past L.  L >= 16 (and the "no delimiter" value 0xFFFFFFFF)
// gives 0: nothing is zeroed.
LC_HD void key_selectors(uint32_t o, uint32_t L, uint32_t &sel, uint32_t &zlo, uint32_t &zhi) {
    sel = 0x3210u + 0x1111u * o;
    zlo = shl_clamp(0x88888888u, 4u * L);
    zhi = shl_clamp(0x88888888u, 4u * (L > 8u ? L : 8u) - 32u);
}
LC_HD void extract_key16_ascii(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4, uint32_t o,
                               uint32_t L, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    uint32_t sel, zlo, zhi;
    key_selectors(o, L, sel, zlo, zhi);
    a = byte_perm(w0, w1, sel | (zlo & 0xFFFFu));
    b = byte_perm(w1, w2, sel | (zlo >> 16));
    c = byte_perm(w2, w3, sel | (zhi & 0xFFFFu));
    d = byte_perm(w3, w4, sel | (zhi >> 16));
}
// Any byte values: align with PRMT, then AND with a byte mask built from the same
// nibble mask (nibble 8 -> selector 1 -> byte 0x00; nibble 0 -> selector 0 -> 0xFF).
LC_HD void extract_key16_any(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4, uint32_t o,
                             uint32_t L, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    uint32_t sel, zlo, zhi;
    key_selectors(o, L, sel, zlo, zhi);
    a = byte_perm(w0, w1, sel) & byte_perm(0xFFu, 0u, (zlo >> 3) & 0xFFFFu);
    b = byte_perm(w1, w2, sel) & byte_perm(0xFFu, 0u, zlo >> 19);
    c = byte_perm(w2, w3, sel) & byte_perm(0xFFu, 0u, (zhi >> 3) & 0xFFFFu);
    d = byte_perm(w3, w4, sel) & byte_perm(0xFFu, 0u, zhi >> 19);
}

LC_HD uint64_t bswap64(uint64_t v) {
    v = ((v & 0x00FF00FF00FF00FFull) << 8) | ((v >> 8) & 0x00FF00FF00FF00FFull);
    v = ((v & 0x0000FFFF0000FFFFull) << 16) | ((v >> 16) & 0x0000FFFF0000FFFFull);
    return (v << 32) | (v >> 32);
}

// 64-bit mix (splitmix64 finaliser) -- table slot hash for (k0, k1) keys.
LC_HD uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
LC_HD uint64_t hash_key128(uint64_t k0, uint64_t k1) {
    return mix64(k0 ^ (k1 * 0x9E3779B97F4A7C15ull) ^ (k1 >> 29));
}
// Home slot of a key in the global combiner table (at most 2^32 slots): a 32-bit multiply-add over the
// key's four words and one round of avalanche -- a dozen instructions where the 64-bit mix above costs
// about forty-five on a GPU without 64-bit integer multipliers.  Every inserter into one table (MAP, the
// flush of the shared-memory table, the shuffle merge) uses this one function.
LC_HD uint32_t slot_hash32(uint64_t k0, uint64_t k1) {
    uint32_t h = (uint32_t)k0 * 0x9E3779B1u + (uint32_t)(k0 >> 32) * 0x85EBCA77u + (uint32_t)k1 * 0xC2B2AE3Du +
                 (uint32_t)(k1 >> 32) * 0x27D4EB2Fu;
    h ^= h >> 15; h *= 0x2C1B3C6Du;
    h ^= h >> 12; h *= 0x297A2D39u;
    h ^= h >> 15;
    return h;
}

// 32 bytes in one access: sm_100 has 256-bit global loads and stores, and the table slots and
// the sort records are 32-byte aligned 32-byte structs (one memory sector each), so moving one
// as a unit halves the load/store instructions and the sector requests of the 128-bit pairs the
// compiler emits for a struct copy.  p must be 32-byte aligned.
#if defined(__CUDACC__)
__device__ __forceinline__ ulonglong4 ld256(const void *p) {
    ulonglong4 v;
    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v.x), "=l"(v.y), "=l"(v.z), "=l"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st256(void *p, const ulonglong4 &v) {
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(v.x), "l"(v.y), "l"(v.z), "l"(v.w) : "memory");
}
#endif

// ---- global combiner table -------------------------------------------------
// One slot = 32 bytes: key (k0, k1) = the token's bytes zero padded to 16,
// little endian (k0 = bytes 0..7).  Tokens contain no NUL, so the padded value
// identifies the token exactly and an all-zero key marks an empty slot.
struct __align__(32) Slot {
    uint64_t k0, k1;
    unsigned long long count;
    uint64_t pad;
};

}  // namespace locust
