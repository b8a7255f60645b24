// Host-side check of the bit tricks in locust_b200/csrc/common.cuh (compiled with g++,
// no GPU): SWAR delimiter classifier, PRMT key extraction, small helpers.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../locust_b200/csrc/common.cuh"
using namespace locust;

static bool naive_delim(unsigned c) { return c && strchr(" ,.-;:'()\"\t\n", (int)c) != nullptr; }
static uint64_t rng_state = 88172645463325252ull;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

int main() {
    // is_delim over all byte values
    for (unsigned c = 0; c < 256; ++c)
        if (is_delim(c) != naive_delim(c)) { printf("is_delim %u\n", c); return 1; }
    // classify4: every byte value in every position with random neighbours
    for (unsigned c = 0; c < 256; ++c)
        for (int pos = 0; pos < 4; ++pos)
            for (int t = 0; t < 64; ++t) {
                uint32_t w = (uint32_t)rnd();
                w = (w & ~(0xFFu << (8 * pos))) | (c << (8 * pos));
                const uint32_t f = classify4(w);
                if (f & ~0x80808080u) { printf("classify4 stray bits\n"); return 1; }
                for (int b = 0; b < 4; ++b) {
                    const unsigned byte = (w >> (8 * b)) & 0xFF;
                    if ((((f >> (8 * b + 7)) & 1u) != 0) != naive_delim(byte)) { printf("classify4 %08x byte %d\n", w, b); return 1; }
                }
            }
    // classify16
    for (int t = 0; t < 200000; ++t) {
        uint8_t b[16];
        for (auto &x : b) { const uint64_t r = rnd(); x = (r & 3) ? (uint8_t)(r >> 8) : (uint8_t)" ,.-;:'()\"\t\n"[(r >> 8) % 12]; }
        uint32_t w[4];
        memcpy(w, b, 16);
        const uint32_t d = classify16(w[0], w[1], w[2], w[3]);
        for (int i = 0; i < 16; ++i)
            if (((d >> i) & 1u) != (naive_delim(b[i]) ? 1u : 0u)) { printf("classify16 bit %d\n", i); return 1; }
        if (d >> 16) { printf("classify16 high bits\n"); return 1; }
    }
    // extract_key16_*: all start offsets and lengths 1..16; the ASCII variant on ASCII data,
    // the general variant on every byte value except NUL
    for (int t = 0; t < 60000; ++t) {
        uint8_t b[20], g[20];
        for (auto &x : b) x = (uint8_t)(1 + rnd() % 127);
        for (auto &x : g) x = (uint8_t)(1 + rnd() % 255);
        uint32_t w[5], v[5];
        memcpy(w, b, 20);
        memcpy(v, g, 20);
        for (uint32_t o = 0; o < 4; ++o)
            for (uint32_t L = 1; L <= 16; ++L) {
                uint32_t k[4], q[4];
                extract_key16_ascii(w[0], w[1], w[2], w[3], w[4], o, L, k[0], k[1], k[2], k[3]);
                extract_key16_any(v[0], v[1], v[2], v[3], v[4], o, L, q[0], q[1], q[2], q[3]);
                uint8_t want[16] = {0}, wantg[16] = {0};
                for (uint32_t i = 0; i < L; ++i) { want[i] = b[o + i]; wantg[i] = g[o + i]; }
                if (memcmp(k, want, 16) != 0) { printf("extract_key16_ascii o=%u L=%u\n", o, L); return 1; }
                if (memcmp(q, wantg, 16) != 0) { printf("extract_key16_any o=%u L=%u\n", o, L); return 1; }
            }
    }
    if (bswap64(0x0102030405060708ull) != 0x0807060504030201ull) { printf("bswap64\n"); return 1; }
    printf("ok\n");
    return 0;
}
