// textgen.cu -- synthetic text generator (see textgen.h).  Bench/test infrastructure.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_set>
#include <vector>

#include <cuda_runtime.h>
#include "textgen.h"

#define TG_HD __host__ __device__ __forceinline__

namespace {

struct GenParams {
    int mode;
    uint32_t vocab;
    uint64_t seed;
    uint64_t nblocks_total;
    const uint64_t *cdf;    // vocab inclusive thresholds scaled to 2^64 (last = 2^64 - 1)
    const uint8_t *words;   // vocab * 16 bytes
    const uint8_t *wlen;    // vocab
};

TG_HD uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}

struct Emitter {
    uint8_t *dst;
    uint32_t pos;
    uint64_t acc;
    TG_HD void put(uint32_t c) {
        acc |= (uint64_t)c << (8u * (pos & 7u));
        ++pos;
        if ((pos & 7u) == 0u) { *reinterpret_cast<uint64_t *>(dst + pos - 8u) = acc; acc = 0ull; }
    }
};

TG_HD uint32_t zipf_rank(const GenParams &p, uint64_t u) {
    // smallest r with u <= cdf[r]
    uint32_t lo = 0, hi = p.vocab - 1;
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (u <= p.cdf[mid]) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// C4 words: 7 base-26 letters of a 32-bit bijection of the rank, then 1-3 more letters
TG_HD uint32_t uniform_word(uint32_t rank, uint8_t *w) {
    uint64_t x = (uint32_t)(rank * 2654435761u);
    for (int i = 0; i < 7; ++i) { w[i] = (uint8_t)('a' + x % 26u); x /= 26u; }
    uint64_t h = mix64(0x5bd1e995ull + rank);
    const uint32_t extra = 1u + (uint32_t)(h % 3u);
    h >>= 8;
    for (uint32_t i = 0; i < extra; ++i) { w[7 + i] = (uint8_t)('a' + h % 26u); h /= 26u; }
    return 7u + extra;
}

TG_HD void gen_block(const GenParams &p, uint64_t b, uint8_t *dst) {
    Emitter e{dst, 0u, 0ull};
    uint64_t s = mix64(p.seed ^ mix64(b + 0x1234567ull));
    const uint32_t eighth = (uint32_t)((b * 8ull) / (p.nblocks_total ? p.nblocks_total : 1ull));
    const uint32_t heavy_permille[8] = {950, 900, 800, 700, 600, 500, 300, 50};
    const bool short_lines = (p.mode != TEXTGEN_SKEWED) || eighth == 4u;
    uint32_t tok = 0;
    for (;;) {
        s += 0x9E3779B97F4A7C15ull;
        const uint64_t r = mix64(s);
        s += 0x9E3779B97F4A7C15ull;
        const uint64_t r2 = mix64(s);
        uint8_t tmp[16];
        const uint8_t *w = tmp;
        uint32_t wl;
        if (p.mode == TEXTGEN_UNIFORM) {
            wl = uniform_word((uint32_t)((r >> 11) % p.vocab), tmp);
        } else if (p.mode == TEXTGEN_SKEWED && (uint32_t)(r2 >> 40) % 1000u < heavy_permille[eighth & 7u]) {
            // bursts: the heavy word only changes every 16 tokens
            const uint32_t which = (uint32_t)(mix64(p.seed + b * 4096ull + (tok >> 4)) % 3u);
            const char *hw = which == 0 ? "a" : (which == 1 ? "the" : "xylophonists");
            wl = which == 0 ? 1u : (which == 1 ? 3u : 12u);
            for (uint32_t i = 0; i < wl; ++i) tmp[i] = (uint8_t)hw[i];
        } else {
            const uint32_t rank = zipf_rank(p, r);
            w = p.words + (uint64_t)rank * 16u;
            wl = p.wlen[rank];
        }
        if (e.pos + wl + 2u > TEXTGEN_BLOCK - 1u) break;
        for (uint32_t i = 0; i < wl; ++i) e.put(w[i]);
        const uint32_t u = (uint32_t)(r2 % 100ull);
        if (u < 85u) e.put(' ');
        else if (u < 92u) { e.put(','); e.put(' '); }
        else if (u < 96u) { e.put('.'); e.put(' '); }
        else e.put(short_lines ? '\n' : ' ');
        ++tok;
    }
    while (e.pos < TEXTGEN_BLOCK - 1u) e.put(' ');
    // skewed text: a line ends at a block boundary once in about 4096 blocks (16 MiB)
    const bool line_end = short_lines || (mix64(p.seed ^ (b * 0xD6E8FEB86659FD93ull)) & 4095ull) == 0ull;
    e.put(line_end ? '\n' : ' ');
}

__global__ void __launch_bounds__(128) gen_kernel(GenParams p, uint64_t first_block, uint64_t nblocks, uint8_t *dst) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nblocks) gen_block(p, first_block + i, dst + i * TEXTGEN_BLOCK);
}

std::string g_err;

}  // namespace

struct textgen {
    GenParams host{};
    GenParams dev{};
    int device = -1;
    std::vector<uint64_t> cdf;
    std::vector<uint8_t> words, wlen;
    uint64_t *d_cdf = nullptr;
    uint8_t *d_words = nullptr, *d_wlen = nullptr;
};

static void build_vocab(textgen *g, uint32_t vocab, double s, uint64_t vocab_seed) {
    g->words.assign((size_t)vocab * 16, 0);
    g->wlen.assign(vocab, 0);
    std::unordered_set<std::string> seen;
    seen.reserve((size_t)vocab * 2);
    for (uint32_t r = 0; r < vocab; ++r) {
        // frequent ranks are short: length grows with log(rank), +-1 jitter, 1..14
        int base = 1 + (int)std::floor(std::log((double)r + 2.0) / std::log(3.2));
        for (uint32_t attempt = 0;; ++attempt) {
            uint64_t h = mix64(vocab_seed ^ mix64(((uint64_t)r << 20) + attempt));
            int len = base + (int)(h % 3u) - 1 + (int)(attempt / 8u);
            h >>= 4;
            if (len < 1) len = 1;
            if (len > 14) len = 14;
            char w[16];
            for (int i = 0; i < len; ++i) { w[i] = (char)('a' + h % 26u); h /= 26u; if (i == 9) h = mix64(h + r); }
            std::string str(w, (size_t)len);
            if (seen.insert(str).second) {
                memcpy(&g->words[(size_t)r * 16], w, (size_t)len);
                g->wlen[r] = (uint8_t)len;
                break;
            }
        }
    }
    // Zipf(s) inverse CDF, thresholds scaled to 2^64
    g->cdf.resize(vocab);
    long double total = 0.0L;
    for (uint32_t r = 0; r < vocab; ++r) total += 1.0L / powl((long double)(r + 1), (long double)s);
    long double acc = 0.0L;
    for (uint32_t r = 0; r < vocab; ++r) {
        acc += 1.0L / powl((long double)(r + 1), (long double)s);
        long double f = acc / total;
        g->cdf[r] = f >= 1.0L ? ~0ull : (uint64_t)(f * 18446744073709551615.0L);
    }
    g->cdf[vocab - 1] = ~0ull;
}

extern "C" {

const char *textgen_last_error(void) { return g_err.c_str(); }

int textgen_create(int mode, uint32_t vocab, double zipf_s, uint64_t seed, uint64_t vocab_seed, uint64_t total_bytes,
                   int device, textgen **out) {
    if (!out || vocab == 0 || total_bytes % TEXTGEN_BLOCK) { g_err = "textgen_create: bad argument"; return 1; }
    textgen *g = new textgen();
    g->device = device;
    g->host.mode = mode; g->host.vocab = vocab; g->host.seed = seed;
    g->host.nblocks_total = total_bytes / TEXTGEN_BLOCK;
    if (mode != TEXTGEN_UNIFORM) build_vocab(g, vocab, zipf_s, vocab_seed);
    g->host.cdf = g->cdf.data(); g->host.words = g->words.data(); g->host.wlen = g->wlen.data();
    g->dev = g->host;
    if (device >= 0) {
        cudaError_t e = cudaSetDevice(device);
        if (e == cudaSuccess && mode != TEXTGEN_UNIFORM) {
            e = cudaMalloc(&g->d_cdf, g->cdf.size() * 8);
            if (e == cudaSuccess) e = cudaMalloc(&g->d_words, g->words.size());
            if (e == cudaSuccess) e = cudaMalloc(&g->d_wlen, g->wlen.size());
            if (e == cudaSuccess) e = cudaMemcpy(g->d_cdf, g->cdf.data(), g->cdf.size() * 8, cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(g->d_words, g->words.data(), g->words.size(), cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(g->d_wlen, g->wlen.data(), g->wlen.size(), cudaMemcpyHostToDevice);
        }
        if (e != cudaSuccess) { g_err = cudaGetErrorString(e); textgen_destroy(g); return 2; }
        g->dev.cdf = g->d_cdf; g->dev.words = g->d_words; g->dev.wlen = g->d_wlen;
    }
    *out = g;
    return 0;
}

void textgen_destroy(textgen *g) {
    if (!g) return;
    if (g->device >= 0) { cudaSetDevice(g->device); cudaFree(g->d_cdf); cudaFree(g->d_words); cudaFree(g->d_wlen); }
    delete g;
}

int textgen_fill_host(textgen *g, uint8_t *dst, uint64_t nbytes, uint64_t byte_offset) {
    if (!g || !dst || nbytes % TEXTGEN_BLOCK || byte_offset % TEXTGEN_BLOCK) { g_err = "textgen_fill_host: bad argument"; return 1; }
    const uint64_t b0 = byte_offset / TEXTGEN_BLOCK, nb = nbytes / TEXTGEN_BLOCK;
    for (uint64_t i = 0; i < nb; ++i) gen_block(g->host, b0 + i, dst + i * TEXTGEN_BLOCK);
    return 0;
}

int textgen_fill_device(textgen *g, uint8_t *dev_dst, uint64_t nbytes, uint64_t byte_offset, void *stream) {
    if (!g || g->device < 0 || !dev_dst || nbytes % TEXTGEN_BLOCK || byte_offset % TEXTGEN_BLOCK) {
        g_err = "textgen_fill_device: bad argument"; return 1;
    }
    cudaSetDevice(g->device);
    const uint64_t b0 = byte_offset / TEXTGEN_BLOCK, nb = nbytes / TEXTGEN_BLOCK;
    if (nb == 0) return 0;
    gen_kernel<<<(unsigned)((nb + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(g->dev, b0, nb, dev_dst);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return 2; }
    return 0;
}

}  // extern "C"
