/*
 * TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * CPU restatement of the word-count hot path of wuyan33/Locust, CPU branch
 * (MapReduce/src/main.cu with GPU_IMPLEMENTATION 0).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this; the product (liblocust.so) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against
 *   (1) the golden vectors in tests/golden/ that were produced by running the
 *       reference's own CPU branch (oracle/_ref/MapReduce_cpu, built by
 *       oracle/build_ref.sh from the sources under /root/reference) on
 *       hamlet.txt and on hand-made edge files, and
 *   (2) when oracle/_ref/MapReduce_cpu is present, that binary itself on
 *       random capped-size ASCII files.
 *
 * What is restated (reference file:line):
 *   loadFile      main.cu:242-259   split the input into lines at '\n'
 *   map / cpuMap  main.cu:295-317   tokens of a line under the delimiter
 *                                   literal " ,.-;:'()\"\t" (main.cu:297,:308)
 *   my_strtok_r   util.cu:54-89     skip leading delimiters, cut at the next one
 *   process       main.cu:498 + KeyValue.h:50-68
 *                                   comparison sort of every token record by
 *                                   key bytes (prefix sorts first)
 *   cpuReduce     main.cu:337-355, reduce main.cu:326-334
 *                                   run-length count over the sorted records
 *   printKeyValues main.cu:280-289  "print key: %s \t value: %s\n" (:286)
 *
 * What is deliberately lifted: the fixed
 * capacities (5 800 lines, 20 tokens per line, 99-byte lines, int counts).
 * None of them fires on hamlet.txt.  Token records are (offset, length) views
 * into the input instead of 204-byte copies; counts are 64-bit.
 *
 * Order contract: unsigned byte order, i.e. memcmp then length.  For 7-bit
 * ASCII this equals the reference's signed-char comparator
 * (KeyValue.h:60-66).  A NUL byte ends its line exactly as it does in the
 * reference (the line is handled as a C string, main.cu:251,:255).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#include "wordcount_oracle.h"

/* main.cu:297 -- the delimiter literal, byte for byte. */
static const char k_delims[] = " ,.-;:'()\"\t";

/* util.cu:62-67 / :74-85 compare the byte against every delimiter in turn. */
static int is_delim(unsigned char c) {
    const char *d;
    for (d = k_delims; *d; ++d)
        if ((unsigned char)*d == c) return 1;
    return 0;
}

typedef struct { uint64_t off; uint32_t len; } tok_t;

typedef struct {
    tok_t *v; size_t n, cap;
} tokvec_t;

static int tv_push(tokvec_t *t, uint64_t off, uint32_t len) {
    if (t->n == t->cap) {
        size_t nc = t->cap ? t->cap * 2 : 1u << 16;
        tok_t *nv = (tok_t *)realloc(t->v, nc * sizeof(tok_t));
        if (!nv) return -1;
        t->v = nv; t->cap = nc;
    }
    t->v[t->n].off = off; t->v[t->n].len = len; t->n++;
    return 0;
}

/* loadFile (main.cu:247 getline) + map (main.cu:295-311) + my_strtok_r
 * (util.cu:54-89), with the per-line token cap lifted.  When
 * emits_per_line > 0 the cap of main.cu:300-303 is reproduced (tokens beyond
 * it on a line are dropped), which is what lets this be checked against the
 * verbatim binary on inputs where the cap fires. */
static int tokenise(const uint8_t *buf, size_t n, int emits_per_line, tokvec_t *out) {
    size_t i = 0;
    while (i < n) {
        /* one line: [i, e) without the '\n'; a NUL ends the C string early */
        size_t e = i;
        while (e < n && buf[e] != '\n') e++;
        size_t s_end = i;
        while (s_end < e && buf[s_end] != 0) s_end++;
        size_t p = i;
        int emitted = 0;
        for (;;) {
            while (p < s_end && is_delim(buf[p])) p++;          /* util.cu:62-67 */
            if (p >= s_end) break;                               /* util.cu:69-72 */
            size_t q = p;
            while (q < s_end && !is_delim(buf[q])) q++;          /* util.cu:75-88 */
            if (emits_per_line > 0 && emitted >= emits_per_line) break; /* main.cu:300-303 */
            if (q - p > 0xffffffffu) return -2;
            if (tv_push(out, (uint64_t)p, (uint32_t)(q - p))) return -1;
            emitted++;
            p = q;
        }
        i = e + 1;
    }
    return 0;
}

/* KeyValue.h:50-68 restated on views: first differing byte decides; a key that
 * is a prefix of the other sorts first (its terminator, 0, is the smaller byte). */
static inline int key_cmp(const uint8_t *buf, const tok_t *a, const tok_t *b) {
    uint32_t m = a->len < b->len ? a->len : b->len;
    int c = memcmp(buf + a->off, buf + b->off, m);
    if (c) return c;
    return (a->len > b->len) - (a->len < b->len);
}

/* main.cu:498 std::sort -- an introsort-style comparison sort (quicksort with
 * median of three, heapsort when the depth budget runs out, insertion sort for
 * short ranges).  Like std::sort it is not stable; equal keys are
 * indistinguishable here, so the result is the same. */
static void ins_sort(const uint8_t *buf, tok_t *v, size_t n) {
    for (size_t i = 1; i < n; i++) {
        tok_t x = v[i]; size_t j = i;
        while (j > 0 && key_cmp(buf, &x, &v[j - 1]) < 0) { v[j] = v[j - 1]; j--; }
        v[j] = x;
    }
}
static void sift(const uint8_t *buf, tok_t *v, size_t i, size_t n) {
    for (;;) {
        size_t l = 2 * i + 1, r = l + 1, m = i;
        if (l < n && key_cmp(buf, &v[m], &v[l]) < 0) m = l;
        if (r < n && key_cmp(buf, &v[m], &v[r]) < 0) m = r;
        if (m == i) return;
        tok_t t = v[i]; v[i] = v[m]; v[m] = t; i = m;
    }
}
static void heap_sort(const uint8_t *buf, tok_t *v, size_t n) {
    for (size_t i = n / 2; i-- > 0;) sift(buf, v, i, n);
    for (size_t k = n; k-- > 1;) { tok_t t = v[0]; v[0] = v[k]; v[k] = t; sift(buf, v, 0, k); }
}
static void intro(const uint8_t *buf, tok_t *v, size_t n, int depth) {
    while (n > 24) {
        if (depth-- == 0) { heap_sort(buf, v, n); return; }
        size_t mid = n / 2;
        if (key_cmp(buf, &v[mid], &v[0]) < 0) { tok_t t = v[0]; v[0] = v[mid]; v[mid] = t; }
        if (key_cmp(buf, &v[n - 1], &v[0]) < 0) { tok_t t = v[0]; v[0] = v[n - 1]; v[n - 1] = t; }
        if (key_cmp(buf, &v[n - 1], &v[mid]) < 0) { tok_t t = v[mid]; v[mid] = v[n - 1]; v[n - 1] = t; }
        tok_t piv = v[mid];
        size_t i = 0, j = n - 1;
        for (;;) {
            while (key_cmp(buf, &v[i], &piv) < 0) i++;
            while (key_cmp(buf, &piv, &v[j]) < 0) j--;
            if (i >= j) break;
            tok_t t = v[i]; v[i] = v[j]; v[j] = t; i++; j--;
        }
        /* [0, j] and [j+1, n) ; recurse on the smaller, loop on the larger */
        size_t ln = j + 1, rn = n - ln;
        if (ln < rn) { intro(buf, v, ln, depth); v += ln; n = rn; }
        else { intro(buf, v + ln, rn, depth); n = ln; }
    }
    ins_sort(buf, v, n);
}
static void sort_tokens(const uint8_t *buf, tok_t *v, size_t n) {
    int depth = 2; size_t t = n;
    while (t > 1) { depth += 2; t >>= 1; }
    intro(buf, v, n, depth);
}

static int emit_result(const uint8_t *buf, const tok_t *u, const uint64_t *cnt, size_t nu,
                       uint64_t ntok, wc_result *res) {
    uint64_t blob = 0;
    for (size_t i = 0; i < nu; i++) blob += u[i].len;
    res->n = nu; res->tokens = ntok;
    res->key_blob = (uint8_t *)malloc(blob ? blob : 1);
    res->key_offsets = (uint64_t *)malloc((nu + 1) * sizeof(uint64_t));
    res->counts = (uint64_t *)malloc((nu ? nu : 1) * sizeof(uint64_t));
    if (!res->key_blob || !res->key_offsets || !res->counts) return -1;
    uint64_t o = 0;
    for (size_t i = 0; i < nu; i++) {
        res->key_offsets[i] = o;
        memcpy(res->key_blob + o, buf + u[i].off, u[i].len);
        o += u[i].len;
        res->counts[i] = cnt[i];
    }
    res->key_offsets[nu] = o;
    return 0;
}

/* The path as the reference runs it: map -> sort every token -> run-length reduce. */
int wc_oracle_sort(const uint8_t *buf, size_t n, int emits_per_line, wc_result *res) {
    tokvec_t tv = {0, 0, 0};
    memset(res, 0, sizeof(*res));
    int rc = tokenise(buf, n, emits_per_line, &tv);
    if (rc) { free(tv.v); return rc; }
    sort_tokens(buf, tv.v, tv.n);                              /* main.cu:498 */
    /* cpuReduce main.cu:337-355: emit (key, end - start) at every key change */
    size_t nu = 0;
    uint64_t *cnt = (uint64_t *)malloc((tv.n ? tv.n : 1) * sizeof(uint64_t));
    if (!cnt) { free(tv.v); return -1; }
    size_t start = 0;
    for (size_t i = 1; i <= tv.n; i++) {
        if (i == tv.n || key_cmp(buf, &tv.v[start], &tv.v[i]) != 0) {
            if (tv.n) { tv.v[nu] = tv.v[start]; cnt[nu] = (uint64_t)(i - start); nu++; }
            start = i;
        }
    }
    rc = emit_result(buf, tv.v, cnt, nu, (uint64_t)tv.n, res);
    free(cnt); free(tv.v);
    return rc;
}

/* Fast checker for the multi-GB configurations: same tokeniser, counts kept in
 * an open-addressed table keyed by the full key bytes, then the unique keys
 * sorted with the same comparator.  tests/test_oracle.py proves it equal to
 * wc_oracle_sort on every fixture and on random inputs. */
static uint64_t fnv1a(const uint8_t *p, uint32_t len) {
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < len; i++) { h ^= p[i]; h *= 1099511628211ull; }
    h ^= h >> 29; h *= 0xbf58476d1ce4e5b9ull; h ^= h >> 32;
    return h;
}
int wc_oracle_fast(const uint8_t *buf, size_t n, wc_result *res) {
    memset(res, 0, sizeof(*res));
    size_t cap = 1u << 16, used = 0;
    tok_t *keys = (tok_t *)calloc(cap, sizeof(tok_t));
    uint64_t *cnt = (uint64_t *)calloc(cap, sizeof(uint64_t));
    if (!keys || !cnt) return -1;
    uint64_t ntok = 0;
    size_t i = 0;
    while (i < n) {
        size_t e = i;
        while (e < n && buf[e] != '\n') e++;
        size_t s_end = i;
        while (s_end < e && buf[s_end] != 0) s_end++;
        size_t p = i;
        for (;;) {
            while (p < s_end && is_delim(buf[p])) p++;
            if (p >= s_end) break;
            size_t q = p;
            while (q < s_end && !is_delim(buf[q])) q++;
            uint32_t len = (uint32_t)(q - p);
            ntok++;
            uint64_t h = fnv1a(buf + p, len);
            size_t s = (size_t)h & (cap - 1);
            for (;;) {
                if (cnt[s] == 0) { keys[s].off = p; keys[s].len = len; cnt[s] = 1; used++; break; }
                if (keys[s].len == len && memcmp(buf + keys[s].off, buf + p, len) == 0) { cnt[s]++; break; }
                s = (s + 1) & (cap - 1);
            }
            if (used * 2 > cap) {                                   /* grow x4 and rehash */
                size_t nc = cap * 4;
                tok_t *nk = (tok_t *)calloc(nc, sizeof(tok_t));
                uint64_t *ncnt = (uint64_t *)calloc(nc, sizeof(uint64_t));
                if (!nk || !ncnt) return -1;
                for (size_t j = 0; j < cap; j++) if (cnt[j]) {
                    size_t t = (size_t)fnv1a(buf + keys[j].off, keys[j].len) & (nc - 1);
                    while (ncnt[t]) t = (t + 1) & (nc - 1);
                    nk[t] = keys[j]; ncnt[t] = cnt[j];
                }
                free(keys); free(cnt); keys = nk; cnt = ncnt; cap = nc;
            }
            p = q;
        }
        i = e + 1;
    }
    /* compact, then sort (key, count) pairs together by key */
    tok_t *u = (tok_t *)malloc((used ? used : 1) * sizeof(tok_t));
    uint64_t *uc = (uint64_t *)malloc((used ? used : 1) * sizeof(uint64_t));
    if (!u || !uc) return -1;
    size_t k = 0;
    for (size_t j = 0; j < cap; j++) if (cnt[j]) { u[k] = keys[j]; k++; }
    sort_tokens(buf, u, used);
    for (size_t j = 0; j < used; j++) {
        uint64_t h = fnv1a(buf + u[j].off, u[j].len);
        size_t s = (size_t)h & (cap - 1);
        while (!(keys[s].len == u[j].len && memcmp(buf + keys[s].off, buf + u[j].off, u[j].len) == 0))
            s = (s + 1) & (cap - 1);
        uc[j] = cnt[s];
    }
    int rc = emit_result(buf, u, uc, used, ntok, res);
    free(u); free(uc); free(keys); free(cnt);
    return rc;
}

void wc_result_free(wc_result *res) {
    free(res->key_blob); free(res->key_offsets); free(res->counts);
    memset(res, 0, sizeof(*res));
}

/* printKeyValues main.cu:286, CPU-branch line format. */
size_t wc_format_cpu(const wc_result *res, char *out, size_t cap) {
    size_t w = 0;
    for (uint64_t i = 0; i < res->n; i++) {
        uint64_t kl = res->key_offsets[i + 1] - res->key_offsets[i];
        char num[32];
        int nl = snprintf(num, sizeof num, "%llu", (unsigned long long)res->counts[i]);
        size_t need = 11 + kl + 10 + (size_t)nl + 1;
        if (out && w + need <= cap) {
            memcpy(out + w, "print key: ", 11);
            memcpy(out + w + 11, res->key_blob + res->key_offsets[i], kl);
            memcpy(out + w + 11 + kl, " \t value: ", 10);
            memcpy(out + w + 21 + kl, num, (size_t)nl);
            out[w + 21 + kl + (size_t)nl] = '\n';
        }
        w += need;
    }
    return w;
}

/* canonical "word\tcount\n" */
size_t wc_format_tsv(const wc_result *res, char *out, size_t cap) {
    size_t w = 0;
    for (uint64_t i = 0; i < res->n; i++) {
        uint64_t kl = res->key_offsets[i + 1] - res->key_offsets[i];
        char num[32];
        int nl = snprintf(num, sizeof num, "%llu", (unsigned long long)res->counts[i]);
        size_t need = kl + 1 + (size_t)nl + 1;
        if (out && w + need <= cap) {
            memcpy(out + w, res->key_blob + res->key_offsets[i], kl);
            out[w + kl] = '\t';
            memcpy(out + w + kl + 1, num, (size_t)nl);
            out[w + kl + 1 + (size_t)nl] = '\n';
        }
        w += need;
    }
    return w;
}

#ifdef WC_ORACLE_MAIN
#include <time.h>
static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: wc_oracle <file> [sort|fast] [cpu|tsv|none] [emits_per_line]\n"); return 2; }
    const char *mode = argc > 2 ? argv[2] : "sort";
    const char *fmt = argc > 3 ? argv[3] : "cpu";
    int cap = argc > 4 ? atoi(argv[4]) : 0;
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t *buf = (uint8_t *)malloc(sz ? (size_t)sz : 1);
    if (fread(buf, 1, (size_t)sz, f) != (size_t)sz) { fprintf(stderr, "short read\n"); return 1; }
    fclose(f);
    wc_result r;
    double t0 = now_s();
    int rc = strcmp(mode, "fast") == 0 ? wc_oracle_fast(buf, (size_t)sz, &r) : wc_oracle_sort(buf, (size_t)sz, cap, &r);
    double t1 = now_s();
    if (rc) { fprintf(stderr, "oracle failed: %d\n", rc); return 1; }
    fprintf(stderr, "oracle %s: %ld bytes, %llu tokens, %llu uniques, %.3f s\n", mode, sz,
            (unsigned long long)r.tokens, (unsigned long long)r.n, t1 - t0);
    if (strcmp(fmt, "none") != 0) {
        size_t need = strcmp(fmt, "tsv") == 0 ? wc_format_tsv(&r, NULL, 0) : wc_format_cpu(&r, NULL, 0);
        char *o = (char *)malloc(need ? need : 1);
        if (strcmp(fmt, "tsv") == 0) wc_format_tsv(&r, o, need); else wc_format_cpu(&r, o, need);
        fwrite(o, 1, need, stdout);
        free(o);
    }
    wc_result_free(&r); free(buf);
    return 0;
}
#endif
