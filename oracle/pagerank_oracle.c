/*
 * TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * CPU statement of the second MapReduce job of locust-b200: PageRank over an edge list, in 64-bit
 * fixed point, so that results are exact integers and independent of summation order.
 *
 * Parity status: UNPINNED.  The reference names "PageRank or similar algorithm that has more work per
 * thread than word count" as a milestone (docs/PROPOSAL.md:21; README.md:6) but holds no code, test or
 * golden vector for it.  This file is therefore the definition of the job, not a restatement; the GPU
 * path (locust_pagerank) is checked against it bit for bit.
 *
 * The job, on the reference's operator surface (lines of text in, (key, value) records in strcmp order out):
 *   input     one edge per line, "src dst": the first two tokens of the line under the word-count delimiter
 *             set (main.cu:297); lines with fewer than two tokens are skipped, further tokens ignored;
 *             a repeated edge counts as often as it appears; node names have at most 16 bytes
 *   map       (degrees)  (src, 1) per edge; every node, also one that only receives, gets a record
 *   iterate   M = 2^48 is the whole rank mass, N the number of nodes, D the damping factor in Q32.
 *             rank_0[v] = M / N.  Per iteration:
 *               contrib[u] = outdeg[u] ? rank[u] / outdeg[u] : 0
 *               dangling   = sum of rank[u] over nodes without out-edges
 *               acc[v]     = sum of contrib[u] over edges u -> v
 *               teleport   = ((M - mulq32(M, D)) + mulq32(dangling, D)) / N
 *               rank'[v]   = teleport + mulq32(acc[v], D)           mulq32(x, D) = (x * D) >> 32, 128-bit product
 *   reduce    records (node, rank) in ascending byte order of the node names
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "wordcount_oracle.h"

static int pr_is_delim(unsigned char c) {
    static const char d[] = " ,.-;:'()\"\t";
    return c != 0 && strchr(d, c) != NULL;
}

typedef struct { uint8_t key[16]; uint32_t len; uint64_t outdeg, rank, acc, contrib; } pr_node;

static uint64_t pr_hash(const uint8_t *k, uint32_t len) {
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < len; ++i) { h ^= k[i]; h *= 1099511628211ull; }
    return h;
}

static uint64_t mulq32(uint64_t x, uint64_t d) { return (uint64_t)(((__uint128_t)x * d) >> 32); }

static int key_less(const pr_node *a, const pr_node *b) {
    const uint32_t m = a->len < b->len ? a->len : b->len;
    const int c = memcmp(a->key, b->key, m);
    return c ? c < 0 : a->len < b->len;
}
static int node_cmp(const void *a, const void *b) {
    const pr_node *x = (const pr_node *)a, *y = (const pr_node *)b;
    return key_less(x, y) ? -1 : (key_less(y, x) ? 1 : 0);
}

/* returns 0, 1 (out of memory) or 2 (a node name longer than 16 bytes) */
int pr_oracle(const uint8_t *buf, size_t n, uint32_t iterations, uint64_t damping_q32, wc_result *res) {
    memset(res, 0, sizeof *res);
    size_t cap = 1024, used = 0, ecap = 1024, ne = 0;
    pr_node *nodes = NULL;
    uint32_t *tab = (uint32_t *)calloc(cap * 2, sizeof(uint32_t));        /* node index + 1 */
    uint32_t *es = (uint32_t *)malloc(ecap * sizeof(uint32_t)), *ed = (uint32_t *)malloc(ecap * sizeof(uint32_t));
    size_t ncap = 1024;
    nodes = (pr_node *)malloc(ncap * sizeof(pr_node));
    if (!tab || !es || !ed || !nodes) return 1;
    size_t tslots = cap * 2;
    size_t pos = 0;
    while (pos < n) {
        size_t eol = pos;
        while (eol < n && buf[eol] != '\n') ++eol;
        /* first two tokens of buf[pos, eol) */
        size_t p = pos;
        uint32_t idx[2];
        int got = 0;
        while (got < 2) {
            while (p < eol && pr_is_delim(buf[p])) ++p;
            if (p >= eol) break;
            size_t q = p;
            while (q < eol && !pr_is_delim(buf[q])) ++q;
            if (q - p > 16) { free(tab); free(es); free(ed); free(nodes); return 2; }
            /* find or add */
            const uint32_t len = (uint32_t)(q - p);
            size_t s = pr_hash(buf + p, len) & (tslots - 1);
            for (;;) {
                if (tab[s] == 0) {
                    if (used == ncap) { ncap *= 2; nodes = (pr_node *)realloc(nodes, ncap * sizeof(pr_node)); if (!nodes) return 1; }
                    memset(&nodes[used], 0, sizeof(pr_node));
                    memcpy(nodes[used].key, buf + p, len);
                    nodes[used].len = len;
                    tab[s] = (uint32_t)(used + 1);
                    ++used;
                    break;
                }
                const pr_node *c = &nodes[tab[s] - 1];
                if (c->len == len && memcmp(c->key, buf + p, len) == 0) break;
                s = (s + 1) & (tslots - 1);
            }
            idx[got++] = tab[s] - 1;
            if (used * 2 > tslots) {            /* grow the index */
                tslots *= 2;
                free(tab);
                tab = (uint32_t *)calloc(tslots, sizeof(uint32_t));
                if (!tab) return 1;
                for (size_t i = 0; i < used; ++i) {
                    size_t t = pr_hash(nodes[i].key, nodes[i].len) & (tslots - 1);
                    while (tab[t]) t = (t + 1) & (tslots - 1);
                    tab[t] = (uint32_t)(i + 1);
                }
            }
            p = q;
        }
        if (got == 2) {
            if (ne == ecap) {
                ecap *= 2;
                es = (uint32_t *)realloc(es, ecap * sizeof(uint32_t)); ed = (uint32_t *)realloc(ed, ecap * sizeof(uint32_t));
                if (!es || !ed) return 1;
            }
            es[ne] = idx[0]; ed[ne] = idx[1]; ++ne;
            nodes[idx[0]].outdeg += 1;
        } else if (got == 1) {
            /* a line with one token names no edge; its token is not a node either: undo nothing -- the node was
             * added to the index above, so take it out again only if nothing else refers to it.  Simpler and
             * equivalent: such lines are not allowed to introduce nodes, so mark and drop below. */
        }
        pos = eol + 1;
    }
    /* nodes that appear only on lines without a second token are not part of the graph */
    uint8_t *live = (uint8_t *)calloc(used ? used : 1, 1);
    if (!live) return 1;
    for (size_t e = 0; e < ne; ++e) { live[es[e]] = 1; live[ed[e]] = 1; }
    size_t N = 0;
    for (size_t i = 0; i < used; ++i) N += live[i];
    const uint64_t M = 1ull << 48;
    for (size_t i = 0; i < used; ++i) nodes[i].rank = (live[i] && N) ? M / N : 0;
    for (uint32_t it = 0; it < iterations && N; ++it) {
        uint64_t dangling = 0;
        for (size_t i = 0; i < used; ++i) {
            if (!live[i]) continue;
            nodes[i].contrib = nodes[i].outdeg ? nodes[i].rank / nodes[i].outdeg : 0;
            if (!nodes[i].outdeg) dangling += nodes[i].rank;
            nodes[i].acc = 0;
        }
        for (size_t e = 0; e < ne; ++e) nodes[ed[e]].acc += nodes[es[e]].contrib;
        const uint64_t teleport = ((M - mulq32(M, damping_q32)) + mulq32(dangling, damping_q32)) / N;
        for (size_t i = 0; i < used; ++i)
            if (live[i]) nodes[i].rank = teleport + mulq32(nodes[i].acc, damping_q32);
    }
    /* records in key order */
    pr_node *out = (pr_node *)malloc((N ? N : 1) * sizeof(pr_node));
    if (!out) return 1;
    size_t k = 0;
    for (size_t i = 0; i < used; ++i) if (live[i]) out[k++] = nodes[i];
    qsort(out, N, sizeof(pr_node), node_cmp);
    res->n = N;
    res->tokens = ne;                               /* edges */
    res->key_offsets = (uint64_t *)malloc((N + 1) * sizeof(uint64_t));
    res->counts = (uint64_t *)malloc((N ? N : 1) * sizeof(uint64_t));
    uint64_t blob = 0;
    for (size_t i = 0; i < N; ++i) blob += out[i].len;
    res->key_blob = (uint8_t *)malloc(blob ? blob : 1);
    if (!res->key_offsets || !res->counts || !res->key_blob) return 1;
    blob = 0;
    for (size_t i = 0; i < N; ++i) {
        res->key_offsets[i] = blob;
        memcpy(res->key_blob + blob, out[i].key, out[i].len);
        blob += out[i].len;
        res->counts[i] = out[i].rank;
    }
    res->key_offsets[N] = blob;
    free(out); free(live); free(tab); free(es); free(ed); free(nodes);
    return 0;
}
