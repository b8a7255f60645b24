/* TEST INFRASTRUCTURE -- not part of the product.  See wordcount_oracle.c. */
#ifndef WORDCOUNT_ORACLE_H
#define WORDCOUNT_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* n unique keys in ascending byte order; key i is
 * key_blob[key_offsets[i] .. key_offsets[i+1]); counts[i] occurrences. */
typedef struct {
    uint64_t n;
    uint64_t tokens;
    uint8_t *key_blob;
    uint64_t *key_offsets;   /* n + 1 entries */
    uint64_t *counts;        /* n entries */
} wc_result;

/* map -> comparison sort of every token -> run-length reduce (the reference's
 * CPU stages, main.cu:495-504).  emits_per_line = 0 lifts the per-line cap;
 * 20 reproduces main.cu:19. */
int wc_oracle_sort(const uint8_t *buf, size_t n, int emits_per_line, wc_result *res);
/* same answer through a hash table; for multi-GB inputs */
int wc_oracle_fast(const uint8_t *buf, size_t n, wc_result *res);
void wc_result_free(wc_result *res);
/* "print key: %s \t value: %s\n" lines (main.cu:286); returns bytes needed */
size_t wc_format_cpu(const wc_result *res, char *out, size_t cap);
/* canonical "word\tcount\n" lines */
size_t wc_format_tsv(const wc_result *res, char *out, size_t cap);

/* Second job (oracle/pagerank_oracle.c): PageRank over "src dst" lines in 64-bit fixed point; result
 * records are (node, rank) in key order, res->tokens = number of edges.  0 ok, 1 out of memory, 2 a node
 * name longer than 16 bytes. */
int pr_oracle(const uint8_t *buf, size_t n, uint32_t iterations, uint64_t damping_q32, wc_result *res);

#ifdef __cplusplus
}
#endif
#endif
