/* textgen.h -- deterministic synthetic text for the benchmark configurations of
 * BASELINE.json.  Test and bench infrastructure, built as its
 * own library (liblocust_textgen.so); liblocust.so does not depend on it.
 *
 * The text is a sequence of independent 4096-byte blocks: block b depends only on
 * (seed, b), never on its neighbours, and always ends in a delimiter.  The same
 * block function runs on the host and in a CUDA kernel, so a rank can generate
 * exactly its byte range of a multi-GB text on its own GPU, and tests can check
 * device output against host output byte for byte.
 */
#ifndef LOCUST_TEXTGEN_H
#define LOCUST_TEXTGEN_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TEXTGEN_BLOCK 4096u

#define TEXTGEN_ZIPF     0   /* C2 / C3: Zipf(s) over a vocabulary of English-like words   */
#define TEXTGEN_UNIFORM  1   /* C4: uniform over `vocab` distinct words of 8-10 letters    */
#define TEXTGEN_SKEWED   2   /* C5: heavy hitters in bursts, very long lines, Zipf tail    */

typedef struct textgen textgen;

/* device < 0: host only.  total_bytes (multiple of TEXTGEN_BLOCK) is the size of the
 * whole text; TEXTGEN_SKEWED uses it to place its regions. */
int textgen_create(int mode, uint32_t vocab, double zipf_s, uint64_t seed, uint64_t vocab_seed,
                   uint64_t total_bytes, int device, textgen **out);
void textgen_destroy(textgen *g);
/* bytes [byte_offset, byte_offset + nbytes) of the text; both multiples of TEXTGEN_BLOCK */
int textgen_fill_host(textgen *g, uint8_t *dst, uint64_t nbytes, uint64_t byte_offset);
int textgen_fill_device(textgen *g, uint8_t *dev_dst, uint64_t nbytes, uint64_t byte_offset, void *stream);
const char *textgen_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
