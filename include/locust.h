/*
 * locust.h -- C ABI of liblocust.so, the B200-native word-count path of
 * wuyan33/Locust ("MapReduce" binary, map -> process -> reduce).
 *
 * The reference has no FFI layer: its operator surface is the stage split inside
 * main() (MapReduce/src/main.cu:358-532).  Each entry point below names the
 * reference code it stands in for.  All entry points are plain C: opaque handle,
 * pointers and sizes, an int status, no C++ types and no exceptions.  Nothing in
 * here exits the process; nothing in here computes on the CPU -- if the CUDA
 * device is missing every call fails with LOCUST_E_CUDA.
 *
 * Record contract (KeyIntValuePair, KeyValue.h:13-18, with the fixed widths
 * lifted): the result is n records (key bytes, count) in ascending unsigned byte
 * order ("strcmp order", a key that is a prefix of another sorts first), keys in
 * one blob addressed by n+1 offsets, counts 64-bit.
 *
 * Input contract: bytes are split at the reference's delimiter literal
 * " ,.-;:'()\"\t" (main.cu:138,:150) plus '\n' (getline, main.cu:43).  Everything
 * else, including '\r' and bytes >= 0x80, is token content.  The input must not
 * contain NUL bytes (the reference treats each line as a C string).
 */
#ifndef LOCUST_H
#define LOCUST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes (the reference checks no status at all) */
#define LOCUST_OK          0
#define LOCUST_E_ARG       1   /* bad argument / NULL pointer                   */
#define LOCUST_E_CUDA      2   /* a CUDA runtime call failed (text in last_error) */
#define LOCUST_E_CAPACITY  3   /* a fixed-size buffer was too small even after growth */
#define LOCUST_E_STATE     4   /* stage called out of order                      */
#define LOCUST_E_NOMEM     5   /* host allocation failed                         */

/* where the bytes handed to locust_map live */
#define LOCUST_HOST    0
#define LOCUST_DEVICE  1

/* stages, for locust_stage_ms (the three chrono intervals of main.cu:405-418,:453-468) */
#define LOCUST_STAGE_MAP      0
#define LOCUST_STAGE_PROCESS  1
#define LOCUST_STAGE_REDUCE   2
#define LOCUST_STAGE_SHUFFLE  3   /* multi-GPU only: first locust_shuffle_sizes .. last locust_shuffle_merge */

/* config flags */
#define LOCUST_F_GENERIC_TOKENISER  1u  /* test switch: byte-serial token path for every token */
#define LOCUST_F_NO_SMEM_COMBINER   2u  /* test switch: every token goes to the global table     */
#define LOCUST_F_WEAK_HASH          4u  /* test switch: 8-bit hashes, forces probe collisions    */
#define LOCUST_F_SORT_PATH          8u  /* alternative path: no combiner; every token becomes a
                                           record, PROCESS radix-sorts them all and REDUCE
                                           run-length counts them -- what the reference does
                                           (main.cu:406-465) without its fixed capacities.  Needs
                                           table_log2 >= log2(tokens); slower, kept for comparison */

typedef struct locust_ctx locust_ctx;

typedef struct locust_config {
    int32_t  device;         /* CUDA device ordinal                                             */
    uint32_t table_log2;     /* global combiner table holds 1 << table_log2 slots of 32 B;
                                0 = default (20).  It grows by itself when it fills.            */
    uint32_t flags;          /* LOCUST_F_*                                                      */
    uint32_t reserved;
    uint64_t long_capacity;  /* occurrences of tokens longer than 16 bytes the job may hold
                                before the buffer is grown; 0 = default (1 << 20)               */
    void    *stream;         /* cudaStream_t to run on; NULL = a private non-blocking stream    */
} locust_config;

/* Replaces the allocations at the top of main() (main.cu:392-393, :402, :451):
 * creates the per-device context that owns every device buffer of the path. */
int locust_create(const locust_config *cfg, locust_ctx **out);

/* Replaces the cudaFree/free tail of main() (main.cu:477-485). */
int locust_destroy(locust_ctx *ctx);

/* Forget the current job (counts, result) but keep the buffers: the equivalent of
 * starting the reference binary again on another file. */
int locust_reset(locust_ctx *ctx);

/* MAP.  Replaces loadFile's record building + kernMap/map (main.cu:40-64,
 * :136-159, launch :406) and the cudaMemcpy in front of it (:403): tokenises
 * `nbytes` bytes and combines (word, 1) pairs into per-device (word, count)
 * partials.  May be called several times before locust_process; every call must
 * begin at a token boundary (start of input or right after a delimiter).
 * LOCUST_DEVICE: `data` is a device pointer on ctx's device and must stay valid
 * and unchanged until locust_reduce has returned (long keys are kept as pointers
 * into it).  LOCUST_HOST: the bytes are copied to the device in chunks that
 * overlap with tokenising; pinned host memory makes that copy asynchronous. */
int locust_map(locust_ctx *ctx, const void *data, uint64_t nbytes, int where);

/* Blocks until every LOCUST_HOST buffer handed to locust_map so far has been copied to the device:
 * from then on the caller may overwrite those buffers (a reader that refills a ring of pinned
 * chunks while earlier chunks are tokenised, as loadFile + cudaMemcpy would if they overlapped:
 * main.cu:242-259, :403).  Tokenising itself keeps running. */
int locust_wait_input(locust_ctx *ctx);

/* PROCESS.  Replaces thrust::partition + thrust::sort with KIVComparator
 * (main.cu:411-415; KeyValue.h:20-33, :79-84): compacts the combined records
 * out of the table and sorts them by key bytes. */
int locust_process(locust_ctx *ctx);

/* REDUCE.  Replaces kernFindUniqBool -> thrust::partition -> kernGetCount
 * (main.cu:161-238, host sequence :447-465): run-length merge over the sorted
 * records, final counts, `val` (exclusive prefix sum of counts, the value the
 * reference prints as "val", main.cu:132) and the key blob. */
int locust_reduce(locust_ctx *ctx);

/* map + process + reduce in one call (single mode, main.cu:397-468). */
int locust_wordcount(locust_ctx *ctx, const void *data, uint64_t nbytes, int where);

/* A second MapReduce job on the same machinery: PageRank over an edge list.  The reference names it as a
 * milestone ("PageRank or similar algorithm that has more work per thread than word count", docs/PROPOSAL.md:21,
 * README.md:6) and holds no code for it, so there is no reference interface to replace; the job is defined by
 * oracle/pagerank_oracle.c (parity unpinned) and follows the word-count conventions:
 *   input   one edge per line, "src dst" = the first two tokens of the line under the word-count delimiter set
 *           (main.cu:297); lines with fewer tokens are skipped, further tokens ignored, repeated edges count;
 *           node names have at most 16 bytes (longer: LOCUST_E_ARG)
 *   ranks   64-bit fixed point, the whole mass is 2^48; damping_q32 = damping * 2^32 (0.85 -> 3650722201);
 *           rank0 = 2^48 / N; per iteration rank'[v] = teleport + (D * sum over u->v of rank[u] / outdeg[u]) >> 32,
 *           teleport = ((2^48 - (2^48 * D >> 32)) + (D * mass of nodes without out-edges >> 32)) / N.
 *           All sums are integer, so the result is independent of their order.
 *   result  read with locust_result_host / _device / locust_format_result: keys = node names in byte order,
 *           counts = ranks, vals = their exclusive prefix sums.
 * Replaces whatever job was in the context (like locust_wordcount).  Grows the table when the graph has more
 * nodes than half its slots. */
int locust_pagerank(locust_ctx *ctx, const void *edges, uint64_t nbytes, int where, uint32_t iterations,
                    uint64_t damping_q32);
/* nodes and edges of the last locust_pagerank */
int locust_pagerank_stats(locust_ctx *ctx, uint64_t *nodes, uint64_t *edges);

/* Replaces the cudaMemcpy device->host of the result (main.cu:470-472).  Pointers
 * are borrowed: valid until the next locust_reset / locust_map / locust_destroy.
 * offsets has n + 1 entries; vals may be NULL. */
int locust_result_host(locust_ctx *ctx, const uint8_t **key_blob, const uint64_t **key_offsets,
                       const uint64_t **counts, const uint64_t **vals, uint64_t *n);

/* Same records, left in device memory (for callers that keep going on the GPU). */
int locust_result_device(locust_ctx *ctx, const uint8_t **key_blob, const uint64_t **key_offsets,
                         const uint64_t **counts, const uint64_t **vals, uint64_t *n);

/* printKeyIntValues (main.cu:126-134, line format :132)
 *   LOCUST_FORMAT_GPU   "print key: %s \t val: %d \t count: %d\n"
 * printKeyValues of the CPU branch (main.cu:286)
 *   LOCUST_FORMAT_CPU   "print key: %s \t value: %s\n"
 * writeKeyIntValues, the intermediate file of multi mode (main.cu:116-124, :121), with
 * the combined count as the value
 *   LOCUST_FORMAT_INTERMEDIATE   "%s \t%d\n".
 * Lines are formatted on the device and copied out in one piece.  Returns the
 * text through *text (borrowed, host memory, not NUL-terminated) and *nbytes. */
#define LOCUST_FORMAT_GPU           0
#define LOCUST_FORMAT_CPU           1
#define LOCUST_FORMAT_INTERMEDIATE  2
int locust_format_result(locust_ctx *ctx, int format, const char **text, uint64_t *nbytes);

/* Replaces loadIntermediateFile (main.cu:66-103): adds the records of an intermediate
 * text ("key \tvalue\n" lines, host memory) to the job, summing the values per key.  A
 * file written by the reference's stage 1 has one line per token occurrence with value 1;
 * a file written with LOCUST_FORMAT_INTERMEDIATE has one line per word with its count.
 * Both reduce to the same result.  The single space the reference writes in front of the
 * TAB is not part of the key (the reference keeps it, main.cu:94: its stage-2 keys end in a
 * blank).  Lines are parsed on the host, the records are merged on the GPU.  Can be mixed
 * with locust_map calls; follow with locust_process / locust_reduce. */
int locust_map_intermediate(locust_ctx *ctx, const char *text, uint64_t nbytes);

/* Device time of a stage in milliseconds (CUDA events on the context's stream):
 * what main.cu:408,:418,:468 print, but measured on the device. */
float locust_stage_ms(locust_ctx *ctx, int stage);

/* The cudaStream_t every call of this context enqueues its work on (cfg.stream, or the
 * private stream locust_create made).  A caller that mixes liblocust calls with its own
 * device work (a collective between locust_shuffle_pack and locust_shuffle_merge, say) issues
 * that work on this stream, or orders its own stream against it with events.  No counterpart
 * in the reference, which runs everything on the default stream (main.cu:403-472). */
void *locust_stream(locust_ctx *ctx);

/* Number of tokens counted (sum of all counts) and of kernels launched by the job
 * so far. */
int locust_stats(locust_ctx *ctx, uint64_t *tokens, uint64_t *unique_keys, uint64_t *kernel_launches);

/* Text of the last error on this context ("" if none). */
const char *locust_last_error(locust_ctx *ctx);

/* ---- multi-GPU shuffle: replaces writeKeyIntValues -> /tmp/out.txt ->
 * loadIntermediateFile (main.cu:116-124, :428-430, :66-103, :441) -------------- */

/* After locust_map: splits this device's partial (word, count) records by
 * hash(word) % nparts and reports how many bytes each part packs to.
 * bytes_per_part: host array of nparts. */
int locust_shuffle_sizes(locust_ctx *ctx, uint32_t nparts, uint64_t *bytes_per_part);

/* The two numbers behind each bytes_per_part of the last locust_shuffle_sizes call: short
 * records (keys of <= 16 bytes, 32 bytes each) and bytes of long records per part.  A caller
 * that sizes the receive slots of locust_shuffle_pack_peers from the counts instead of from
 * the sender's table size needs them separately (short_cap, and slot_bytes - 32 - 32 * short_cap). */
int locust_shuffle_counts(locust_ctx *ctx, uint32_t nparts, uint64_t *short_records, uint64_t *long_bytes);

/* Packs part p at dev_out + part_offsets[p] (device memory, caller-owned; offsets
 * and sizes as reported by locust_shuffle_sizes, each a multiple of 32). */
int locust_shuffle_pack(locust_ctx *ctx, uint32_t nparts, void *dev_out, const uint64_t *part_offsets);

/* ---- pack and send in one kernel, over peer memory (NVLink) ---------------------------------
 * Instead of packing into a local buffer and handing that to a collective, part p is written
 * straight into a slot of rank p's receive buffer by the packing kernel itself.  No sizes are
 * exchanged: a slot is `slot_bytes` long, holds up to `short_cap` short records at byte 32 and
 * long records behind them, and its 32-byte header (written last) says how much arrived.  The
 * caller separates pack and merge by a barrier over all ranks on the context's stream (e.g. a
 * one-element all-reduce), then merges every slot of its own buffer with locust_shuffle_merge
 * (dev_records = slot, nbytes = slot_bytes).  short_cap = the sender's table size cannot
 * overflow; long records that do not fit make locust_reduce fail with LOCUST_E_CAPACITY.
 * Note: MAP's own buffers are not re-grown on this path -- size table_log2 for the shard.
 *
 * locust_exchange_alloc: a zeroed receive buffer on ctx's device, owned by ctx.
 * locust_ipc_export / _open / _close: the 64-byte handle another process on the same box needs
 * to map that buffer (cudaIpcGetMemHandle / OpenMemHandle / CloseMemHandle). */
int locust_exchange_alloc(locust_ctx *ctx, uint64_t nbytes, void **dev_buf);
/* Gives a buffer of locust_exchange_alloc back before the context is destroyed (a caller that re-sizes
 * its receive slots); peers must have closed their mappings of it. */
int locust_exchange_free(locust_ctx *ctx, void *dev_buf);
int locust_ipc_export(locust_ctx *ctx, const void *dev_buf, uint8_t handle[64]);
int locust_ipc_open(locust_ctx *ctx, const uint8_t handle[64], void **dev_buf);
int locust_ipc_close(locust_ctx *ctx, void *dev_buf);
int locust_shuffle_pack_peers(locust_ctx *ctx, uint32_t nparts, void *const *dst_slots, uint64_t slot_bytes,
                              uint64_t short_cap);

/* Drops the local partials (they have been packed) and starts the final table. */
int locust_shuffle_begin_merge(locust_ctx *ctx);

/* Adds packed records received from a peer (device memory; must stay valid until
 * locust_reduce has returned). */
int locust_shuffle_merge(locust_ctx *ctx, const void *dev_records, uint64_t nbytes);

/* ---- ordered shuffle: ranks end up with consecutive key ranges --------------------------
 * The reference's master node sorts everything again after reading its nodes' partial
 * outputs (loadIntermediateFile + thrust::sort, main.cu:441, :415).  With parts that are key
 * ranges instead of hash classes that last sort is not needed: rank r sorts and reduces range
 * r alone, and the whole result is the ranks' results one after the other.
 *
 * After locust_map on every rank:
 *   locust_partition_histogram adds, into dev_hist (device memory, uint64[LOCUST_PARTITION_BINS],
 *     zeroed by the caller), how many of this rank's keys start with each two-byte prefix (an
 *     estimate from every eighth slot when the table is large: the counts only steer the balance);
 *   the caller sums dev_hist over all ranks (all-reduce);
 *   locust_partition_by_range builds the prefix -> part map from the summed histogram (the same
 *     on every rank): ranges follow key order and hold about the same number of keys, as far as
 *     whole prefixes allow.  From then on locust_shuffle_sizes / _pack / _pack_peers with the
 *     same nparts split by range instead of by hash, until locust_reset (or nparts = 0 here).
 * After the exchange, locust_process and locust_reduce on every rank:
 *   locust_result_part_bytes / locust_result_part_pack move a rank's result into one buffer
 *     {n, key_bytes, total, part_bytes | key_offsets[n+1] | counts[n] | vals[n] | key bytes};
 *   locust_result_concat, on the rank that wants the whole result, replaces its own result by
 *     the given parts in order (its own part included), rebasing offsets and vals.  n_keys,
 *     blob_bytes, totals: the first three header words of each part, host arrays.
 *
 * On one box the parts need not be sent by a collective: the receiver owns a buffer of one
 * slot per rank (locust_exchange_alloc, mapped by the others with locust_ipc_*), every rank
 * packs its part straight into its slot over NVLink (locust_result_part_pack with the slot's
 * peer address; a rank whose part is larger than the slot writes only its header there,
 * locust_result_header), and after a barrier the receiver strings the slots together with
 * locust_result_adopt_slots (LOCUST_E_CAPACITY if a part did not fit). */
#define LOCUST_PARTITION_BINS 65536u
int locust_partition_histogram(locust_ctx *ctx, void *dev_hist);
int locust_partition_by_range(locust_ctx *ctx, const void *dev_hist, uint32_t nparts);
int locust_result_part_bytes(locust_ctx *ctx, uint64_t *bytes);
int locust_result_part_pack(locust_ctx *ctx, void *dev_out);
int locust_result_concat(locust_ctx *ctx, uint32_t nparts, const void *const *dev_parts, const uint64_t *n_keys,
                         const uint64_t *blob_bytes, const uint64_t *totals);
int locust_result_header(locust_ctx *ctx, void *dev_header);
int locust_result_adopt_slots(locust_ctx *ctx, const void *dev_slots, uint64_t slot_bytes, uint32_t nslots);

/* Small helpers over the same peer mappings, so that a job on one box needs no collective call:
 * locust_peer_put copies nbytes (multiple of 16) from dev_src to each of the ndst destinations
 *   (a rank's histogram into its slot on every rank);
 * locust_partition_by_range_sum is locust_partition_by_range over nhists histograms that lie
 *   one after the other, added up on the fly (the slots the ranks put theirs into);
 * locust_peer_barrier orders the ranks' streams: peer_flags[p] is rank p's array of `world`
 *   uint32 flags (zero before first use, mapped by every rank); work queued after it on this
 *   stream starts only when every rank's stream has reached its barrier with the same epoch
 *   (1, 2, 3, ... per use).  It waits on the GPU, not on the host; after five seconds without
 *   a peer it gives up and the job fails at locust_reduce / locust_result_adopt_slots. */
int locust_peer_put(locust_ctx *ctx, uint32_t ndst, void *const *dsts, const void *dev_src, uint64_t nbytes);
int locust_partition_by_range_sum(locust_ctx *ctx, const void *dev_hists, uint32_t nhists, uint32_t nparts);
int locust_peer_barrier(locust_ctx *ctx, void *const *peer_flags, uint32_t rank, uint32_t world, uint32_t epoch);

/* ---- several GPUs of one box from one process -----------------------------------
 * Replaces the reference's multi mode as a whole: the per-node line windows and stage
 * arguments of its command line (main.cu:367-385), the sorted partials handed over in
 * /tmp/out.txt (:421-446) and the command runner that starts the nodes
 * (Distributor/slave.py:30-32).  One call runs the job on every listed GPU: shards cut on
 * delimiters, map on every GPU, two-byte-prefix histograms and range map, records packed
 * straight into the owning GPU's memory (NVLink peer stores; receive slots sized from a
 * count pass, so a table that overflowed in MAP still grows), merge + PROCESS + REDUCE per
 * key range, sorted slices strung together on devices[0].  One host thread per GPU; the
 * streams meet through events, never through kernels that wait for each other, so the same
 * device may be listed several times (devices = {0, 0, 0} runs three ranks on GPU 0).
 *
 *   locust_multi_create            contexts, streams and peer access for devices[0..n)
 *                                  (cfg->device and cfg->stream are ignored; NULL cfg = defaults)
 *   locust_multi_wordcount         host_data[0..nbytes): cut, copied and counted; the result is
 *                                  on locust_multi_context(m, 0)
 *   locust_multi_wordcount_device  shard r already lies in device memory of devices[r]; every shard
 *                                  must begin at a token boundary and stay valid until the next job
 *   locust_multi_context           the per-rank context: rank 0 for locust_result_host /
 *                                  locust_result_device / locust_format_result after a job, any rank
 *                                  for locust_stage_ms / locust_stats
 *   locust_multi_stats             device time of the last job on rank 0's stream (first kernel to
 *                                  result in place, ms) and the record bytes that left their GPU */
typedef struct locust_multi locust_multi;
int locust_multi_create(const int *devices, int n, const locust_config *cfg, locust_multi **out);
int locust_multi_destroy(locust_multi *m);
int locust_multi_wordcount(locust_multi *m, const void *host_data, uint64_t nbytes);
int locust_multi_wordcount_device(locust_multi *m, const void *const *dev_shards, const uint64_t *nbytes);
locust_ctx *locust_multi_context(locust_multi *m, int rank);
int locust_multi_stats(locust_multi *m, float *job_ms, uint64_t *record_bytes_exchanged);
const char *locust_multi_last_error(locust_multi *m);

/* ---- for callers that cannot allocate ------------------------------------------
 * No counterpart in the reference (its main() allocates freely, main.cu:393,:402,:451).
 *
 * locust_count_static runs the MAP stage -- tokenise and combine (word, 1) pairs,
 * exactly as locust_map does -- without calling cudaMalloc, cudaFree or any other
 * allocating routine.  It exists for hosts that cannot obtain device memory while
 * they run: a worker sharing its GPU with a process that already holds the heap, a
 * plug-in inside a framework whose pooling allocator owns it, or a latency path
 * that cannot take the device-wide synchronisation an allocation implies.  The
 * combiner's scratch (job counters + hash table, about 1 MiB) is a fixed array
 * inside the library image; every GPU that loads liblocust.so holds it until the
 * process exits.
 *
 * Limits (fixed, nothing grows): the input is at most LOCUST_STATIC_MAX_BYTES, it
 * holds at most LOCUST_STATIC_MAX_KEYS distinct words, and no word is longer than
 * 16 bytes.  A call beyond them fails with LOCUST_E_ARG (nbytes passed in) or
 * LOCUST_E_CAPACITY (found while counting) and says why in errbuf.
 *
 *   dev_data, nbytes   input text in device memory of the current device
 *   dev_records        device buffer for max_records records of 32 bytes:
 *                      {k0, k1, count, 0} as four little-endian uint64; k0/k1 are
 *                      the word's bytes zero padded to 16.  Order is unspecified
 *                      (these are combined partials, not the sorted result).
 *   n_records          out: records written (on "dev_records is too small": needed)
 *   stream             cudaStream_t to run on (NULL = default stream); the call
 *                      returns after the work on it has finished
 *   errbuf, errlen     optional buffer for an error message
 *
 * Runs on the calling thread's current device.  The first call on a device also
 * does a one-time set-up (it resolves the address of the scratch array and sets a
 * kernel attribute); make that first call before any stream capture begins.
 * Every call waits on the host until its work on `stream` has finished, because it
 * reads the job counters back to decide between LOCUST_OK and LOCUST_E_CAPACITY.
 * max_records beyond LOCUST_STATIC_MAX_KEYS is allowed (no more than that many
 * records are ever written).
 * All calls on a device share the one scratch array: the table is cleared at the
 * start of every call, and two calls whose GPU work overlaps (different streams or
 * threads) race on it -- serialise them. */
#define LOCUST_STATIC_MAX_KEYS   16384u
#define LOCUST_STATIC_MAX_BYTES  (1ull << 32)
int locust_count_static(const void *dev_data, uint64_t nbytes, void *dev_records, uint64_t max_records,
                        uint64_t *n_records, void *stream, char *errbuf, size_t errlen);

/* Library version string. */
const char *locust_version(void);

#ifdef __cplusplus
}
#endif
#endif /* LOCUST_H */
