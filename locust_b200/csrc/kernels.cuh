// kernels.cuh -- launch interfaces between the C ABI (locust_api.cu) and the
// hand-written sm_100a kernels of the word-count path.
//
// Every kernel after MAP reads its problem size from device memory (Counters), so
// the three stages run back to back on one stream without a host round trip; the
// host reads Counters once, at the end of REDUCE.
#pragma once
#include "common.cuh"

namespace locust {

// status bits written by kernels
constexpr uint32_t kStatusTableFull = 1u;   // global combiner table ran out of probes
constexpr uint32_t kStatusLongFull  = 2u;   // long-token buffer too small
constexpr uint32_t kStatusEntryFull = 4u;   // candidate-record buffers too small
constexpr uint32_t kStatusBlobFull  = 8u;   // key blob too small
constexpr uint32_t kStatusTextFull  = 16u;  // formatted-text buffer too small
constexpr uint32_t kStatusFarFull   = 32u;  // list of long-token positions of one MAP launch too small
constexpr uint32_t kStatusBadPart   = 64u;  // a received part does not have the layout of a packed part
constexpr uint32_t kStatusTieFull   = 128u;
constexpr uint32_t kStatusNodeLong  = 256u; // PageRank job: a node name of more than 16 bytes // more records with a shared 16-byte prefix than the tie arrays hold

// One entry of the long-key list: `count` occurrences of the key at ptr[0..len), len > 16.
// Entries are (all but) distinct keys: long_keys.cuh finds an existing entry through a hash
// index and adds to its count.  hash = 64-bit hash of the key bytes.
struct LongOcc {
    const uint8_t *ptr;
    uint64_t len;
    unsigned long long count;
    uint64_t hash;
};

// Device-side scalars of one job.
struct Counters {
    unsigned long long n_long;      // entries of the long-key list (can exceed the buffer: then kStatusLongFull)
    unsigned long long n_entries;   // candidate records after extract + append
    unsigned long long n_out;       // unique keys
    unsigned long long blob_bytes;  // sum of key lengths
    unsigned long long total;       // sum of counts = tokens
    unsigned long long text_bytes;  // formatted output size
    unsigned long long n_tied;      // records whose 16-byte prefix is shared and that need ordering by their tails
    unsigned long long n_far;       // long-token occurrences the current MAP launch has listed (can exceed the list)
    unsigned long long far_peak;    // largest n_far of the job
    uint32_t status;
    uint32_t sort_parity;           // 0: sorted records in buffer A, 1: in buffer B
    uint32_t pass_active[16];       // radix pass p moves data
    uint32_t pass_src[16];          // buffer pass p reads (0 = A, 1 = B)
};

// A candidate record of PROCESS/REDUCE, 32 bytes so that one record is one memory sector:
// the radix passes scatter whole records, and an 8-byte-per-array layout would turn every
// scattered record into three or four partly used sectors.
struct __align__(32) Rec {
    uint64_t be0, be1;            // first 16 key bytes, big endian, zero padded: integer order = byte order, prefix first
    unsigned long long count;     // occurrences combined so far
    uint32_t len;                 // key length in bytes
    uint32_t lidx;                // keys longer than 16 bytes: entry of the long-token list that holds the bytes
};

// Double buffer of records for the sort; `longs` is the long-token list Rec::lidx points into.
#if defined(__CUDACC__)
__device__ __forceinline__ Rec load_rec(const Rec *p) {
    const ulonglong4 v = ld256(p);
    Rec r;
    r.be0 = v.x; r.be1 = v.y; r.count = v.z; r.len = (uint32_t)v.w; r.lidx = (uint32_t)(v.w >> 32);
    return r;
}
__device__ __forceinline__ void store_rec(Rec *p, const Rec &r) {
    st256(p, make_ulonglong4(r.be0, r.be1, r.count, (unsigned long long)r.len | ((unsigned long long)r.lidx << 32)));
}
#endif

struct Records {
    Rec *a, *b;
    const LongOcc *longs;
    uint64_t cap;
};

struct MapParams {
    const uint8_t *base;        // 16-byte aligned; valid bytes are base[lo .. hi)
    uint64_t lo, hi;
    Slot *table;
    uint64_t table_mask;        // slots - 1
    LongOcc *long_list;
    uint64_t long_cap;
    uint32_t *long_index;       // open-addressed index over long_list: position + 1, 0 = empty
    uint64_t long_index_mask;   // its slots - 1
    ulonglong2 *far_list;       // {start, end of input} addresses of tokens longer than 16 bytes met by this launch
    uint64_t far_cap;
    Counters *ctr;
    uint32_t flags;             // LOCUST_F_*
    Records rec;                // LOCUST_F_SORT_PATH only: every token becomes a record
};

// ---- second job: PageRank over an edge list (pagerank_kernels.cu) ----
struct PrState {
    unsigned long long n_nodes;     // distinct node names
    unsigned long long n_edges;     // lines with two tokens
    unsigned long long dangling[2]; // rank mass of the nodes without out-edges, by iteration parity
    unsigned long long n_hot;       // edges into heavily linked nodes (marked in the edge list)
};
struct PrArrays {                   // one value per table slot
    unsigned long long *rank, *contrib, *acc;
};
struct PrParams {
    const uint8_t *base;            // 16-byte aligned; valid bytes are base[lo .. hi)
    uint64_t lo, hi;
    Slot *table;                    // Slot::count = out-degree while the job runs, the rank at its end
    uint64_t table_mask;
    uint2 *edges;                   // (slot of src, slot of dst; top bit of dst: a heavily linked node, set after the parse)
    uint64_t edge_cap;
    PrState *st;
    Counters *ctr;
};

struct SortScratch {
    uint32_t *hist;          // 16 * 256 digit histograms
    uint32_t *bases;         // 16 * 256 exclusive bases
    uint64_t *lb_state;      // look-back words: tiles * 256
    uint32_t *tickets;       // pairs of {next tile, finished blocks}
    uint32_t *tie_k, *tie_q; // positions of records with a shared 16-byte prefix: by key / by position
    uint64_t tie_cap;        // their capacity (a power of two, >= 2 x long-key capacity)
};

struct ReduceBufs {
    uint8_t *head;                   // cap flags
    unsigned long long *rank;        // cap
    unsigned long long *csum;        // cap + 1: prefix sum over the counts in sorted order
    uint32_t *src_pos;               // cap: sorted position of each output key
    uint64_t *key_offsets;           // cap + 1
    unsigned long long *counts;      // cap
    unsigned long long *vals;        // cap
    uint8_t *key_blob;
    uint64_t blob_cap;
};

constexpr uint32_t kSortTile = 2048;
constexpr uint32_t kScanTile = 2048;

cudaError_t map_kernel_setup();
int launch_tokenize_combine(const MapParams &p, int num_sms, cudaStream_t st);
// production MAP kernel (flags == 0), map_stream_kernel.cu; launch_tokenize_combine dispatches to it
cudaError_t map_stream_setup();
int launch_tokenize_stream(const MapParams &p, int variant, int num_sms, cudaStream_t st);

// Counts the long tokens a MAP launch has listed (MapParams::far_list) into the long-key list and empties
// the list for the next launch.  Runs right behind every MAP launch.
int launch_long_tokens(const MapParams &p, int num_sms, cudaStream_t st);
// zeroes the long-key index if the counters say the last job used it
int launch_long_index_clear(const Counters *ctr, uint32_t *index, uint64_t slots, int num_sms, cudaStream_t st);

// PageRank job: parse the edge list into the table and the edge array; start ranks, `iterations` rounds of
// scatter + node update, ranks stored in Slot::count at the end
int launch_pagerank_parse(const PrParams &p, int num_sms, cudaStream_t st);
int launch_pagerank_iterate(Slot *table, uint64_t slots, uint2 *edges, uint64_t edge_cap, PrArrays a, PrState *state,
                            uint32_t iterations, unsigned long long damping, int num_sms, cudaStream_t st);

// PROCESS
int launch_extract(const Slot *table, uint64_t slots, Records r, Counters *ctr, int num_sms, cudaStream_t st);
int launch_append_long(const LongOcc *list, uint64_t long_cap, Records r, Counters *ctr, int num_sms, cudaStream_t st);
// epoch: first of 16 consecutive look-back epochs this sort may use
int launch_sort_records(Records r, SortScratch s, Counters *ctr, uint32_t epoch, int num_sms, cudaStream_t st);
int launch_tie_fix(Records r, SortScratch s, Counters *ctr, int num_sms, cudaStream_t st);

// REDUCE.  epoch: first of 3 consecutive epochs.
int launch_reduce(Records r, SortScratch s, ReduceBufs o, Counters *ctr, uint32_t epoch, int num_sms, cudaStream_t st);

// OUTPUT: "print key: ..." lines (main.cu:132 / :286).  Two steps so the host can
// size the text buffer in between.
int launch_format_sizes(ReduceBufs o, Counters *ctr, int cpu_format, unsigned long long *line_off,
                        SortScratch s, uint32_t epoch, int num_sms, cudaStream_t st);
int launch_format_write(ReduceBufs o, Counters *ctr, int cpu_format, const unsigned long long *line_off,
                        char *text, uint64_t text_cap, int num_sms, cudaStream_t st);

// Device-wide exclusive scan of 64-bit values; n is read from *n_ptr.  `out` may
// equal `in`.  The grand total goes to *total (if non-null) and, when
// append_total is set, to out[n].
int launch_exclusive_scan_u64(const unsigned long long *in, unsigned long long *out, const unsigned long long *n_ptr,
                              uint64_t n_max, uint64_t *lb_state, uint32_t *ticket, uint32_t epoch,
                              unsigned long long *total, bool append_total, int num_sms, cudaStream_t st);

// SHUFFLE (multi-GPU)
// Part layout: 32-byte header {n_short, long_bytes, long_off, 0}, n_short records of 32 bytes
// {k0, k1, count, 0} from byte 32, long records {len, count, key bytes padded to 16} from
// byte long_off (right behind the short records when packed for NCCL, at a fixed offset
// when packed straight into a peer's slot).
// part_counts: 2 * nparts scalars (short records, long bytes per part).
int launch_shuffle_count(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                         const Counters *ctr, uint32_t nparts, const uint8_t *partmap, unsigned long long *part_counts,
                         int num_sms, cudaStream_t st);
int launch_shuffle_pack(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                        const Counters *ctr, uint32_t nparts, const uint8_t *partmap, uint8_t *out,
                        const unsigned long long *part_offsets, const unsigned long long *part_counts,
                        unsigned long long *part_cursors, int num_sms, cudaStream_t st);
int launch_shuffle_pack_peers(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                              const Counters *ctr, uint32_t nparts, const uint8_t *partmap, uint8_t *const *dev_dst,
                              uint64_t short_cap, uint64_t long_room, unsigned long long *part_cursors, uint32_t *xflag,
                              int num_sms, cudaStream_t st);
// ordered shuffle: histogram of the keys' first two bytes (65 536 bins), bin -> part map
int launch_partition_hist(const Slot *table, uint64_t slots, const LongOcc *long_list, uint64_t long_cap,
                          const Counters *ctr, unsigned long long *hist, int num_sms, cudaStream_t st);
int launch_partition_map(const unsigned long long *hist, uint32_t nhists, unsigned long long *sum_scratch,
                         uint32_t nparts, uint8_t *partmap, cudaStream_t st);
// peer memory: the same words to several destinations; a barrier between the ranks' streams
int launch_peer_put(uint4 *const *dev_dsts, uint32_t ndst, const uint4 *src, uint64_t n16, int num_sms, cudaStream_t st);
int launch_peer_barrier(uint32_t *const *dev_flags, uint32_t rank, uint32_t world, uint32_t epoch, uint32_t *xflag,
                        cudaStream_t st);
// a sorted result as one part {header, offsets, counts, vals, key bytes}; appending parts to a result
int launch_result_pack(ReduceBufs red, const Counters *ctr, uint8_t *out, uint64_t n_hint, int num_sms, cudaStream_t st);
int launch_result_header(const Counters *ctr, unsigned long long *dev_hdr, cudaStream_t st);
int launch_result_append(ReduceBufs red, const uint8_t *part, uint64_t n_base, uint64_t blob_base, uint64_t total_base,
                         uint64_t n_hint, int num_sms, cudaStream_t st);
int launch_shuffle_merge(const uint8_t *part, uint64_t nbytes, MapParams p, int num_sms, cudaStream_t st);

}  // namespace locust
