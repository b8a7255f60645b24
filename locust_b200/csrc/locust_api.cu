// locust_api.cu -- the C ABI of liblocust.so (include/locust.h): per-device context,
// stage sequencing, buffer ownership, growth and the host copies of the result.
// Plays the role of main() in MapReduce/src/main.cu:358-532 minus argv and stdout.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "../../include/locust.h"

using namespace locust;

namespace {

struct Segment {
    const uint8_t *dev;   // device bytes
    uint64_t nbytes;
    uint8_t *owned;       // non-null when the context made the device copy
    uint64_t owned_cap;   // size of that allocation
};

struct Spare {            // a device copy of an earlier job, kept for the next one
    uint8_t *ptr;
    uint64_t cap;
};

constexpr uint64_t kHostChunk = 32ull << 20;   // host->device copy granularity

}  // namespace

struct locust_ctx {
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    bool own_stream = false;
    uint32_t flags = 0;

    Slot *table = nullptr;
    uint64_t slots = 0;
    LongOcc *long_list = nullptr;
    uint64_t long_cap = 0;
    ulonglong2 *far_list = nullptr;   // positions of long tokens, per MAP launch
    uint64_t far_cap = 0;
    uint32_t *long_index = nullptr;   // hash index over long_list (long_keys.cuh)
    uint64_t long_index_slots = 0;
    bool long_index_dirty = true;     // freshly allocated: contents undefined
    Counters *ctr = nullptr;        // device
    Counters *h_ctr = nullptr;      // pinned host mirror
    Records rec{};
    SortScratch sort{};
    ReduceBufs red{};
    uint64_t lb_words = 0;
    uint32_t epoch = 1;

    std::vector<Segment> segments;
    std::vector<Spare> spares;      // cudaMalloc / cudaFree per job is slow, more so with peer mappings open
    bool cleared = false;           // table/counters zeroed for the current job
    bool shuffled = false;          // partials were packed; the table now holds merged records
    bool replayable = true;         // the job can be redone from `segments` alone (false once records were merged in)
    int stage = 0;                  // 0 mapping, 1 processed, 2 reduced

    // host copies of the result (pinned)
    uint8_t *h_blob = nullptr;  uint64_t h_blob_cap = 0;
    uint64_t *h_offsets = nullptr, *h_counts = nullptr, *h_vals = nullptr;  uint64_t h_rec_cap = 0;
    bool host_valid = false;
    // formatted text
    unsigned long long *line_off = nullptr;  uint64_t line_cap = 0;
    char *d_text = nullptr;  uint64_t d_text_cap = 0;
    char *h_text = nullptr;  uint64_t h_text_cap = 0;
    // shuffle scratch: device [counts 2n | cursors 2n | offsets n], host copies of counts and offsets
    unsigned long long *part_dev = nullptr;  uint32_t part_cap = 0;
    std::vector<unsigned long long> part_host, part_off_host;
    uint32_t part_n = 0;
    unsigned long long *peer_ptrs = nullptr;  // device copies of pointer lists: [0,256) put destinations, [256,512) barrier flags
    std::vector<unsigned long long> peer_put_host, peer_flag_host;
    unsigned long long *hist_sum = nullptr;  // ordered shuffle: the ranks' histograms added up
    uint8_t *partmap = nullptr;              // ordered shuffle: first two key bytes -> part (65 536 entries)
    uint32_t partmap_n = 0;                  // parts the map was built for; 0 = shuffle by hash
    // direct exchange: overflow flag of pack_peers (device + pinned mirror), receive buffers we own
    uint32_t *xflag = nullptr, *h_xflag = nullptr;
    std::vector<void *> exchange_bufs;

    // second job (locust_pagerank): per-slot arrays, edge list, device scalars and their pinned mirror
    PrArrays pr{};  uint64_t pr_slots = 0;
    uint2 *pr_edges = nullptr;  uint64_t pr_edge_cap = 0;
    PrState *pr_state = nullptr, *h_pr_state = nullptr;
    bool pr_done = false;           // the current result is a PageRank result

    cudaEvent_t ev[8]{};            // map, process, reduce, shuffle: begin/end
    bool ev_set[8]{};
    uint64_t launches = 0;
    std::string err;
};

namespace {

int fail(locust_ctx *c, int code, const char *what, cudaError_t e = cudaSuccess) {
    if (c) {
        c->err = what;
        if (e != cudaSuccess) { c->err += ": "; c->err += cudaGetErrorString(e); }
    }
    return code;
}

#define CU(call)                                                              \
    do {                                                                      \
        cudaError_t e_ = (call);                                              \
        if (e_ != cudaSuccess) return fail(c, LOCUST_E_CUDA, #call, e_);      \
    } while (0)

template <class T>
cudaError_t dalloc(T **p, uint64_t n) { return cudaMalloc(reinterpret_cast<void **>(p), n * sizeof(T)); }

void free_record_buffers(locust_ctx *c) {
    cudaFree(c->rec.a); cudaFree(c->rec.b);
    cudaFree(c->sort.lb_state);
    cudaFree(c->red.head); cudaFree(c->red.rank); cudaFree(c->red.csum); cudaFree(c->red.src_pos); cudaFree(c->red.key_offsets);
    cudaFree(c->red.counts); cudaFree(c->red.vals); cudaFree(c->red.key_blob);
    c->rec = Records{};
    c->red = ReduceBufs{};
    c->sort.lb_state = nullptr;
}

// (re)allocates everything whose size follows the table and long-token capacities
int alloc_sized(locust_ctx *c, uint64_t slots, uint64_t long_cap) {
    cudaFree(c->table); c->table = nullptr;
    cudaFree(c->long_list); c->long_list = nullptr;
    cudaFree(c->long_index); c->long_index = nullptr;
    cudaFree(c->sort.tie_k); cudaFree(c->sort.tie_q); c->sort.tie_k = c->sort.tie_q = nullptr;
    free_record_buffers(c);
    c->slots = slots;
    c->long_cap = long_cap;
    const uint64_t cap = slots + long_cap;
    CU(dalloc(&c->table, slots));
    CU(dalloc(&c->long_list, long_cap));
    c->long_index_slots = 4;
    while (c->long_index_slots < 2 * long_cap) c->long_index_slots <<= 1;
    CU(dalloc(&c->long_index, c->long_index_slots));
    c->long_index_dirty = true;
    c->sort.tie_cap = c->long_index_slots;          // >= 2 x long_cap: every long key and one 16-byte key per group
    CU(dalloc(&c->sort.tie_k, c->sort.tie_cap));
    CU(dalloc(&c->sort.tie_q, c->sort.tie_cap));
    CU(dalloc(&c->rec.a, cap)); CU(dalloc(&c->rec.b, cap));
    c->rec.longs = c->long_list;
    c->rec.cap = cap;
    c->lb_words = ((cap + kSortTile - 1) / kSortTile) * 256 + 256;
    CU(dalloc(&c->sort.lb_state, c->lb_words));
    CU(cudaMemsetAsync(c->sort.lb_state, 0, c->lb_words * sizeof(uint64_t), c->stream));
    c->epoch = 1;
    CU(dalloc(&c->red.head, cap)); CU(dalloc(&c->red.rank, cap)); CU(dalloc(&c->red.csum, cap + 1)); CU(dalloc(&c->red.src_pos, cap));
    CU(dalloc(&c->red.key_offsets, cap + 1)); CU(dalloc(&c->red.counts, cap)); CU(dalloc(&c->red.vals, cap));
    c->red.blob_cap = 16 * slots + 64 * long_cap;
    CU(dalloc(&c->red.key_blob, c->red.blob_cap));
    return LOCUST_OK;
}

uint32_t take_epochs(locust_ctx *c, uint32_t k) {
    if (c->epoch + k > 255u) {
        cudaMemsetAsync(c->sort.lb_state, 0, c->lb_words * sizeof(uint64_t), c->stream);
        c->epoch = 1;
    }
    const uint32_t e = c->epoch;
    c->epoch += k;
    return e;
}

// The long-key index is zeroed for the next job only if the last one put anything into it
// (decided on the device: the host does not know n_long without a synchronisation).
cudaError_t clear_long_index(locust_ctx *c) {
    if (c->long_index_dirty) {
        c->long_index_dirty = false;
        return cudaMemsetAsync(c->long_index, 0, c->long_index_slots * sizeof(uint32_t), c->stream);
    }
    c->launches += (uint64_t)launch_long_index_clear(c->ctr, c->long_index, c->long_index_slots, c->num_sms, c->stream);
    return cudaGetLastError();
}

int clear_job(locust_ctx *c) {
    CU(cudaMemsetAsync(c->table, 0, c->slots * sizeof(Slot), c->stream));
    CU(clear_long_index(c));                         // reads the previous job's counters: before they are zeroed
    CU(cudaMemsetAsync(c->ctr, 0, sizeof(Counters), c->stream));
    CU(cudaMemsetAsync(c->xflag, 0, sizeof(uint32_t), c->stream));
    c->cleared = true;
    c->shuffled = false;
    c->stage = 0;
    c->host_valid = false;
    return LOCUST_OK;
}

void drop_segments(locust_ctx *c) {
    for (auto &s : c->segments) {
        if (!s.owned) continue;
        if (c->spares.size() < 4) c->spares.push_back(Spare{s.owned, s.owned_cap});
        else cudaFree(s.owned);
    }
    c->segments.clear();
}

// device memory for a copy of `nbytes` input bytes: a spare of an earlier job if one is large enough
cudaError_t take_device_copy(locust_ctx *c, uint64_t nbytes, uint8_t **dev, uint64_t *cap) {
    for (size_t i = 0; i < c->spares.size(); ++i) {
        if (c->spares[i].cap >= nbytes) {
            *dev = c->spares[i].ptr; *cap = c->spares[i].cap;
            c->spares.erase(c->spares.begin() + (long)i);
            return cudaSuccess;
        }
    }
    if (c->spares.size() >= 4) { cudaFree(c->spares.front().ptr); c->spares.erase(c->spares.begin()); }
    *cap = nbytes;
    return cudaMalloc(reinterpret_cast<void **>(dev), nbytes);
}

MapParams map_params(locust_ctx *c, const uint8_t *dev, uint64_t lo_abs, uint64_t hi_abs) {
    // dev + lo_abs .. dev + hi_abs, expressed against a 16-byte aligned base
    MapParams p{};
    const uintptr_t a = reinterpret_cast<uintptr_t>(dev);
    const uintptr_t base = a & ~uintptr_t(15);
    p.base = reinterpret_cast<const uint8_t *>(base);
    p.lo = (a - base) + lo_abs;
    p.hi = (a - base) + hi_abs;
    p.table = c->table;
    p.table_mask = c->slots - 1;
    p.long_list = c->long_list;
    p.long_cap = c->long_cap;
    p.far_list = c->far_list;
    p.far_cap = c->far_cap;
    p.long_index = c->long_index;
    p.long_index_mask = c->long_index_slots - 1;
    p.ctr = c->ctr;
    p.flags = c->flags;
    p.rec = c->rec;
    return p;
}

int map_device_range(locust_ctx *c, const uint8_t *dev, uint64_t lo, uint64_t hi) {
    if (hi <= lo) return LOCUST_OK;
    const MapParams mp = map_params(c, dev, lo, hi);
    c->launches += (uint64_t)launch_tokenize_combine(mp, c->num_sms, c->stream);
    if (mp.flags == 0u) c->launches += (uint64_t)launch_long_tokens(mp, c->num_sms, c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;
}

bool host_is_delim(uint8_t ch) { return is_delim(ch); }

// HOST input: copy in chunks that end right after a delimiter and tokenise chunk k
// while chunk k+1 is on the wire.
int map_host(locust_ctx *c, const uint8_t *host, uint64_t nbytes, uint8_t *dev) {
    uint64_t done = 0;
    cudaEvent_t ev;
    CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    // the copy stream must not overtake work already queued on the compute stream
    CU(cudaEventRecord(ev, c->stream));
    CU(cudaStreamWaitEvent(c->copy_stream, ev, 0));
    int rc = LOCUST_OK;
    while (done < nbytes) {
        uint64_t end = done + kHostChunk;
        if (end >= nbytes) end = nbytes;
        else {
            // cut after the last delimiter of the chunk so no token straddles two launches
            uint64_t e = end;
            while (e > done && !host_is_delim(host[e - 1])) --e;
            if (e > done) end = e;
            else {                               // one token longer than the chunk: extend to its end
                while (end < nbytes && !host_is_delim(host[end])) ++end;
            }
        }
        cudaError_t e1 = cudaMemcpyAsync(dev + done, host + done, end - done, cudaMemcpyHostToDevice, c->copy_stream);
        if (e1 != cudaSuccess) { rc = fail(c, LOCUST_E_CUDA, "cudaMemcpyAsync(H2D)", e1); break; }
        cudaEventRecord(ev, c->copy_stream);
        cudaStreamWaitEvent(c->stream, ev, 0);
        rc = map_device_range(c, dev, done, end);
        if (rc) break;
        done = end;
    }
    cudaEventDestroy(ev);
    return rc;
}

int run_map_segment(locust_ctx *c, const Segment &s) { return map_device_range(c, s.dev, 0, s.nbytes); }

int run_process(locust_ctx *c) {
    CU(cudaEventRecord(c->ev[2], c->stream)); c->ev_set[2] = true;
    c->launches += (uint64_t)launch_extract(c->table, c->slots, c->rec, c->ctr, c->num_sms, c->stream);
    c->launches += (uint64_t)launch_append_long(c->long_list, c->long_cap, c->rec, c->ctr, c->num_sms, c->stream);
    c->launches += (uint64_t)launch_sort_records(c->rec, c->sort, c->ctr, take_epochs(c, 16), c->num_sms, c->stream);
    c->launches += (uint64_t)launch_tie_fix(c->rec, c->sort, c->ctr, c->num_sms, c->stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev[3], c->stream)); c->ev_set[3] = true;
    c->stage = 1;
    return LOCUST_OK;
}

int run_reduce_async(locust_ctx *c) {
    CU(cudaEventRecord(c->ev[4], c->stream)); c->ev_set[4] = true;
    c->launches += (uint64_t)launch_reduce(c->rec, c->sort, c->red, c->ctr, take_epochs(c, 3), c->num_sms, c->stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev[5], c->stream)); c->ev_set[5] = true;
    CU(cudaMemcpyAsync(c->h_ctr, c->ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(c->h_xflag, c->xflag, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    return LOCUST_OK;
}

// Grows whatever the status bits name and runs MAP again over the recorded inputs.
int grow_and_remap(locust_ctx *c, uint32_t st) {
    if (c->shuffled || !c->replayable) {
        static thread_local char msg[160];
        snprintf(msg, sizeof msg, "table too small for the merged records; raise table_log2 / long_capacity (status bits %u, "
                                  "%llu long keys of %llu)", st, (unsigned long long)c->h_ctr->n_long, (unsigned long long)c->long_cap);
        return fail(c, LOCUST_E_CAPACITY, msg);
    }
    int rc;
    if (st & kStatusFarFull) {                       // more long tokens in one MAP launch than the position list holds
        cudaFree(c->far_list); c->far_list = nullptr;
        uint64_t want = c->h_ctr->far_peak + (c->h_ctr->far_peak >> 3) + 1024;
        if (want < 2 * c->far_cap) want = 2 * c->far_cap;
        c->far_cap = want;
        CU(dalloc(&c->far_list, c->far_cap));
    }
    if (st & (kStatusTableFull | kStatusLongFull | kStatusEntryFull)) {
        uint64_t slots = c->slots, long_cap = c->long_cap;
        if (st & kStatusTableFull) slots *= 4;
        if (st & kStatusLongFull) {                  // n_long stops counting soon after the list is full: at least double
            long_cap = c->h_ctr->n_long + (c->h_ctr->n_long >> 3) + 1024;
            if (long_cap < 2 * c->long_cap) long_cap = 2 * c->long_cap;
        }
        if (st & kStatusEntryFull) {
            // the record buffers hold slots + long_cap entries; n_entries kept counting past the end
            const uint64_t need = c->h_ctr->n_entries + c->h_ctr->n_long + 1024;
            while (slots + long_cap < need && slots < (1ull << 32)) slots *= 2;
        }
        if (slots > (1ull << 32)) return fail(c, LOCUST_E_CAPACITY, "combiner table would exceed 2^32 slots");
        rc = alloc_sized(c, slots, long_cap);
        if (rc) return rc;
    } else if (st & kStatusBlobFull) {   // the key blob was too small
        cudaFree(c->red.key_blob); c->red.key_blob = nullptr;
        c->red.blob_cap = c->h_ctr->blob_bytes + 64;
        CU(dalloc(&c->red.key_blob, c->red.blob_cap));
    }
    rc = clear_job(c);
    if (rc) return rc;
    CU(cudaEventRecord(c->ev[0], c->stream));
    for (const auto &s : c->segments) { rc = run_map_segment(c, s); if (rc) return rc; }
    CU(cudaEventRecord(c->ev[1], c->stream));
    return LOCUST_OK;
}

// Runs REDUCE, reads the counters back and, if a fixed-size buffer was too small,
// grows it and redoes the job from the recorded input segments.
int finish_reduce(locust_ctx *c) {
    for (int attempt = 0; attempt < 8; ++attempt) {
        int rc = run_reduce_async(c);
        if (rc) return rc;
        CU(cudaStreamSynchronize(c->stream));
        if (*c->h_xflag & 4u)
            return fail(c, LOCUST_E_STATE, "locust_peer_barrier: a rank did not arrive within five seconds");
        if (*c->h_xflag & 2u)
            return fail(c, LOCUST_E_CAPACITY, "MAP overflowed the combiner table or the long-token list before "
                                              "locust_shuffle_pack_peers (raise table_log2 / long_capacity)");
        if (*c->h_xflag)
            return fail(c, LOCUST_E_CAPACITY, "locust_shuffle_pack_peers: records did not fit the destination slot "
                                              "(raise slot_bytes); the job's result is incomplete");
        const uint32_t st = c->h_ctr->status;
        if (!(st & (kStatusTableFull | kStatusLongFull | kStatusEntryFull | kStatusBlobFull | kStatusFarFull))) {
            c->stage = 2;
            return LOCUST_OK;
        }
        if (c->shuffled && (st & kStatusBlobFull) && !(st & (kStatusTableFull | kStatusLongFull | kStatusEntryFull | kStatusFarFull))) {
            // merged records stay in the table: only the blob has to grow, then PROCESS + REDUCE again
            cudaFree(c->red.key_blob); c->red.key_blob = nullptr;
            c->red.blob_cap = c->h_ctr->blob_bytes + 64;
            CU(dalloc(&c->red.key_blob, c->red.blob_cap));
            CU(cudaMemsetAsync(&c->ctr->n_entries, 0, sizeof(unsigned long long), c->stream));
            CU(cudaMemsetAsync(&c->ctr->status, 0, sizeof(uint32_t), c->stream));
        } else {
            rc = grow_and_remap(c, st);
            if (rc) return rc;
        }
        rc = run_process(c);
        if (rc) return rc;
    }
    return fail(c, LOCUST_E_CAPACITY, "buffers still too small after repeated growth");
}

int ensure_host_result(locust_ctx *c) {
    if (c->host_valid) return LOCUST_OK;
    const uint64_t n = c->h_ctr->n_out, blob = c->h_ctr->blob_bytes;
    if (n + 1 > c->h_rec_cap) {
        cudaFreeHost(c->h_offsets); cudaFreeHost(c->h_counts); cudaFreeHost(c->h_vals);
        c->h_offsets = c->h_counts = c->h_vals = nullptr;
        c->h_rec_cap = n + 1 + (n >> 2);
        CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_offsets), c->h_rec_cap * 8));
        CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_counts), c->h_rec_cap * 8));
        CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_vals), c->h_rec_cap * 8));
    }
    if (blob + 1 > c->h_blob_cap) {
        cudaFreeHost(c->h_blob); c->h_blob = nullptr;
        c->h_blob_cap = blob + 1 + (blob >> 2);
        CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_blob), c->h_blob_cap));
    }
    CU(cudaMemcpyAsync(c->h_offsets, c->red.key_offsets, (n + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    if (n) {
        CU(cudaMemcpyAsync(c->h_counts, c->red.counts, n * 8, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(c->h_vals, c->red.vals, n * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    if (blob) CU(cudaMemcpyAsync(c->h_blob, c->red.key_blob, blob, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->host_valid = true;
    return LOCUST_OK;
}

}  // namespace

extern "C" {

const char *locust_version(void) { return "locust-b200 0.1 (sm_100a)"; }

int locust_create(const locust_config *cfg, locust_ctx **out) {
    if (!out) return LOCUST_E_ARG;
    *out = nullptr;
    locust_config def{};
    if (!cfg) cfg = &def;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return LOCUST_E_CUDA;     // no CPU fallback, by design
    if (cfg->device < 0 || cfg->device >= ndev) return LOCUST_E_ARG;
    if (cfg->table_log2 > 32) return LOCUST_E_ARG;
    locust_ctx *c = new (std::nothrow) locust_ctx();
    if (!c) return LOCUST_E_NOMEM;
    *out = c;                      // handed out even on failure so last_error can be read; destroy frees it
    c->device = cfg->device;
    c->flags = cfg->flags;
    CU(cudaSetDevice(c->device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, c->device));
    if (prop.major < 10) return fail(c, LOCUST_E_CUDA, "device is not sm_100 class (this library is built for sm_100a only)");
    c->num_sms = prop.multiProcessorCount;
    if (cfg->stream) c->stream = static_cast<cudaStream_t>(cfg->stream);
    else { CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }
    CU(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (auto &ev : c->ev) CU(cudaEventCreate(&ev));
    CU(map_kernel_setup());
    CU(dalloc(&c->ctr, 1));
    CU(dalloc(&c->xflag, 1));
    CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_xflag), sizeof(uint32_t)));
    *c->h_xflag = 0;
    CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_ctr), sizeof(Counters)));
    memset(c->h_ctr, 0, sizeof(Counters));
    CU(dalloc(&c->sort.hist, 16 * 256 + 16));
    CU(dalloc(&c->sort.bases, 16 * 256));
    CU(dalloc(&c->sort.tickets, 64));
    CU(cudaMemsetAsync(c->sort.tickets, 0, 64 * sizeof(uint32_t), c->stream));
    c->far_cap = 1ull << 21;
    CU(dalloc(&c->far_list, c->far_cap));
    const uint32_t lg = cfg->table_log2 ? cfg->table_log2 : 20;
    const uint64_t long_cap = cfg->long_capacity ? cfg->long_capacity : (1ull << 20);
    int rc = alloc_sized(c, 1ull << lg, long_cap);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    return LOCUST_OK;
}

int locust_destroy(locust_ctx *c) {
    if (!c) return LOCUST_E_ARG;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    drop_segments(c);
    for (auto &sp : c->spares) cudaFree(sp.ptr);
    cudaFree(c->table); cudaFree(c->long_list); cudaFree(c->long_index); cudaFree(c->far_list); cudaFree(c->ctr);
    cudaFree(c->sort.tie_k); cudaFree(c->sort.tie_q);
    free_record_buffers(c);
    cudaFree(c->sort.hist); cudaFree(c->sort.bases); cudaFree(c->sort.tickets);
    cudaFree(c->line_off); cudaFree(c->d_text); cudaFree(c->part_dev); cudaFree(c->xflag); cudaFree(c->partmap); cudaFree(c->peer_ptrs); cudaFree(c->hist_sum);
    for (void *b : c->exchange_bufs) cudaFree(b);
    cudaFree(c->pr.rank); cudaFree(c->pr.contrib); cudaFree(c->pr.acc); cudaFree(c->pr_edges); cudaFree(c->pr_state);
    cudaFreeHost(c->h_pr_state);
    cudaFreeHost(c->h_xflag);
    cudaFreeHost(c->h_ctr); cudaFreeHost(c->h_blob); cudaFreeHost(c->h_offsets);
    cudaFreeHost(c->h_counts); cudaFreeHost(c->h_vals); cudaFreeHost(c->h_text);
    for (auto &ev : c->ev) if (ev) cudaEventDestroy(ev);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return LOCUST_OK;
}

int locust_reset(locust_ctx *c) {
    if (!c) return LOCUST_E_ARG;
    CU(cudaSetDevice(c->device));
    // device copies of host inputs may still be read by queued kernels
    if (!c->segments.empty()) CU(cudaStreamSynchronize(c->stream));
    drop_segments(c);
    c->cleared = false;
    c->shuffled = false;
    c->replayable = true;
    c->stage = 0;
    c->host_valid = false;
    c->partmap_n = 0;
    c->pr_done = false;
    for (auto &b : c->ev_set) b = false;
    c->launches = 0;
    c->err.clear();
    return LOCUST_OK;
}

int locust_map(locust_ctx *c, const void *data, uint64_t nbytes, int where) {
    if (!c) return LOCUST_E_ARG;
    if (!data && nbytes) return fail(c, LOCUST_E_ARG, "locust_map: data is NULL");
    if (where != LOCUST_HOST && where != LOCUST_DEVICE) return fail(c, LOCUST_E_ARG, "locust_map: bad location");
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_map after locust_process; call locust_reset first");
    CU(cudaSetDevice(c->device));
    if (!c->cleared) {
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        int rc = clear_job(c);
        if (rc) return rc;
    }
    int rc = LOCUST_OK;
    if (nbytes) {
        if (where == LOCUST_DEVICE) {
            Segment s{static_cast<const uint8_t *>(data), nbytes, nullptr, 0};
            c->segments.push_back(s);
            rc = run_map_segment(c, s);
        } else {
            uint8_t *dev = nullptr;
            uint64_t cap = 0;
            CU(take_device_copy(c, nbytes + 64, &dev, &cap));
            c->segments.push_back(Segment{dev, nbytes, dev, cap});
            rc = map_host(c, static_cast<const uint8_t *>(data), nbytes, dev);
        }
    }
    if (rc) return rc;
    CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    return LOCUST_OK;
}

int locust_wait_input(locust_ctx *c) {
    if (!c) return LOCUST_E_ARG;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->copy_stream));
    return LOCUST_OK;
}

int locust_process(locust_ctx *c) {
    if (!c) return LOCUST_E_ARG;
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_process called twice");
    CU(cudaSetDevice(c->device));
    if (!c->cleared) {            // no input at all: still a valid, empty job
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        int rc = clear_job(c);
        if (rc) return rc;
        CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    }
    return run_process(c);
}

int locust_reduce(locust_ctx *c) {
    if (!c) return LOCUST_E_ARG;
    if (c->stage != 1) return fail(c, LOCUST_E_STATE, "locust_reduce needs locust_process first");
    CU(cudaSetDevice(c->device));
    return finish_reduce(c);
}

int locust_wordcount(locust_ctx *c, const void *data, uint64_t nbytes, int where) {
    int rc = locust_reset(c);
    if (rc) return rc;
    rc = locust_map(c, data, nbytes, where);
    if (rc) return rc;
    rc = locust_process(c);
    if (rc) return rc;
    return locust_reduce(c);
}

// ---- second job: PageRank over an edge list (pagerank_kernels.cu; defined by oracle/pagerank_oracle.c) ----
int locust_pagerank(locust_ctx *c, const void *data, uint64_t nbytes, int where, uint32_t iterations, uint64_t damping_q32) {
    int rc = locust_reset(c);
    if (rc) return rc;
    if (!data && nbytes) return fail(c, LOCUST_E_ARG, "locust_pagerank: data is NULL");
    if (where != LOCUST_HOST && where != LOCUST_DEVICE) return fail(c, LOCUST_E_ARG, "locust_pagerank: bad location");
    if (damping_q32 > (1ull << 32)) return fail(c, LOCUST_E_ARG, "locust_pagerank: damping is Q32 and at most 2^32 (1.0)");
    if (c->flags) return fail(c, LOCUST_E_ARG, "locust_pagerank: contexts with test flags run word count only");
    if (c->slots > (1ull << 31)) return fail(c, LOCUST_E_ARG, "locust_pagerank: needs a table of at most 2^31 slots (edges carry slot numbers in 31 bits)");
    CU(cudaSetDevice(c->device));
    if (!c->pr_state) {
        CU(dalloc(&c->pr_state, 1));
        CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_pr_state), sizeof(PrState)));
    }
    const uint8_t *dev = static_cast<const uint8_t *>(data);
    if (nbytes && where == LOCUST_HOST) {
        uint8_t *copy = nullptr;
        uint64_t cap = 0;
        CU(take_device_copy(c, nbytes + 64, &copy, &cap));
        c->segments.push_back(Segment{copy, nbytes, copy, cap});
        CU(cudaMemcpyAsync(copy, data, nbytes, cudaMemcpyHostToDevice, c->stream));
        dev = copy;
    } else if (nbytes) {
        c->segments.push_back(Segment{dev, nbytes, nullptr, 0});
    }
    c->replayable = false;                      // growth below is this job's own; the word-count replay must never run
    const uint64_t want_edges = nbytes / 4 + 1;  // an edge line has at least four bytes
    if (want_edges > c->pr_edge_cap) {
        cudaFree(c->pr_edges); c->pr_edges = nullptr; c->pr_edge_cap = 0;
        CU(dalloc(&c->pr_edges, want_edges));
        c->pr_edge_cap = want_edges;
    }
    for (int attempt = 0; ; ++attempt) {
        if (c->pr_slots != c->slots) {
            cudaFree(c->pr.rank); cudaFree(c->pr.contrib); cudaFree(c->pr.acc);
            c->pr = PrArrays{}; c->pr_slots = 0;
            CU(dalloc(&c->pr.rank, c->slots)); CU(dalloc(&c->pr.contrib, c->slots)); CU(dalloc(&c->pr.acc, c->slots));
            c->pr_slots = c->slots;
        }
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        rc = clear_job(c);
        if (rc) return rc;
        CU(cudaMemsetAsync(c->pr_state, 0, sizeof(PrState), c->stream));
        PrParams p{};
        const uintptr_t a = reinterpret_cast<uintptr_t>(dev), base = a & ~uintptr_t(15);
        p.base = reinterpret_cast<const uint8_t *>(base);
        p.lo = a - base; p.hi = (a - base) + nbytes;
        p.table = c->table; p.table_mask = c->slots - 1;
        p.edges = c->pr_edges; p.edge_cap = c->pr_edge_cap;
        p.st = c->pr_state; p.ctr = c->ctr;
        c->launches += (uint64_t)launch_pagerank_parse(p, c->num_sms, c->stream);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(c->h_ctr, c->ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(c->h_pr_state, c->pr_state, sizeof(PrState), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        const uint32_t st = c->h_ctr->status;
        if (st & kStatusNodeLong) return fail(c, LOCUST_E_ARG, "locust_pagerank: a node name has more than 16 bytes");
        if (st & kStatusEntryFull) return fail(c, LOCUST_E_STATE, "locust_pagerank: edge list overflow");
        // redo with a larger table when it filled up, or when it is more than half full (long probe chains)
        if (!(st & kStatusTableFull) && c->h_pr_state->n_nodes * 2 <= c->slots) break;
        if (attempt >= 8 || c->slots >= (1ull << 31)) return fail(c, LOCUST_E_CAPACITY, "locust_pagerank: more nodes than a table of 2^31 slots holds");
        uint64_t slots = c->slots * 4;
        while (slots < 2 * c->h_pr_state->n_nodes && slots < (1ull << 31)) slots *= 2;
        if (slots > (1ull << 31)) slots = 1ull << 31;
        rc = alloc_sized(c, slots, c->long_cap);
        if (rc) return rc;
    }
    CU(cudaMemsetAsync(c->pr.acc, 0, c->slots * sizeof(unsigned long long), c->stream));      // counts in-degrees first
    c->launches += (uint64_t)launch_pagerank_iterate(c->table, c->slots, c->pr_edges, c->h_pr_state->n_edges, c->pr, c->pr_state,
                                                     iterations, damping_q32, c->num_sms, c->stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    rc = run_process(c);
    if (rc) return rc;
    rc = finish_reduce(c);
    if (rc) return rc;
    c->pr_done = true;
    return LOCUST_OK;
}

int locust_pagerank_stats(locust_ctx *c, uint64_t *nodes, uint64_t *edges) {
    if (!c) return LOCUST_E_ARG;
    if (!c->pr_done) return fail(c, LOCUST_E_STATE, "locust_pagerank_stats: the current result is not a PageRank result");
    if (nodes) *nodes = c->h_pr_state->n_nodes;
    if (edges) *edges = c->h_pr_state->n_edges;
    return LOCUST_OK;
}

int locust_result_device(locust_ctx *c, const uint8_t **key_blob, const uint64_t **key_offsets,
                         const uint64_t **counts, const uint64_t **vals, uint64_t *n) {
    if (!c) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "no result yet: run locust_reduce");
    if (key_blob) *key_blob = c->red.key_blob;
    if (key_offsets) *key_offsets = c->red.key_offsets;
    if (counts) *counts = reinterpret_cast<const uint64_t *>(c->red.counts);
    if (vals) *vals = reinterpret_cast<const uint64_t *>(c->red.vals);
    if (n) *n = c->h_ctr->n_out;
    return LOCUST_OK;
}

int locust_result_host(locust_ctx *c, const uint8_t **key_blob, const uint64_t **key_offsets,
                       const uint64_t **counts, const uint64_t **vals, uint64_t *n) {
    if (!c) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "no result yet: run locust_reduce");
    CU(cudaSetDevice(c->device));
    int rc = ensure_host_result(c);
    if (rc) return rc;
    if (key_blob) *key_blob = c->h_blob;
    if (key_offsets) *key_offsets = c->h_offsets;
    if (counts) *counts = c->h_counts;
    if (vals) *vals = c->h_vals;
    if (n) *n = c->h_ctr->n_out;
    return LOCUST_OK;
}

int locust_format_result(locust_ctx *c, int cpu_format, const char **text, uint64_t *nbytes) {
    if (!c || !text || !nbytes || cpu_format < 0 || cpu_format > 2) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "no result yet: run locust_reduce");
    CU(cudaSetDevice(c->device));
    const uint64_t n = c->h_ctr->n_out;
    if (n + 1 > c->line_cap) {
        cudaFree(c->line_off); c->line_off = nullptr;
        c->line_cap = n + 1;
        CU(dalloc(&c->line_off, c->line_cap));
    }
    c->launches += (uint64_t)launch_format_sizes(c->red, c->ctr, cpu_format, c->line_off, c->sort, take_epochs(c, 1),
                                                 c->num_sms, c->stream);
    CU(cudaMemcpyAsync(c->h_ctr, c->ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    const uint64_t tb = c->h_ctr->text_bytes;
    if (tb + 1 > c->d_text_cap) {
        cudaFree(c->d_text); c->d_text = nullptr;
        c->d_text_cap = tb + 1;
        CU(dalloc(&c->d_text, c->d_text_cap));
    }
    if (tb + 1 > c->h_text_cap) {
        cudaFreeHost(c->h_text); c->h_text = nullptr;
        c->h_text_cap = tb + 1;
        CU(cudaMallocHost(reinterpret_cast<void **>(&c->h_text), c->h_text_cap));
    }
    c->launches += (uint64_t)launch_format_write(c->red, c->ctr, cpu_format, c->line_off, c->d_text, c->d_text_cap,
                                                 c->num_sms, c->stream);
    CU(cudaGetLastError());
    if (tb) CU(cudaMemcpyAsync(c->h_text, c->d_text, tb, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *text = c->h_text;
    *nbytes = tb;
    return LOCUST_OK;
}

float locust_stage_ms(locust_ctx *c, int stage) {
    if (!c || stage < 0 || stage > 3) return -1.0f;
    if (!c->ev_set[2 * stage] || !c->ev_set[2 * stage + 1]) return -1.0f;
    cudaSetDevice(c->device);
    if (cudaEventSynchronize(c->ev[2 * stage + 1]) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, c->ev[2 * stage], c->ev[2 * stage + 1]) != cudaSuccess) return -1.0f;
    return ms;
}

int locust_stats(locust_ctx *c, uint64_t *tokens, uint64_t *unique_keys, uint64_t *kernel_launches) {
    if (!c) return LOCUST_E_ARG;
    if (tokens) *tokens = c->stage == 2 ? c->h_ctr->total : 0;
    if (unique_keys) *unique_keys = c->stage == 2 ? c->h_ctr->n_out : 0;
    if (kernel_launches) *kernel_launches = c->launches;
    return LOCUST_OK;
}

const char *locust_last_error(locust_ctx *c) { return c ? c->err.c_str() : "null context"; }

void *locust_stream(locust_ctx *c) { return c ? static_cast<void *>(c->stream) : nullptr; }

// ---- intermediate text of multi mode (main.cu:66-103) -----------------------------------------
int locust_map_intermediate(locust_ctx *c, const char *text, uint64_t nbytes) {
    if (!c || (!text && nbytes)) return LOCUST_E_ARG;
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_map_intermediate after locust_process; call locust_reset first");
    CU(cudaSetDevice(c->device));
    if (!c->cleared) {
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        int rc = clear_job(c);
        if (rc) return rc;
    }
    // host: lines -> one packed part (header, 32-byte short records, long records)
    std::vector<uint64_t> shorts;
    std::vector<uint8_t> longs;
    uint64_t i = 0;
    while (i < nbytes) {
        const char *nl = static_cast<const char *>(memchr(text + i, '\n', nbytes - i));
        const uint64_t e = nl ? (uint64_t)(nl - text) : nbytes;
        const char *tab = static_cast<const char *>(memchr(text + i, '\t', e - i));
        if (tab) {
            uint64_t klen = (uint64_t)(tab - (text + i));
            if (klen && text[i + klen - 1] == ' ') --klen;              // the blank of "%s \t%d"
            const long long v = strtoll(std::string(tab + 1, text + e).c_str(), nullptr, 10);
            if (klen && v > 0) {
                if (klen <= 16) {
                    uint64_t k[2] = {0, 0};
                    memcpy(k, text + i, klen);
                    shorts.push_back(k[0]); shorts.push_back(k[1]); shorts.push_back((uint64_t)v); shorts.push_back(0);
                } else {
                    const uint64_t padded = (klen + 15) & ~15ull, at = longs.size();
                    longs.resize(at + 16 + padded, 0);
                    memcpy(&longs[at], &klen, 8);
                    const uint64_t cnt = (uint64_t)v;
                    memcpy(&longs[at + 8], &cnt, 8);
                    memcpy(&longs[at + 16], text + i, klen);
                }
            }
        }
        i = e + 1;
    }
    const uint64_t n_short = shorts.size() / 4, total = (32 + 32 * n_short + longs.size() + 31) & ~31ull;
    std::vector<uint8_t> part(total, 0);
    const uint64_t hdr[4] = {n_short, (uint64_t)longs.size(), 32 + 32 * n_short, 0};
    memcpy(part.data(), hdr, 32);
    if (n_short) memcpy(part.data() + 32, shorts.data(), 32 * n_short);
    if (!longs.empty()) memcpy(part.data() + 32 + 32 * n_short, longs.data(), longs.size());
    uint8_t *dev = nullptr;
    uint64_t cap = 0;
    CU(take_device_copy(c, total, &dev, &cap));
    c->segments.push_back(Segment{dev, 0, dev, cap});      // owned: long keys point into it until reset
    CU(cudaMemcpyAsync(dev, part.data(), total, cudaMemcpyHostToDevice, c->stream));
    MapParams p = map_params(c, dev, 0, 0);
    c->launches += (uint64_t)launch_shuffle_merge(dev, total, p, c->num_sms, c->stream);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));                    // `part` is about to go out of scope
    c->replayable = false;
    CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    return LOCUST_OK;
}

// ---- multi-GPU shuffle ------------------------------------------------------------------
static const uint8_t *part_map_for(const locust_ctx *c) { return c->partmap_n ? c->partmap : nullptr; }
static uint64_t part_bytes(uint64_t n_short, uint64_t long_bytes) {
    return (32ull + 32ull * n_short + long_bytes + 31ull) & ~31ull;
}

int locust_shuffle_sizes(locust_ctx *c, uint32_t nparts, uint64_t *bytes_per_part) {
    if (!c || !bytes_per_part || nparts == 0) return LOCUST_E_ARG;
    if (nparts > 256) return fail(c, LOCUST_E_ARG, "locust_shuffle_sizes: at most 256 parts");
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_shuffle_sizes after locust_process");
    if (c->partmap_n && c->partmap_n != nparts)
        return fail(c, LOCUST_E_STATE, "locust_shuffle_sizes: the key ranges were built for a different number of parts");
    CU(cudaSetDevice(c->device));
    if (!c->cleared) {            // nothing mapped on this rank: an empty job still takes part in the exchange
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        int rc0 = clear_job(c);
        if (rc0) return rc0;
        CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    }
    if (!c->ev_set[6]) { CU(cudaEventRecord(c->ev[6], c->stream)); c->ev_set[6] = true; }
    if (nparts > c->part_cap) {
        cudaFree(c->part_dev); c->part_dev = nullptr;
        c->part_cap = nparts;
        CU(dalloc(&c->part_dev, 5ull * nparts));
    }
    c->part_n = nparts;
    c->part_host.assign(2ull * nparts, 0ull);
    // count optimistically and read the job counters back in the same round trip; if MAP
    // overflowed a buffer, grow, redo MAP and count again
    for (int attempt = 0;; ++attempt) {
        CU(cudaMemsetAsync(c->part_dev, 0, 4ull * nparts * sizeof(unsigned long long), c->stream));
        c->launches += (uint64_t)launch_shuffle_count(c->table, c->slots, c->long_list, c->long_cap, c->ctr, nparts,
                                                      part_map_for(c), c->part_dev, c->num_sms, c->stream);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(c->part_host.data(), c->part_dev, 2ull * nparts * sizeof(unsigned long long),
                           cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(c->h_ctr, c->ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        const uint32_t st = c->h_ctr->status;
        if (!(st & (kStatusTableFull | kStatusLongFull | kStatusEntryFull | kStatusFarFull))) break;
        if (attempt >= 8) return fail(c, LOCUST_E_CAPACITY, "buffers still too small after repeated growth");
        int rc = grow_and_remap(c, st);
        if (rc) return rc;
    }
    for (uint32_t i = 0; i < nparts; ++i) bytes_per_part[i] = part_bytes(c->part_host[i], c->part_host[nparts + i]);
    return LOCUST_OK;
}

int locust_shuffle_counts(locust_ctx *c, uint32_t nparts, uint64_t *short_records, uint64_t *long_bytes) {
    if (!c || nparts == 0 || !short_records || !long_bytes) return LOCUST_E_ARG;
    if (c->part_n != nparts || c->part_host.size() < 2ull * nparts)
        return fail(c, LOCUST_E_STATE, "locust_shuffle_counts needs locust_shuffle_sizes with the same nparts first");
    for (uint32_t i = 0; i < nparts; ++i) { short_records[i] = c->part_host[i]; long_bytes[i] = c->part_host[nparts + i]; }
    return LOCUST_OK;
}

int locust_shuffle_pack(locust_ctx *c, uint32_t nparts, void *dev_out, const uint64_t *part_offsets) {
    if (!c || !dev_out || !part_offsets || nparts == 0) return LOCUST_E_ARG;
    if (c->stage != 0 || c->part_n != nparts) return fail(c, LOCUST_E_STATE, "locust_shuffle_pack needs locust_shuffle_sizes with the same nparts first");
    CU(cudaSetDevice(c->device));
    c->part_off_host.assign(part_offsets, part_offsets + nparts);
    for (uint32_t i = 0; i < nparts; ++i)
        if (part_offsets[i] % 32) return fail(c, LOCUST_E_ARG, "part offsets must be multiples of 32");
    CU(cudaMemcpyAsync(c->part_dev + 4ull * nparts, c->part_off_host.data(), nparts * sizeof(unsigned long long),
                       cudaMemcpyHostToDevice, c->stream));
    c->launches += (uint64_t)launch_shuffle_pack(c->table, c->slots, c->long_list, c->long_cap, c->ctr, nparts,
                                                 part_map_for(c), static_cast<uint8_t *>(dev_out), c->part_dev + 4ull * nparts, c->part_dev,
                                                 c->part_dev + 2ull * nparts, c->num_sms, c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;      // (the offsets came from pageable memory: that copy was staged before it returned)
}

int locust_exchange_alloc(locust_ctx *c, uint64_t nbytes, void **dev_buf) {
    if (!c || !dev_buf || nbytes == 0) return LOCUST_E_ARG;
    CU(cudaSetDevice(c->device));
    void *b = nullptr;
    CU(cudaMalloc(&b, nbytes));
    CU(cudaMemsetAsync(b, 0, nbytes, c->stream));          // empty headers: a slot nobody wrote merges as nothing
    CU(cudaStreamSynchronize(c->stream));
    c->exchange_bufs.push_back(b);
    *dev_buf = b;
    return LOCUST_OK;
}

int locust_exchange_free(locust_ctx *c, void *dev_buf) {
    if (!c || !dev_buf) return LOCUST_E_ARG;
    CU(cudaSetDevice(c->device));
    for (size_t i = 0; i < c->exchange_bufs.size(); ++i) {
        if (c->exchange_bufs[i] != dev_buf) continue;
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaFree(dev_buf));
        c->exchange_bufs.erase(c->exchange_bufs.begin() + (long)i);
        return LOCUST_OK;
    }
    return fail(c, LOCUST_E_ARG, "locust_exchange_free: not a buffer of locust_exchange_alloc on this context");
}

int locust_ipc_export(locust_ctx *c, const void *dev_buf, uint8_t handle[64]) {
    if (!c || !dev_buf || !handle) return LOCUST_E_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CU(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, const_cast<void *>(dev_buf)));
    memcpy(handle, &h, 64);
    return LOCUST_OK;
}

int locust_ipc_open(locust_ctx *c, const uint8_t handle[64], void **dev_buf) {
    if (!c || !handle || !dev_buf) return LOCUST_E_ARG;
    CU(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    CU(cudaIpcOpenMemHandle(dev_buf, h, cudaIpcMemLazyEnablePeerAccess));
    return LOCUST_OK;
}

int locust_ipc_close(locust_ctx *c, void *dev_buf) {
    if (!c || !dev_buf) return LOCUST_E_ARG;
    CU(cudaSetDevice(c->device));
    CU(cudaIpcCloseMemHandle(dev_buf));
    return LOCUST_OK;
}

int locust_shuffle_pack_peers(locust_ctx *c, uint32_t nparts, void *const *dst_slots, uint64_t slot_bytes,
                              uint64_t short_cap) {
    if (!c || !dst_slots || nparts == 0) return LOCUST_E_ARG;
    if (nparts > 256) return fail(c, LOCUST_E_ARG, "locust_shuffle_pack_peers: at most 256 parts");
    if (slot_bytes % 32 || slot_bytes < 32 + 32 * short_cap)
        return fail(c, LOCUST_E_ARG, "locust_shuffle_pack_peers: slot_bytes must be a multiple of 32 and hold short_cap records");
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_shuffle_pack_peers after locust_process");
    if (c->partmap_n && c->partmap_n != nparts)
        return fail(c, LOCUST_E_STATE, "locust_shuffle_pack_peers: the key ranges were built for a different number of parts");
    CU(cudaSetDevice(c->device));
    if (!c->cleared) {
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        int rc0 = clear_job(c);
        if (rc0) return rc0;
        CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    }
    if (!c->ev_set[6]) { CU(cudaEventRecord(c->ev[6], c->stream)); c->ev_set[6] = true; }
    if (nparts > c->part_cap) {
        cudaFree(c->part_dev); c->part_dev = nullptr;
        c->part_cap = nparts;
        CU(dalloc(&c->part_dev, 5ull * nparts));
    }
    // device copy of the destination pointers lives behind the cursors (same layout as the offsets of the NCCL path)
    c->part_off_host.resize(nparts);
    for (uint32_t i = 0; i < nparts; ++i) {
        if (!dst_slots[i]) return fail(c, LOCUST_E_ARG, "locust_shuffle_pack_peers: NULL destination");
        c->part_off_host[i] = reinterpret_cast<unsigned long long>(dst_slots[i]);
    }
    CU(cudaMemsetAsync(c->part_dev, 0, 4ull * nparts * sizeof(unsigned long long), c->stream));
    CU(cudaMemcpyAsync(c->part_dev + 4ull * nparts, c->part_off_host.data(), nparts * sizeof(unsigned long long),
                       cudaMemcpyHostToDevice, c->stream));
    c->launches += (uint64_t)launch_shuffle_pack_peers(
        c->table, c->slots, c->long_list, c->long_cap, c->ctr, nparts, part_map_for(c),
        reinterpret_cast<uint8_t *const *>(c->part_dev + 4ull * nparts), short_cap, slot_bytes - 32 - 32 * short_cap,
        c->part_dev + 2ull * nparts, c->xflag, c->num_sms, c->stream);
    CU(cudaGetLastError());
    c->replayable = false;      // a MAP overflow can no longer be repaired by redoing MAP: the parts are gone
    return LOCUST_OK;
}

// ---- ordered shuffle: parts are key ranges, so each rank's sorted result is one slice of the whole ----
int locust_partition_histogram(locust_ctx *c, void *dev_hist) {
    if (!c || !dev_hist) return LOCUST_E_ARG;
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_partition_histogram after locust_process");
    CU(cudaSetDevice(c->device));
    if (!c->cleared) {            // nothing mapped on this rank: it still takes part with an empty table
        CU(cudaEventRecord(c->ev[0], c->stream)); c->ev_set[0] = true;
        int rc0 = clear_job(c);
        if (rc0) return rc0;
        CU(cudaEventRecord(c->ev[1], c->stream)); c->ev_set[1] = true;
    }
    if (!c->ev_set[6]) { CU(cudaEventRecord(c->ev[6], c->stream)); c->ev_set[6] = true; }
    c->launches += (uint64_t)launch_partition_hist(c->table, c->slots, c->long_list, c->long_cap, c->ctr,
                                                   static_cast<unsigned long long *>(dev_hist), c->num_sms, c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;
}

int locust_partition_by_range(locust_ctx *c, const void *dev_hist, uint32_t nparts) {
    return locust_partition_by_range_sum(c, dev_hist, 1, nparts);
}

int locust_partition_by_range_sum(locust_ctx *c, const void *dev_hist, uint32_t nhists, uint32_t nparts) {
    if (!c) return LOCUST_E_ARG;
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_partition_by_range after locust_process");
    if (nparts == 0) { c->partmap_n = 0; return LOCUST_OK; }       // back to the hash shuffle
    if (!dev_hist) return LOCUST_E_ARG;
    if (nparts > 256 || nhists == 0 || nhists > 256)
        return fail(c, LOCUST_E_ARG, "locust_partition_by_range: at most 256 parts, 1 to 256 histograms");
    CU(cudaSetDevice(c->device));
    if (!c->partmap) CU(dalloc(&c->partmap, (uint64_t)LOCUST_PARTITION_BINS));
    if (nhists > 1 && !c->hist_sum) CU(dalloc(&c->hist_sum, (uint64_t)LOCUST_PARTITION_BINS));
    c->launches += (uint64_t)launch_partition_map(static_cast<const unsigned long long *>(dev_hist), nhists, c->hist_sum, nparts, c->partmap, c->stream);
    CU(cudaGetLastError());
    c->partmap_n = nparts;
    return LOCUST_OK;
}

// device copy of a short pointer list, refreshed only when the list changes
static int stage_pointers(locust_ctx *c, std::vector<unsigned long long> &cache, uint32_t at, uint32_t n, void *const *ptrs) {
    if (!c->peer_ptrs) CU(dalloc(&c->peer_ptrs, 512));
    bool same = cache.size() == n;
    for (uint32_t i = 0; same && i < n; ++i) same = cache[i] == reinterpret_cast<unsigned long long>(ptrs[i]);
    if (same) return LOCUST_OK;
    CU(cudaStreamSynchronize(c->stream));            // an earlier kernel may still be reading the old list
    cache.resize(n);
    for (uint32_t i = 0; i < n; ++i) cache[i] = reinterpret_cast<unsigned long long>(ptrs[i]);
    CU(cudaMemcpyAsync(c->peer_ptrs + at, cache.data(), n * sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
    return LOCUST_OK;
}

int locust_peer_put(locust_ctx *c, uint32_t ndst, void *const *dsts, const void *dev_src, uint64_t nbytes) {
    if (!c || !dsts || !dev_src || ndst == 0 || ndst > 256) return LOCUST_E_ARG;
    if (nbytes % 16 || reinterpret_cast<uintptr_t>(dev_src) % 16) return fail(c, LOCUST_E_ARG, "locust_peer_put: 16-byte units");
    for (uint32_t i = 0; i < ndst; ++i)
        if (!dsts[i] || reinterpret_cast<uintptr_t>(dsts[i]) % 16) return fail(c, LOCUST_E_ARG, "locust_peer_put: destinations must be 16-byte aligned");
    CU(cudaSetDevice(c->device));
    int rc = stage_pointers(c, c->peer_put_host, 0, ndst, dsts);
    if (rc) return rc;
    c->launches += (uint64_t)launch_peer_put(reinterpret_cast<uint4 *const *>(c->peer_ptrs), ndst, static_cast<const uint4 *>(dev_src),
                                             nbytes / 16, c->num_sms, c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;
}

int locust_peer_barrier(locust_ctx *c, void *const *peer_flags, uint32_t rank, uint32_t world, uint32_t epoch) {
    if (!c || !peer_flags || world == 0 || world > 256 || rank >= world) return LOCUST_E_ARG;
    for (uint32_t i = 0; i < world; ++i)
        if (!peer_flags[i] || reinterpret_cast<uintptr_t>(peer_flags[i]) % 4) return fail(c, LOCUST_E_ARG, "locust_peer_barrier: NULL or unaligned flag array");
    CU(cudaSetDevice(c->device));
    int rc = stage_pointers(c, c->peer_flag_host, 256, world, peer_flags);
    if (rc) return rc;
    c->launches += (uint64_t)launch_peer_barrier(reinterpret_cast<uint32_t *const *>(c->peer_ptrs + 256), rank, world, epoch, c->xflag, c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;
}

static uint64_t result_part_bytes(uint64_t n, uint64_t blob) { return (32ull + 8ull * (n + 1) + 16ull * n + blob + 31ull) & ~31ull; }

int locust_result_part_bytes(locust_ctx *c, uint64_t *bytes) {
    if (!c || !bytes) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "locust_result_part_bytes before locust_reduce");
    *bytes = result_part_bytes(c->h_ctr->n_out, c->h_ctr->blob_bytes);
    return LOCUST_OK;
}

int locust_result_part_pack(locust_ctx *c, void *dev_out) {
    if (!c || !dev_out) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "locust_result_part_pack before locust_reduce");
    if (reinterpret_cast<uintptr_t>(dev_out) % 8) return fail(c, LOCUST_E_ARG, "locust_result_part_pack: dev_out must be 8-byte aligned");
    CU(cudaSetDevice(c->device));
    c->launches += (uint64_t)launch_result_pack(c->red, c->ctr, static_cast<uint8_t *>(dev_out), c->h_ctr->n_out + (c->h_ctr->blob_bytes >> 4),
                                                c->num_sms, c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;
}

int locust_result_concat(locust_ctx *c, uint32_t nparts, const void *const *dev_parts, const uint64_t *n_keys,
                         const uint64_t *blob_bytes, const uint64_t *totals) {
    if (!c || !nparts || !dev_parts || !n_keys || !blob_bytes || !totals) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "locust_result_concat before locust_reduce");
    CU(cudaSetDevice(c->device));
    uint64_t n = 0, blob = 0, total = 0;
    for (uint32_t i = 0; i < nparts; ++i) {
        if (!dev_parts[i] || reinterpret_cast<uintptr_t>(dev_parts[i]) % 8)
            return fail(c, LOCUST_E_ARG, "locust_result_concat: parts must be non-NULL and 8-byte aligned");
        n += n_keys[i]; blob += blob_bytes[i]; total += totals[i];
    }
    // room for the whole result: the arrays of REDUCE are replaced if the sum does not fit
    if (n > c->rec.cap) {
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->red.key_offsets); cudaFree(c->red.counts); cudaFree(c->red.vals);
        c->red.key_offsets = nullptr; c->red.counts = nullptr; c->red.vals = nullptr;
        CU(dalloc(&c->red.key_offsets, n + 1)); CU(dalloc(&c->red.counts, n)); CU(dalloc(&c->red.vals, n));
    }
    if (blob > c->red.blob_cap) {
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->red.key_blob); c->red.key_blob = nullptr;
        c->red.blob_cap = blob + 64;
        CU(dalloc(&c->red.key_blob, c->red.blob_cap));
    }
    uint64_t nb = 0, bb = 0, tb = 0;
    for (uint32_t i = 0; i < nparts; ++i) {
        c->launches += (uint64_t)launch_result_append(c->red, static_cast<const uint8_t *>(dev_parts[i]), nb, bb, tb,
                                                      n_keys[i] + (blob_bytes[i] >> 4), c->num_sms, c->stream);
        nb += n_keys[i]; bb += blob_bytes[i]; tb += totals[i];
    }
    CU(cudaGetLastError());
    c->h_ctr->n_out = n; c->h_ctr->blob_bytes = blob; c->h_ctr->total = total;
    CU(cudaMemcpyAsync(&c->ctr->n_out, &c->h_ctr->n_out, 3 * sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->host_valid = false;
    return LOCUST_OK;
}

int locust_result_header(locust_ctx *c, void *dev_header) {
    if (!c || !dev_header) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "locust_result_header before locust_reduce");
    if (reinterpret_cast<uintptr_t>(dev_header) % 8) return fail(c, LOCUST_E_ARG, "locust_result_header: dev_header must be 8-byte aligned");
    CU(cudaSetDevice(c->device));
    c->launches += (uint64_t)launch_result_header(c->ctr, static_cast<unsigned long long *>(dev_header), c->stream);
    CU(cudaGetLastError());
    return LOCUST_OK;
}

int locust_result_adopt_slots(locust_ctx *c, const void *dev_slots, uint64_t slot_bytes, uint32_t nslots) {
    if (!c || !dev_slots || nslots == 0 || nslots > 256) return LOCUST_E_ARG;
    if (c->stage != 2) return fail(c, LOCUST_E_STATE, "locust_result_adopt_slots before locust_reduce");
    if (slot_bytes < 64 || slot_bytes % 32 || reinterpret_cast<uintptr_t>(dev_slots) % 8)
        return fail(c, LOCUST_E_ARG, "locust_result_adopt_slots: slots are multiples of 32 bytes, 8-byte aligned");
    CU(cudaSetDevice(c->device));
    uint64_t h[256 * 4];
    CU(cudaMemcpy2DAsync(h, 32, dev_slots, slot_bytes, 32, nslots, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(c->h_xflag, c->xflag, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (*c->h_xflag & 4u) return fail(c, LOCUST_E_STATE, "locust_peer_barrier: a rank did not arrive within five seconds");
    const void *parts[256];
    uint64_t n[256], blob[256], total[256];
    for (uint32_t i = 0; i < nslots; ++i) {
        if (h[4 * i + 3] > slot_bytes)
            return fail(c, LOCUST_E_CAPACITY, "locust_result_adopt_slots: a rank's slice did not fit its slot (raise the slot size)");
        if (h[4 * i + 3] != result_part_bytes(h[4 * i], h[4 * i + 1]))
            return fail(c, LOCUST_E_STATE, "locust_result_adopt_slots: a slot does not hold a packed result");
        parts[i] = static_cast<const uint8_t *>(dev_slots) + (uint64_t)i * slot_bytes;
        n[i] = h[4 * i]; blob[i] = h[4 * i + 1]; total[i] = h[4 * i + 2];
    }
    return locust_result_concat(c, nslots, parts, n, blob, total);
}

int locust_shuffle_begin_merge(locust_ctx *c) {
    if (!c) return LOCUST_E_ARG;
    if (c->stage != 0) return fail(c, LOCUST_E_STATE, "locust_shuffle_begin_merge after locust_process");
    CU(cudaSetDevice(c->device));
    CU(cudaMemsetAsync(c->table, 0, c->slots * sizeof(Slot), c->stream));
    CU(clear_long_index(c));
    CU(cudaMemsetAsync(c->ctr, 0, sizeof(Counters), c->stream));
    c->cleared = true;
    c->shuffled = true;
    c->part_n = 0;
    return LOCUST_OK;
}

int locust_shuffle_merge(locust_ctx *c, const void *dev_records, uint64_t nbytes) {
    if (!c || (!dev_records && nbytes)) return LOCUST_E_ARG;
    if (c->stage != 0 || !c->shuffled) return fail(c, LOCUST_E_STATE, "locust_shuffle_merge needs locust_shuffle_begin_merge first");
    if (nbytes == 0) return LOCUST_OK;
    if (nbytes < 32 || nbytes % 32) return fail(c, LOCUST_E_ARG, "a packed part is a multiple of 32 bytes");
    CU(cudaSetDevice(c->device));
    MapParams p = map_params(c, static_cast<const uint8_t *>(dev_records), 0, 0);
    c->launches += (uint64_t)launch_shuffle_merge(static_cast<const uint8_t *>(dev_records), nbytes, p, c->num_sms, c->stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev[7], c->stream)); c->ev_set[7] = true;
    return LOCUST_OK;
}

}  // extern "C"
