// locust_multi.cu -- one word-count job over several GPUs of one box, driven from ONE process
// through the C ABI (include/locust.h, locust_multi_*).
//
// Stands in for the reference's multi mode: per-node `[line_start, line_end)` windows
// (MapReduce/src/main.cu:367-385), sorted partials handed over in /tmp/out.txt (:421-446) and the TCP
// command runner that starts the nodes (Distributor/slave.py:30-32).  Here a "node" is a GPU:
//
//   shard cuts      the host buffer is cut at r*N/P, moved forward to the byte after a delimiter
//   map             every GPU tokenises and combines its shard (locust_map on its own context)
//   key ranges      two-byte-prefix histograms are put into every GPU's memory (peer stores), summed
//                   locally, and turned into P key ranges of about equal size
//   exchange        counted first (which also lets a full table grow and redo MAP while that is still
//                   possible), then the packing kernel of GPU s stores the records of range d straight
//                   into GPU d's receive slot s over NVLink; slots are sized from the counts
//   merge, sort     every GPU merges what it received and runs PROCESS + REDUCE on its own range
//   result          the sorted slices are packed into slots on GPU 0 and strung together there
//
// One host thread per GPU drives that GPU's context and stream.  Where the streams have to meet
// (histograms in place, records in place, slices in place) every thread records an event, the threads
// meet on the host, and every stream waits for every event: no kernel ever waits for another GPU, so
// the same code runs with several contexts on one device (devices = {0, 0, 0}), which is how the
// single-GPU test box exercises it.
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "common.cuh"
#include "../../include/locust.h"

namespace {

constexpr uint64_t kHistBytes = (uint64_t)LOCUST_PARTITION_BINS * 8ull;

class HostBarrier {
  public:
    explicit HostBarrier(int n) : n_(n) {}
    void wait() {
        std::unique_lock<std::mutex> lk(mu_);
        const uint64_t gen = gen_;
        if (++count_ == n_) { count_ = 0; ++gen_; cv_.notify_all(); }
        else cv_.wait(lk, [&] { return gen_ != gen; });
    }

  private:
    std::mutex mu_;
    std::condition_variable cv_;
    int n_, count_ = 0;
    uint64_t gen_ = 0;
};

}  // namespace

struct locust_multi {
    int n = 0;
    std::vector<int> devices;
    std::vector<locust_ctx *> ctx;
    std::vector<cudaStream_t> stream;
    std::vector<cudaEvent_t> ev;               // one per rank, re-recorded at every meeting point
    cudaEvent_t t0 = nullptr, t1 = nullptr;    // on rank 0's stream
    std::vector<uint8_t *> hbuf;               // per rank: P histogram slots + this rank's own histogram
    std::vector<uint8_t *> xbuf;               // per rank: P receive slots of slot_bytes
    uint64_t short_cap = 0, long_room = 0, slot_bytes = 0;
    uint8_t *rbuf = nullptr;                   // on rank 0: P result slots of result_slot bytes
    uint64_t result_slot = 0;
    // per-job scratch shared by the threads
    std::vector<uint64_t> n_short, long_bytes, part_bytes;     // [sender * P + dest]
    std::vector<uint64_t> slice_bytes;
    std::vector<int> rc;
    std::vector<std::string> errs;
    std::string err;
    float ms = 0.f;
    uint64_t bytes_sent = 0;                   // record bytes that left their GPU in the last job
};

namespace {

int mfail(locust_multi *m, int code, const std::string &what) {
    if (m) m->err = what;
    return code;
}

struct Job {
    locust_multi *m;
    const uint8_t *const *shards;      // per rank
    const uint64_t *nbytes;
    int where;
    HostBarrier *bar;
};

// every stream of the job waits for everything enqueued so far on every other stream
void meet(const Job &j, int r) {
    locust_multi *m = j.m;
    cudaEventRecord(m->ev[r], m->stream[r]);
    j.bar->wait();
    for (int q = 0; q < m->n; ++q)
        if (q != r) cudaStreamWaitEvent(m->stream[r], m->ev[q], 0);
    j.bar->wait();                      // nobody re-records its event before everyone has queued its waits
}

bool anyone_failed(const locust_multi *m) {
    for (int r = 0; r < m->n; ++r)
        if (m->rc[r] != LOCUST_OK) return true;
    return false;
}

#define STEP(call)                                                                       \
    do {                                                                                 \
        if (m->rc[r] == LOCUST_OK) {                                                     \
            const int rc_ = (call);                                                      \
            if (rc_ != LOCUST_OK) { m->rc[r] = rc_; m->errs[r] = std::string(#call) + ": " + locust_last_error(c); } \
        }                                                                                \
    } while (0)
#define CUSTEP(call)                                                                     \
    do {                                                                                 \
        if (m->rc[r] == LOCUST_OK) {                                                     \
            const cudaError_t e_ = (call);                                               \
            if (e_ != cudaSuccess) { m->rc[r] = LOCUST_E_CUDA; m->errs[r] = std::string(#call) + ": " + cudaGetErrorString(e_); } \
        }                                                                                \
    } while (0)

// (re)allocation of the receive slots, done by rank 0's thread between two host barriers
void size_exchange(locust_multi *m) {
    const int P = m->n;
    uint64_t s_max = 0, l_max = 0;
    for (int i = 0; i < P * P; ++i) {
        if (m->n_short[i] > s_max) s_max = m->n_short[i];
        if (m->long_bytes[i] > l_max) l_max = m->long_bytes[i];
    }
    if (s_max <= m->short_cap && l_max <= m->long_room && m->slot_bytes) return;
    for (int q = 0; q < P; ++q) {
        cudaSetDevice(m->devices[q]);
        cudaStreamSynchronize(m->stream[q]);
        cudaFree(m->xbuf[q]);
        m->xbuf[q] = nullptr;
    }
    m->short_cap = s_max + (s_max >> 2) + 1024;
    m->long_room = (l_max + (l_max >> 2) + 4096 + 31) & ~31ull;
    m->slot_bytes = 32 + 32 * m->short_cap + m->long_room;
    for (int q = 0; q < P; ++q) {
        cudaSetDevice(m->devices[q]);
        cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&m->xbuf[q]), m->slot_bytes * (uint64_t)P);
        if (e == cudaSuccess) e = cudaMemset(m->xbuf[q], 0, m->slot_bytes * (uint64_t)P);
        if (e != cudaSuccess) { m->rc[0] = LOCUST_E_NOMEM; m->errs[0] = std::string("exchange buffers: ") + cudaGetErrorString(e); return; }
    }
}

void size_result(locust_multi *m) {
    uint64_t need = 0;
    for (int r = 0; r < m->n; ++r)
        if (m->slice_bytes[r] > need) need = m->slice_bytes[r];
    if (need <= m->result_slot) return;
    cudaSetDevice(m->devices[0]);
    cudaStreamSynchronize(m->stream[0]);
    cudaFree(m->rbuf);
    m->rbuf = nullptr;
    m->result_slot = (need + (need >> 2) + 4096 + 31) & ~31ull;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&m->rbuf), m->result_slot * (uint64_t)m->n);
    if (e != cudaSuccess) { m->rc[0] = LOCUST_E_NOMEM; m->errs[0] = std::string("result slots: ") + cudaGetErrorString(e); }
}

void rank_main(Job j, int r) {
    locust_multi *m = j.m;
    const int P = m->n;
    locust_ctx *c = m->ctx[r];
    cudaSetDevice(m->devices[r]);
    if (r == 0) cudaEventRecord(m->t0, m->stream[0]);
    STEP(locust_reset(c));
    if (j.nbytes[r]) STEP(locust_map(c, j.shards[r], j.nbytes[r], j.where));
    if (P == 1) {
        STEP(locust_process(c));
        STEP(locust_reduce(c));
        cudaEventRecord(m->t1, m->stream[0]);
        return;
    }
    // key ranges: every rank's histogram into every rank's slot r, summed after the meeting point
    uint8_t *own = m->hbuf[r] + (uint64_t)P * kHistBytes;
    CUSTEP(cudaMemsetAsync(own, 0, kHistBytes, m->stream[r]));
    STEP(locust_partition_histogram(c, own));
    {
        std::vector<void *> dst(P);
        for (int q = 0; q < P; ++q) dst[q] = m->hbuf[q] + (uint64_t)r * kHistBytes;
        STEP(locust_peer_put(c, (uint32_t)P, dst.data(), own, kHistBytes));
    }
    meet(j, r);
    STEP(locust_partition_by_range_sum(c, m->hbuf[r], (uint32_t)P, (uint32_t)P));
    // count (a table that overflowed in MAP grows and is redone here), then size the receive slots from the counts
    STEP(locust_shuffle_sizes(c, (uint32_t)P, &m->part_bytes[(size_t)r * P]));
    STEP(locust_shuffle_counts(c, (uint32_t)P, &m->n_short[(size_t)r * P], &m->long_bytes[(size_t)r * P]));
    j.bar->wait();
    if (r == 0 && !anyone_failed(m)) size_exchange(m);
    j.bar->wait();
    if (anyone_failed(m)) { if (m->rc[r] == LOCUST_OK) m->rc[r] = LOCUST_E_STATE; }
    {
        std::vector<void *> dst(P);
        for (int q = 0; q < P; ++q) dst[q] = m->xbuf[q] ? m->xbuf[q] + (uint64_t)r * m->slot_bytes : nullptr;
        STEP(locust_shuffle_pack_peers(c, (uint32_t)P, dst.data(), m->slot_bytes, m->short_cap));
    }
    meet(j, r);
    STEP(locust_shuffle_begin_merge(c));
    for (int s = 0; s < P; ++s) STEP(locust_shuffle_merge(c, m->xbuf[r] + (uint64_t)s * m->slot_bytes, m->slot_bytes));
    STEP(locust_process(c));
    STEP(locust_reduce(c));
    // sorted slices into rank 0's slots
    m->slice_bytes[r] = 0;
    STEP(locust_result_part_bytes(c, &m->slice_bytes[r]));
    j.bar->wait();
    if (r == 0 && !anyone_failed(m)) size_result(m);
    j.bar->wait();
    if (anyone_failed(m)) { if (m->rc[r] == LOCUST_OK) m->rc[r] = LOCUST_E_STATE; }
    STEP(locust_result_part_pack(c, m->rbuf + (uint64_t)r * m->result_slot));
    meet(j, r);
    if (r == 0) {
        STEP(locust_result_adopt_slots(c, m->rbuf, m->result_slot, (uint32_t)P));
        cudaEventRecord(m->t1, m->stream[0]);
    }
}

int run_job(locust_multi *m, const uint8_t *const *shards, const uint64_t *nbytes, int where) {
    const int P = m->n;
    HostBarrier bar(P);
    Job j{m, shards, nbytes, where, &bar};
    for (int r = 0; r < P; ++r) { m->rc[r] = LOCUST_OK; m->errs[r].clear(); }
    std::vector<std::thread> th;
    for (int r = 1; r < P; ++r) th.emplace_back(rank_main, j, r);
    rank_main(j, 0);
    for (auto &t : th) t.join();
    for (int r = 0; r < P; ++r)
        if (m->rc[r] != LOCUST_OK && !m->errs[r].empty())
            return mfail(m, m->rc[r], "rank " + std::to_string(r) + ": " + m->errs[r]);
    for (int r = 0; r < P; ++r)
        if (m->rc[r] != LOCUST_OK) return mfail(m, m->rc[r], "rank " + std::to_string(r) + " stopped because another rank failed");
    cudaSetDevice(m->devices[0]);
    cudaEventSynchronize(m->t1);
    cudaEventElapsedTime(&m->ms, m->t0, m->t1);
    m->bytes_sent = 0;
    if (P > 1)
        for (int s = 0; s < P; ++s)
            for (int d = 0; d < P; ++d)
                if (s != d) m->bytes_sent += m->part_bytes[(size_t)s * P + d];
    return LOCUST_OK;
}

}  // namespace

extern "C" {

int locust_multi_create(const int *devices, int n, const locust_config *cfg, locust_multi **out) {
    if (!devices || n < 1 || n > 256 || !out) return LOCUST_E_ARG;
    locust_multi *m = new (std::nothrow) locust_multi();
    if (!m) return LOCUST_E_NOMEM;
    *out = m;                                    // handed out even on failure so the error can be read
    m->n = n;
    m->devices.assign(devices, devices + n);
    m->ctx.assign(n, nullptr);
    m->stream.assign(n, nullptr);
    m->ev.assign(n, nullptr);
    m->hbuf.assign(n, nullptr);
    m->xbuf.assign(n, nullptr);
    m->n_short.assign((size_t)n * n, 0); m->long_bytes.assign((size_t)n * n, 0); m->part_bytes.assign((size_t)n * n, 0);
    m->slice_bytes.assign(n, 0);
    m->rc.assign(n, LOCUST_OK);
    m->errs.assign(n, std::string());
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return mfail(m, LOCUST_E_CUDA, "no usable CUDA device");
    for (int r = 0; r < n; ++r)
        if (devices[r] < 0 || devices[r] >= ndev) return mfail(m, LOCUST_E_ARG, "locust_multi_create: device ordinal out of range");
    // every GPU stores into every other GPU's receive slots
    for (int a = 0; a < n; ++a) {
        if (cudaSetDevice(devices[a]) != cudaSuccess) return mfail(m, LOCUST_E_CUDA, "cudaSetDevice");
        for (int b = 0; b < n; ++b) {
            if (devices[a] == devices[b]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[a], devices[b]);
            if (!can) return mfail(m, LOCUST_E_CUDA, "locust_multi_create: the devices cannot access each other's memory");
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return mfail(m, LOCUST_E_CUDA, "cudaDeviceEnablePeerAccess");
            cudaGetLastError();
        }
    }
    for (int r = 0; r < n; ++r) {
        locust_config c1{};
        if (cfg) c1 = *cfg;
        c1.device = devices[r];
        c1.stream = nullptr;                    // one private stream per rank
        const int rc = locust_create(&c1, &m->ctx[r]);
        if (rc != LOCUST_OK) return mfail(m, rc, std::string("locust_create: ") + (m->ctx[r] ? locust_last_error(m->ctx[r]) : "no context"));
        m->stream[r] = static_cast<cudaStream_t>(locust_stream(m->ctx[r]));
        cudaSetDevice(devices[r]);
        if (cudaEventCreateWithFlags(&m->ev[r], cudaEventDisableTiming) != cudaSuccess) return mfail(m, LOCUST_E_CUDA, "cudaEventCreate");
        if (cudaMalloc(reinterpret_cast<void **>(&m->hbuf[r]), (uint64_t)(n + 1) * kHistBytes) != cudaSuccess)
            return mfail(m, LOCUST_E_NOMEM, "histogram slots");
    }
    cudaSetDevice(devices[0]);
    if (cudaEventCreate(&m->t0) != cudaSuccess || cudaEventCreate(&m->t1) != cudaSuccess) return mfail(m, LOCUST_E_CUDA, "cudaEventCreate");
    return LOCUST_OK;
}

int locust_multi_destroy(locust_multi *m) {
    if (!m) return LOCUST_E_ARG;
    for (int r = 0; r < m->n; ++r) {
        cudaSetDevice(m->devices[r]);
        if (m->stream[r]) cudaStreamSynchronize(m->stream[r]);
    }
    for (int r = 0; r < m->n; ++r) {
        cudaSetDevice(m->devices[r]);
        if (m->ctx[r]) locust_destroy(m->ctx[r]);
        if (m->ev[r]) cudaEventDestroy(m->ev[r]);
        cudaFree(m->hbuf[r]);
        cudaFree(m->xbuf[r]);
    }
    cudaSetDevice(m->devices[0]);
    cudaFree(m->rbuf);
    if (m->t0) cudaEventDestroy(m->t0);
    if (m->t1) cudaEventDestroy(m->t1);
    delete m;
    return LOCUST_OK;
}

int locust_multi_wordcount(locust_multi *m, const void *host_data, uint64_t nbytes) {
    if (!m || (!host_data && nbytes)) return LOCUST_E_ARG;
    const uint8_t *d = static_cast<const uint8_t *>(host_data);
    const int P = m->n;
    // shard r = [cut[r], cut[r + 1]): nominal r*N/P moved forward to the byte after a delimiter, so that no
    // token straddles two GPUs ('\n' where one is near, any delimiter otherwise -- identical results)
    std::vector<uint64_t> cut(P + 1, 0);
    for (int r = 1; r < P; ++r) {
        uint64_t c = nbytes / (uint64_t)P * (uint64_t)r;
        if (c < cut[r - 1]) c = cut[r - 1];
        while (c > 0 && c < nbytes && !locust::is_delim(d[c - 1])) ++c;
        cut[r] = c;
    }
    cut[P] = nbytes;
    std::vector<const uint8_t *> shards(P);
    std::vector<uint64_t> nb(P);
    for (int r = 0; r < P; ++r) { shards[r] = d + cut[r]; nb[r] = cut[r + 1] - cut[r]; }
    return run_job(m, shards.data(), nb.data(), LOCUST_HOST);
}

int locust_multi_wordcount_device(locust_multi *m, const void *const *dev_shards, const uint64_t *nbytes) {
    if (!m || !dev_shards || !nbytes) return LOCUST_E_ARG;
    for (int r = 0; r < m->n; ++r)
        if (!dev_shards[r] && nbytes[r]) return mfail(m, LOCUST_E_ARG, "locust_multi_wordcount_device: NULL shard");
    return run_job(m, reinterpret_cast<const uint8_t *const *>(dev_shards), nbytes, LOCUST_DEVICE);
}

locust_ctx *locust_multi_context(locust_multi *m, int rank) { return (m && rank >= 0 && rank < m->n) ? m->ctx[rank] : nullptr; }

int locust_multi_stats(locust_multi *m, float *job_ms, uint64_t *record_bytes_exchanged) {
    if (!m) return LOCUST_E_ARG;
    if (job_ms) *job_ms = m->ms;
    if (record_bytes_exchanged) *record_bytes_exchanged = m->bytes_sent;
    return LOCUST_OK;
}

const char *locust_multi_last_error(locust_multi *m) { return m ? m->err.c_str() : "null handle"; }

}  // extern "C"
