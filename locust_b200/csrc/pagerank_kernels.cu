// pagerank_kernels.cu -- second MapReduce job on the word-count machinery: PageRank over an edge list
// ("src dst" per line), in 64-bit fixed point so that the result does not depend on the order of the sums.
// The reference names the job as a milestone (docs/PROPOSAL.md:21, README.md:6) and holds no code for it;
// oracle/pagerank_oracle.c is the definition these kernels are checked against.
//
//   parse    one pass over the text: a thread owns the lines that start in its 16 bytes, takes the first two
//            tokens, finds or adds both in the combiner table (out-degrees are counted from the edge list afterwards) and
//            appends the edge as a pair of slot numbers.  Edges of one 4 KiB tile are placed by a block scan,
//            so the edge list is dense and its counter sees one atomic per tile.
//   nodes    per slot: rank, contribution = rank / out-degree, mass of the nodes without out-edges
//   scatter  per edge: acc[dst] += contrib[src]; edges into heavily linked nodes are added up per block in shared
//            memory first (their atomics would queue on one L2 address)
//   after the last iteration the ranks are stored in Slot::count, and PROCESS + REDUCE of the word-count path
//   put the (node, rank) records in byte order.
#include "kernels.cuh"

namespace locust {

namespace {

constexpr uint32_t kPrProbe = 1u << 14;
constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr unsigned long long kMass = 1ull << 48;

LC_D void pr_cas128(Slot *s, uint64_t n0, uint64_t n1, uint64_t &o0, uint64_t &o1) {
    asm volatile(
        "{\n\t"
        ".reg .b128 c, n, d;\n\t"
        "mov.b128 c, {%3, %3};\n\t"
        "mov.b128 n, {%4, %5};\n\t"
        "atom.global.relaxed.gpu.cas.b128 d, [%2], c, n;\n\t"
        "mov.b128 {%0, %1}, d;\n\t"
        "}\n"
        : "=l"(o0), "=l"(o1) : "l"(s), "l"(0ull), "l"(n0), "l"(n1) : "memory");
}

// slot of node (k0, k1), added if new.  0xFFFFFFFF: the table is full.
__device__ __noinline__ uint32_t pr_find_or_add(Slot *table, uint64_t table_mask, Counters *ctr, uint64_t k0, uint64_t k1) {
    uint64_t slot = slot_hash32(k0, k1) & table_mask;
    const uint64_t lim = table_mask + 1ull < (uint64_t)kPrProbe ? table_mask + 1ull : (uint64_t)kPrProbe;
    for (uint64_t probe = 0; probe < lim; ++probe) {
        Slot *s = table + slot;
        const ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2 *>(s));
        bool mine = cur.x == k0 && cur.y == k1;
        if (!mine && (cur.x | cur.y) == 0ull) {
            uint64_t o0, o1;
            pr_cas128(s, k0, k1, o0, o1);
            mine = (o0 | o1) == 0ull || (o0 == k0 && o1 == k1);
        }
        if (mine) return (uint32_t)slot;
        slot = (slot + 1) & table_mask;
    }
    atomicOr(&ctr->status, kStatusTableFull);
    return 0xFFFFFFFFu;
}

// first two tokens of the line that starts at q; false when the line has fewer
__device__ bool pr_parse_line(const PrParams &p, uint64_t q, uint64_t key[4]) {
    uint64_t at = q;
#pragma unroll 1
    for (int t = 0; t < 2; ++t) {
        uint32_t b = 0;
        while (at < p.hi) {
            b = p.base[at];
            if (b == '\n' || !is_delim(b)) break;
            ++at;
        }
        if (at >= p.hi || b == '\n') return false;
        uint64_t k0 = 0, k1 = 0;
        uint32_t len = 0;
        while (at < p.hi) {
            b = p.base[at];
            if (is_delim(b)) break;
            if (len < 8u) k0 |= (uint64_t)b << (8u * len);
            else if (len < 16u) k1 |= (uint64_t)b << (8u * (len - 8u));
            ++len; ++at;
        }
        if (len > 16u) { atomicOr(&p.ctr->status, kStatusNodeLong); return false; }
        key[2 * t] = k0; key[2 * t + 1] = k1;
    }
    return true;
}

__global__ void __launch_bounds__(256) pr_parse_kernel(PrParams p) {
    __shared__ uint32_t warp_tot[8];
    __shared__ unsigned long long tile_base;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint64_t first = p.lo & ~15ull;
    const uint64_t tiles = (p.hi - first + 4095ull) / 4096ull;
    for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const uint64_t c = first + tile * 4096ull + (uint64_t)threadIdx.x * 16ull;
        uint32_t starts = 0;
        if (c < p.hi) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p.base + c));
            const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
            uint32_t nl = 0;
#pragma unroll
            for (int j = 0; j < 16; ++j)
                if (((wv[j >> 2] >> (8 * (j & 3))) & 0xFFu) == '\n') nl |= 1u << j;
            starts = (nl << 1) & 0xFFFFu;
            if (c > p.lo && p.base[c - 1] == '\n') starts |= 1u;
            if (p.lo >= c && p.lo < c + 16ull) starts |= 1u << (uint32_t)(p.lo - c);
            // only positions inside [lo, hi)
            if (c < p.lo) starts &= ~((1u << (uint32_t)(p.lo - c)) - 1u);
            if (c + 16ull > p.hi) starts &= (1u << (uint32_t)(p.hi - c)) - 1u;
        }
        // an edge line has at least four bytes: at most four of them start in 16 bytes
        uint2 e0 = make_uint2(0, 0), e1 = e0, e2 = e0, e3 = e0;
        uint32_t nv = 0;
        while (starts) {
            const uint32_t j = (uint32_t)__ffs((int)starts) - 1u;
            starts &= starts - 1u;
            uint64_t key[4];
            if (!pr_parse_line(p, c + j, key)) continue;
            const uint32_t s = pr_find_or_add(p.table, p.table_mask, p.ctr, key[0], key[1]);
            const uint32_t d = pr_find_or_add(p.table, p.table_mask, p.ctr, key[2], key[3]);
            if (s == 0xFFFFFFFFu || d == 0xFFFFFFFFu) continue;          // table full: the job is redone
            const uint2 e = make_uint2(s, d);
            if (nv == 0) e0 = e; else if (nv == 1) e1 = e; else if (nv == 2) e2 = e; else e3 = e;
            ++nv;
        }
        // block scan of nv
        uint32_t inc = nv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        if (lane == 31) warp_tot[w] = inc;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t tot = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) tot += warp_tot[k];
            tile_base = tot ? atomicAdd(&p.st->n_edges, (unsigned long long)tot) : 0ull;
        }
        __syncthreads();
        unsigned long long at = tile_base + (inc - nv);
        for (uint32_t k = 0; k < w; ++k) at += warp_tot[k];
        if (nv > 0 && at + nv <= p.edge_cap) {
            p.edges[at] = e0;
            if (nv > 1) p.edges[at + 1] = e1;
            if (nv > 2) p.edges[at + 2] = e2;
            if (nv > 3) p.edges[at + 3] = e3;
        } else if (nv > 0) {
            atomicOr(&p.ctr->status, kStatusEntryFull);
        }
        __syncthreads();            // warp_tot and tile_base are reused by the next tile
    }
}

// nodes = occupied slots, counted once behind the parse (an atomic per new node would queue up on one address)
__global__ void __launch_bounds__(256) pr_count_nodes_kernel(const Slot *__restrict__ table, uint64_t slots, PrState *st) {
    __shared__ uint32_t warp_cnt[8];
    uint32_t mine = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * 256u + threadIdx.x; i < slots; i += (uint64_t)gridDim.x * 256u) {
        const ulonglong2 k = __ldcg(reinterpret_cast<const ulonglong2 *>(table + i));
        mine += (k.x | k.y) != 0ull;
    }
    mine = __reduce_add_sync(kFull, mine);
    if ((threadIdx.x & 31u) == 0) warp_cnt[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += warp_cnt[k];
        if (t) atomicAdd(&st->n_nodes, (unsigned long long)t);
    }
}

// Degrees in one pass over the edge list.  Out-degree: Slot::count of every source += its edges (the lanes of a warp
// that hold the same source send one atomic between them).  In-degree: counted into acc, which the first nodes pass
// empties again; it only serves to find the heavily linked nodes.
__global__ void __launch_bounds__(256) pr_degree_kernel(const uint2 *__restrict__ edges, Slot *table, PrArrays a, const PrState *st) {
    const unsigned long long n = st->n_edges;
    const unsigned long long stride = (unsigned long long)gridDim.x * 256u;
    const unsigned long long rounds = (n + stride - 1) / stride;
    unsigned long long i = (unsigned long long)blockIdx.x * 256u + threadIdx.x;
    for (unsigned long long r = 0; r < rounds; ++r, i += stride) {
        uint2 e = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
        if (i < n) e = __ldg(edges + i);
        const uint32_t m = __match_any_sync(kFull, e.x);
        if ((uint32_t)__ffs((int)m) - 1u == (threadIdx.x & 31u) && e.x != 0xFFFFFFFFu)
            atomicAdd(&table[e.x].count, (unsigned long long)__popc(m));
        if (e.y != 0xFFFFFFFFu) atomicAdd(a.acc + e.y, 1ull);
    }
}

// An edge into a node with at least kHotDegree in-edges gets the top bit of its destination set: the scatter
// passes add such edges up in shared memory before they touch the node's accumulator.
constexpr unsigned long long kHotDegree = 8192;
constexpr uint32_t kHotBit = 0x80000000u;
__global__ void __launch_bounds__(256) pr_mark_hot_kernel(uint2 *edges, PrArrays a, PrState *st) {
    __shared__ uint32_t warp_cnt[8];
    const unsigned long long n = st->n_edges;
    uint32_t mine = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256u + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * 256u) {
        const uint32_t d = edges[i].y;
        if (__ldcg(a.acc + d) >= kHotDegree) { edges[i].y = d | kHotBit; ++mine; }
    }
    mine = __reduce_add_sync(kFull, mine);
    if ((threadIdx.x & 31u) == 0) warp_cnt[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += warp_cnt[k];
        if (t) atomicAdd(&st->n_hot, (unsigned long long)t);
    }
}

LC_D unsigned long long mulq32(unsigned long long x, unsigned long long d) {
    return (__umul64hi(x, d) << 32) | ((x * d) >> 32);
}

// step 0 sets the start ranks; step k > 0 closes iteration k - 1.  `last`: store the ranks in the table.
__global__ void __launch_bounds__(256) pr_nodes_kernel(Slot *table, uint64_t slots, PrArrays a, PrState *st, uint32_t step,
                                                       unsigned long long damping, int last) {
    __shared__ unsigned long long warp_sum[8];
    const unsigned long long n = st->n_nodes;
    if (n == 0) return;
    unsigned long long fixed;          // start rank (step 0) or the teleport term
    if (step == 0) fixed = kMass / n;
    else fixed = ((kMass - mulq32(kMass, damping)) + mulq32(st->dangling[(step - 1u) & 1u], damping)) / n;
    unsigned long long dsum = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * 256u + threadIdx.x; i < slots; i += (uint64_t)gridDim.x * 256u) {
        const ulonglong4 v = ld256(table + i);
        if ((v.x | v.y) == 0ull) continue;
        const unsigned long long r = step == 0 ? fixed : fixed + mulq32(a.acc[i], damping);
        a.acc[i] = 0ull;
        if (last) { table[i].count = r; continue; }
        a.rank[i] = r;
        const unsigned long long deg = v.z;
        a.contrib[i] = deg ? r / deg : 0ull;
        if (!deg) dsum += r;
    }
    if (last) return;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(kFull, dsum, o);
    if ((threadIdx.x & 31u) == 0) warp_sum[threadIdx.x >> 5] = dsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += warp_sum[k];
        if (t) atomicAdd(&st->dangling[step & 1u], t);
    }
}

// iteration `it`: acc[dst] += contrib[src]; also empties the dangling sum the closing nodes pass will fill.
// One atomic per edge, except for edges into heavily linked nodes (marked by pr_mark_hot_kernel): atomics on one
// address queue up in L2 (32 M edges over 1 M nodes: 0.48 ms per iteration with uniform endpoints, 1.75 ms when a few
// hundred nodes receive 40 % of the edges), so a block adds those up in a direct-mapped table in shared memory over
// a chunk of 8192 edges and sends one atomic per node and chunk.  A destination that finds its place taken by another
// node goes to global memory directly.  All sums are integer sums: the result does not depend on any of this.
// (Letting the lanes of a warp that hold the same destination add up first -- MATCH.ANY + REDUX -- was measured and
// dropped: 1.40 / 1.43 ms, the match costs more than the atomics it saves; profiles/r02_pagerank.txt.)
constexpr uint32_t kHotSlots = 1024, kScatterChunk = 256 * 32;
__global__ void __launch_bounds__(256) pr_scatter_kernel(const uint2 *__restrict__ edges, PrArrays a, PrState *st, uint32_t it) {
    __shared__ uint32_t hkey[kHotSlots];
    __shared__ uint32_t hval[3][kHotSlots];          // 16-bit limbs of the sums: 64-bit shared atomics are compare-and-swap loops
    if (blockIdx.x == 0 && threadIdx.x == 0) st->dangling[(it + 1u) & 1u] = 0ull;
    const unsigned long long n = st->n_edges;
    if (st->n_hot == 0ull) {                         // no heavily linked node in this graph
        for (unsigned long long i = (unsigned long long)blockIdx.x * 256u + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * 256u) {
            const uint2 e = __ldg(edges + i);
            const unsigned long long cv = __ldcg(a.contrib + e.x);
            if (cv) atomicAdd(a.acc + e.y, cv);
        }
        return;
    }
    for (unsigned long long base = (unsigned long long)blockIdx.x * kScatterChunk; base < n; base += (unsigned long long)gridDim.x * kScatterChunk) {
        for (uint32_t j = threadIdx.x; j < kHotSlots; j += 256u) { hkey[j] = 0xFFFFFFFFu; hval[0][j] = 0u; hval[1][j] = 0u; hval[2][j] = 0u; }
        __syncthreads();
#pragma unroll 4
        for (uint32_t k = 0; k < kScatterChunk / 256u; ++k) {
            const unsigned long long i = base + k * 256u + threadIdx.x;
            if (i >= n) break;
            const uint2 e = __ldg(edges + i);
            const unsigned long long cv = __ldcg(a.contrib + e.x);
            if (!cv) continue;
            if (e.y & kHotBit) {
                const uint32_t d = e.y & ~kHotBit, h = d & (kHotSlots - 1u);
                const uint32_t old = atomicCAS(&hkey[h], 0xFFFFFFFFu, d);
                if (old == 0xFFFFFFFFu || old == d) {
                    // a contribution is below 2^48 and a chunk has 2^13 edges: three 16-bit limbs cannot overflow 32 bits
                    atomicAdd(&hval[0][h], (uint32_t)cv & 0xFFFFu);
                    atomicAdd(&hval[1][h], (uint32_t)(cv >> 16) & 0xFFFFu);
                    atomicAdd(&hval[2][h], (uint32_t)(cv >> 32));
                } else {
                    atomicAdd(a.acc + d, cv);
                }
            } else {
                atomicAdd(a.acc + e.y, cv);
            }
        }
        __syncthreads();
        for (uint32_t j = threadIdx.x; j < kHotSlots; j += 256u) {
            const uint32_t d = hkey[j];
            if (d != 0xFFFFFFFFu) {
                const unsigned long long v = (unsigned long long)hval[0][j] + ((unsigned long long)hval[1][j] << 16) + ((unsigned long long)hval[2][j] << 32);
                if (v) atomicAdd(a.acc + d, v);
            }
        }
        __syncthreads();
    }
}

}  // namespace

int launch_pagerank_parse(const PrParams &p, int num_sms, cudaStream_t st) {
    if (p.hi <= p.lo) return 0;
    const uint64_t tiles = (p.hi - (p.lo & ~15ull) + 4095ull) / 4096ull;
    const uint64_t want = (uint64_t)num_sms * 8u;
    pr_parse_kernel<<<(unsigned)(tiles < want ? tiles : want), 256, 0, st>>>(p);
    const uint64_t slots = p.table_mask + 1ull, blocks = (slots + 255u) / 256u;
    pr_count_nodes_kernel<<<(unsigned)(blocks < want ? blocks : want), 256, 0, st>>>(p.table, slots, p.st);
    return 2;
}

int launch_pagerank_iterate(Slot *table, uint64_t slots, uint2 *edges, uint64_t edge_cap, PrArrays a, PrState *state,
                            uint32_t iterations, unsigned long long damping, int num_sms, cudaStream_t st) {
    const unsigned gn = (unsigned)((slots + 255u) / 256u < (uint64_t)num_sms * 8u ? (slots + 255u) / 256u : (uint64_t)num_sms * 8u);
    unsigned ge = (unsigned)((edge_cap + 255u) / 256u < (uint64_t)num_sms * 8u ? (edge_cap + 255u) / 256u : (uint64_t)num_sms * 8u);
    if (ge == 0) ge = 1;
    int launches = 3;
    // acc is all zero here (the caller clears it): it counts in-degrees until the first nodes pass empties it again
    pr_degree_kernel<<<ge, 256, 0, st>>>(edges, table, a, state);
    pr_mark_hot_kernel<<<ge, 256, 0, st>>>(edges, a, state);
    pr_nodes_kernel<<<gn, 256, 0, st>>>(table, slots, a, state, 0u, damping, iterations == 0u);
    for (uint32_t it = 0; it < iterations; ++it) {
        pr_scatter_kernel<<<ge, 256, 0, st>>>(edges, a, state, it);
        pr_nodes_kernel<<<gn, 256, 0, st>>>(table, slots, a, state, it + 1u, damping, it + 1u == iterations);
        launches += 2;
    }
    return launches;
}

}  // namespace locust
