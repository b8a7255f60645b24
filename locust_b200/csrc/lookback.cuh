// lookback.cuh -- decoupled look-back (chained scan) building blocks shared by the
// device-wide scan and the onesweep radix passes.
//
// A status word is 64 bits: [63:56] epoch, [55:54] flag, [53:0] value.  The epoch
// is a host-side counter that changes with every launch, so the status array
// never has to be cleared between launches (only when the 8-bit epoch wraps).
#pragma once
#include "common.cuh"

namespace locust {

constexpr uint32_t kLbAggregate = 1u;
constexpr uint32_t kLbInclusive = 2u;
constexpr uint64_t kLbValueMask = (1ull << 54) - 1ull;

LC_D uint64_t lb_pack(uint32_t epoch, uint32_t flag, uint64_t v) {
    return ((uint64_t)epoch << 56) | ((uint64_t)flag << 54) | (v & kLbValueMask);
}
LC_D void lb_store(uint64_t *p, uint64_t w) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
LC_D uint64_t lb_load(const uint64_t *p) {
    uint64_t w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}

// Sum of the values of tiles [0, tile) in column `col` of a status array laid out
// as state[tile * stride + col]; stops at the first inclusive word.  The words of up
// to eight predecessors are requested at once, so the walk costs about one memory
// round trip per eight tiles instead of one per tile.
LC_D uint64_t lb_lookback(const uint64_t *state, uint32_t tile, uint32_t stride, uint32_t col, uint32_t epoch) {
    uint64_t sum = 0;
    uint32_t t = tile;
    while (t > 0) {
        const uint32_t batch = t < 8u ? t : 8u;
        uint64_t w[8];
#pragma unroll
        for (uint32_t i = 0; i < 8; ++i)
            if (i < batch) w[i] = lb_load(state + (uint64_t)(t - 1u - i) * stride + col);
#pragma unroll
        for (uint32_t i = 0; i < 8; ++i) {
            if (i < batch) {
                uint64_t x = w[i];
                uint32_t spins = 0;
                while ((uint32_t)(x >> 56) != epoch || ((x >> 54) & 3ull) == 0ull) {
                    if (++spins > 16) __nanosleep(64);
                    x = lb_load(state + (uint64_t)(t - 1u - i) * stride + col);
                }
                sum += x & kLbValueMask;
                if (((x >> 54) & 3ull) == kLbInclusive) return sum;
            }
        }
        t -= batch;
    }
    return sum;
}

// The same walk done by a whole warp over a single column (stride 1): 32 predecessors per round,
// one per lane; the nearest inclusive word ends the walk.  All 32 lanes call this together and
// get the sum.  With tens of thousands of tiles in flight a serial walk costs more than the
// tile's own work.
LC_D uint64_t lb_lookback_warp(const uint64_t *state, uint32_t tile, uint32_t epoch, uint32_t lane) {
    uint64_t sum = 0;
    uint32_t t = tile;
    while (t > 0) {
        const bool valid = lane < t;
        uint64_t x = 0;
        if (valid) {
            const uint64_t *p = state + (t - 1u - lane);
            x = lb_load(p);
            uint32_t spins = 0;
            while ((uint32_t)(x >> 56) != epoch || ((x >> 54) & 3ull) == 0ull) {
                if (++spins > 16) __nanosleep(64);
                x = lb_load(p);
            }
        }
        const uint32_t incl = __ballot_sync(0xFFFFFFFFu, valid && ((x >> 54) & 3ull) == kLbInclusive);
        const uint32_t last = incl ? (uint32_t)__ffs((int)incl) - 1u : 31u;      // nearest inclusive predecessor
        uint64_t v = (valid && lane <= last) ? (x & kLbValueMask) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
        sum += v;
        if (incl) break;
        t -= t < 32u ? t : 32u;
    }
    return sum;
}

// Dynamic tile index: blocks take tiles in the order they start running, so a tile
// only ever waits on tiles that are already resident.  ticket[0] = next tile,
// ticket[1] = finished blocks; the last block to finish resets both.
LC_D uint32_t take_ticket(uint32_t *ticket, uint32_t *smem_slot) {
    if (threadIdx.x == 0) *smem_slot = atomicAdd(ticket, 1u);
    __syncthreads();
    return *smem_slot;
}
LC_D void return_ticket(uint32_t *ticket) {
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(ticket + 1, 1u) == gridDim.x - 1u) { ticket[0] = 0u; ticket[1] = 0u; __threadfence(); }
    }
}

}  // namespace locust
