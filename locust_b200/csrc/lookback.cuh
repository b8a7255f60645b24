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
// as state[tile * stride + col]; stops at the first inclusive word.
LC_D uint64_t lb_lookback(const uint64_t *state, uint32_t tile, uint32_t stride, uint32_t col, uint32_t epoch) {
    uint64_t sum = 0;
    for (uint32_t t = tile; t-- > 0;) {
        uint64_t w;
        uint32_t spins = 0;
        for (;;) {
            w = lb_load(state + (uint64_t)t * stride + col);
            if ((uint32_t)(w >> 56) == epoch && ((w >> 54) & 3ull) != 0ull) break;
            if (++spins > 16) __nanosleep(64);
        }
        sum += w & kLbValueMask;
        if (((w >> 54) & 3ull) == kLbInclusive) break;
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
