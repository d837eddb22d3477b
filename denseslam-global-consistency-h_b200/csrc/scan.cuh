// scan.cuh -- single-pass ORDERED tile scan (decoupled look-back) over the hash-slot space.
//
// The reference compacts with a per-CTA shared-memory scan plus one atomicAdd per CTA
// (CUDA/ITMCUDAUtils.h:35-76), which makes the visible-list order and the free-list pops
// schedule-dependent.  The CPU engine pops free lists and appends visible slots in ascending
// slot order (CPU/ITMSceneReconstructionEngine_CPU.cpp:207-316); to be bit-identical with it we
// need exclusive prefix sums in slot order.  Each CTA takes a tile through an atomic ticket (so a
// tile's predecessors are always already running), publishes its aggregate, and a warp walks back
// over predecessor tiles 32 at a time.  Two 31-bit counters travel together in one 64-bit word
// with a 2-bit status, so a single relaxed load sees a consistent (status, a, b).
#pragma once
#include "common.cuh"

namespace dsb {

constexpr unsigned long long kScanAggregate = 1ull << 62;   // tile aggregate available
constexpr unsigned long long kScanInclusive = 2ull << 62;   // inclusive prefix available

__device__ __forceinline__ unsigned long long scan_pack(unsigned long long status, unsigned a, unsigned b)
{
    return status | ((unsigned long long)a << 31) | (unsigned long long)b;
}
__device__ __forceinline__ unsigned scan_a(unsigned long long s) { return (unsigned)((s >> 31) & 0x7fffffffull); }
__device__ __forceinline__ unsigned scan_b(unsigned long long s) { return (unsigned)(s & 0x7fffffffull); }

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Called by ALL threads of warp 0 of the CTA owning `tile`; (aggA, aggB) is the CTA aggregate.
// Returns the exclusive prefix of the tile in (exA, exB) on every lane of the warp.
__device__ __forceinline__ void scan_lookback(unsigned long long *state, int tile, unsigned aggA, unsigned aggB,
                                              unsigned &exA, unsigned &exB)
{
    const int lane = threadIdx.x & 31;
    if (tile == 0) {
        if (lane == 0) st_relaxed(&state[0], scan_pack(kScanInclusive, aggA, aggB));
        exA = 0; exB = 0;
        return;
    }
    if (lane == 0) st_relaxed(&state[tile], scan_pack(kScanAggregate, aggA, aggB));
    unsigned sumA = 0, sumB = 0;
    int base = tile - 1;            // lane l inspects tile base - l
    while (true) {
        int j = base - lane;
        unsigned long long s = 0;
        bool valid = j >= 0;
        if (valid) {
            do { s = ld_relaxed(&state[j]); } while ((s >> 62) == 0);
        } else {
            s = kScanInclusive;     // virtual tile -1: inclusive prefix 0
        }
        unsigned incl = __ballot_sync(0xffffffffu, (s >> 62) == 2);
        // lanes up to (and including) the first inclusive one contribute
        int first = incl ? (__ffs(incl) - 1) : 32;
        unsigned a = (lane <= first) ? scan_a(s) : 0, b = (lane <= first) ? scan_b(s) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        sumA += a; sumB += b;
        if (incl) break;
        base -= 32;
    }
    exA = sumA; exB = sumB;
    if (lane == 0) st_relaxed(&state[tile], scan_pack(kScanInclusive, exA + aggA, exB + aggB));
}

// CTA-wide exclusive scan of two per-thread counts (blockDim.x == 256). Returns exclusive prefix
// for the thread and the CTA totals. `smem` must hold 2*8 unsigned.
__device__ __forceinline__ void block_scan2(unsigned a, unsigned b, unsigned *smem, unsigned &exA, unsigned &exB,
                                            unsigned &totA, unsigned &totB)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned ia = a, ib = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned ta = __shfl_up_sync(0xffffffffu, ia, o), tb = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) { ia += ta; ib += tb; }
    }
    if (lane == 31) { smem[warp] = ia; smem[8 + warp] = ib; }
    __syncthreads();
    unsigned wa = 0, wb = 0, ta = 0, tb = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        unsigned va = smem[w], vb = smem[8 + w];
        if (w < warp) { wa += va; wb += vb; }
        ta += va; tb += vb;
    }
    exA = wa + ia - a; exB = wb + ib - b; totA = ta; totB = tb;
    __syncthreads();
}

} // namespace dsb
