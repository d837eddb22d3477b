// common.cuh -- shared device/host definitions for libdsb200 (sm_100a).
//
// Numerics contract: this library is compiled with --fmad=false and WITHOUT --use_fast_math, so
// every float expression is evaluated exactly as written (IEEE add/mul/div/sqrt, no contraction).
// The expressions below follow the operation order of the reference's device-agnostic math
// (SURVEY.md Appendix C), which makes hash / visibility / range-image results bit-identical to
// the reference CPU engine built with -ffp-contract=off, and in practice the voxel and raycast
// results too.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/dsb200.h"

#ifdef __FAST_MATH__
#error "libdsb200 must not be built with --use_fast_math (bit-exact allocation/visibility contract)"
#endif

namespace dsb {

constexpr int kBlock = DSB200_BLOCK_SIZE;     // 8
constexpr int kBlock3 = DSB200_BLOCK_SIZE3;   // 512
constexpr int kNumSMs = 148;                  // B200
constexpr int kMaxShards = 16;                // sharded scene: at most this many raycast results per composite

struct Mat4 { float m[16]; };                 // column-major, ORUtils/Matrix.h:7-11
struct Vec4 { float x, y, z, w; };
struct Vec3 { float x, y, z; };

// Matrix4 * Vector4 with the reference's summation order (ORUtils/Matrix.h:128-135)
__host__ __device__ __forceinline__ Vec4 mul(const Mat4 &M, Vec4 v)
{
    Vec4 r;
    r.x = M.m[0] * v.x + M.m[4] * v.y + M.m[8] * v.z + M.m[12] * v.w;
    r.y = M.m[1] * v.x + M.m[5] * v.y + M.m[9] * v.z + M.m[13] * v.w;
    r.z = M.m[2] * v.x + M.m[6] * v.y + M.m[10] * v.z + M.m[14] * v.w;
    r.w = M.m[3] * v.x + M.m[7] * v.y + M.m[11] * v.z + M.m[15] * v.w;
    return r;
}

// hashIndex (DeviceAgnostic/ITMRepresentationAccess.h:10-12), mask generalised to bucket_num-1
__host__ __device__ __forceinline__ int hash_index(int x, int y, int z, int bucketMask)
{
    return (int)((((unsigned)x * 73856093u) ^ ((unsigned)y * 19349669u) ^ ((unsigned)z * 83492791u)) & (unsigned)bucketMask);
}

// Shard ownership of a voxel block (include/dsb200.h, dsb200_shard): bricks of 2^brickLog2 blocks per axis
// (arithmetic shift = floor division), brick coordinate hashed with the table's primes and a murmur3 finaliser.
__host__ __device__ __forceinline__ int shard_owner(int bx, int by, int bz, int count, int brickLog2)
{
    unsigned h = ((unsigned)(bx >> brickLog2) * 73856093u) ^ ((unsigned)(by >> brickLog2) * 19349669u) ^
                 ((unsigned)(bz >> brickLog2) * 83492791u);
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return (int)(h % (unsigned)count);
}

// holds = owned, or (halo) one of the blocks b - {0,1}^3 is owned.  b - d lies in another brick only when b sits on the
// low face of its brick along every axis where d is 1, so interior blocks cost one owner evaluation.
__host__ __device__ __forceinline__ bool shard_holds(int bx, int by, int bz, int rank, int count, int brickLog2, int halo)
{
    if (shard_owner(bx, by, bz, count, brickLog2) == rank) return true;
    if (!halo) return false;
    const int mask = (1 << brickLog2) - 1;
    const int fx = (bx & mask) == 0, fy = (by & mask) == 0, fz = (bz & mask) == 0;
    for (int d = 1; d < 8; d++) {
        const int dx = d & 1, dy = (d >> 1) & 1, dz = d >> 2;
        if ((dx && !fx) || (dy && !fy) || (dz && !fz)) continue;
        if (shard_owner(bx - dx, by - dy, bz - dz, count, brickLog2) == rank) return true;
    }
    return false;
}

// 20-byte hash entry viewed as 5 ints: {pos.x|pos.y<<16, pos.z|pad<<16, offset, ptr, allocatedTime}
struct HashEntry {
    short px, py, pz, pad;
    int offset, ptr, allocatedTime;
};
static_assert(sizeof(HashEntry) == 20, "ITMHashEntry layout");
static_assert(sizeof(dsb200_hash_entry) == 20, "ITMHashEntry layout");
static_assert(sizeof(dsb200_voxel) == 8, "ITMVoxel_s_rgb layout");

__device__ __forceinline__ HashEntry load_entry(const dsb200_hash_entry *table, int idx)
{
    // 20 B records are 4-byte aligned: five 32-bit read-only loads
    const int *p = reinterpret_cast<const int *>(table) + (size_t)idx * 5;
    int a = __ldg(p), b = __ldg(p + 1);
    HashEntry e;
    e.px = (short)(a & 0xffff); e.py = (short)(a >> 16); e.pz = (short)(b & 0xffff); e.pad = 0;
    e.offset = __ldg(p + 2); e.ptr = __ldg(p + 3); e.allocatedTime = 0;
    return e;
}
// same, but through the coherent path (for kernels that also modify the table)
__device__ __forceinline__ HashEntry load_entry_cg(const dsb200_hash_entry *table, int idx)
{
    const int *p = reinterpret_cast<const int *>(table) + (size_t)idx * 5;
    int a = __ldcg(p), b = __ldcg(p + 1);
    HashEntry e;
    e.px = (short)(a & 0xffff); e.py = (short)(a >> 16); e.pz = (short)(b & 0xffff); e.pad = 0;
    e.offset = __ldcg(p + 2); e.ptr = __ldcg(p + 3); e.allocatedTime = 0;
    return e;
}
__device__ __forceinline__ void store_entry(dsb200_hash_entry *table, int idx, short x, short y, short z, int offset,
                                            int ptr, int allocatedTime)
{
    int *p = reinterpret_cast<int *>(table) + (size_t)idx * 5;
    p[0] = (int)((unsigned)(unsigned short)x | ((unsigned)(unsigned short)y << 16));
    p[1] = (int)(unsigned)(unsigned short)z;
    p[2] = offset; p[3] = ptr; p[4] = allocatedTime;
}

// Device-resident counters shared by the stages of one scene (authoritative while a frame is in
// flight; mirrored to dsb200_scene / dsb200_render_state on sync).
struct Counters {
    int lastFreeBlock;     // scene->localVBA.lastFreeBlockId
    int lastFreeExcess;    // scene->index.lastFreeExcessListId
    int noVisible;         // renderState->noVisibleEntries
    int noVisiblePrev;     // value at the start of AllocateSceneFromDepth
    int renderBlocks;      // rendering-block (16x16 tile) count of CreateExpectedDepths
    int freedBlocks;       // blocks released by the last pruning pass
    int lastFreeBlockNew;  // k_allocate result, committed by k_visible_list
    int lastFreeExcessNew;
    int totalFreed;        // blocks released by all pruning passes since ResetScene (GetDecayedBlockCount)
    int snapshotOverflow;  // a visible-list snapshot did not fit its asynchronous reservation (sticky until the next sync)
    int noVisibleNext;     // deep look-ahead: length of the NEXT frame's visible list (built beside this frame's raycast)
    int numNew;            // deep look-ahead: allocation requests ranked by k_allocate (entries of the undo / commit list)
};

extern unsigned long long g_launches;   // kernels launched by this library (bench.py gpu_launches)

#define DSB_LAUNCH(kernel, grid, block, smem, stream, ...)                    \
    do {                                                                       \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);            \
        ++dsb::g_launches;                                                     \
    } while (0)

inline int div_up(int a, int b) { return (a + b - 1) / b; }

} // namespace dsb
