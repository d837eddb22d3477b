// recon_prune.cu -- voxel decay and sliding-window pruning for sm_100a.
//
// Reference (CUDA only): decay_device / slidewindow_device -> decayVoxel / deleteVoxel ->
// deleteBlock, CUDA/ITMSceneReconstructionEngine_CUDA.cu:501-847.  One 512-thread CTA per
// snapshot entry; thread 0 of an emptied block unlinks it under a per-bucket try-lock and pushes
// the VBA block with atomicAdd -- the outcome (which deletions are skipped on lock contention, the
// free-list order) depends on CTA scheduling.
//
// Here the same work runs under ONE fixed, valid schedule (DESIGN.md "pruning schedule"):
//   phase A  k_prune_voxels   every listed block: reset voxels, test emptiness (table read-only)
//   phase B1 k_prune_rank     ordered scan over the list: rank of each emptied block -> its
//                             free-list position (list order), leader election per bucket
//   phase B2 k_prune_unlink   per bucket, ONE thread applies that bucket's deletions in list
//                             order (deleteBlock's three unlink cases), so no lock is needed and no
//                             deletion is ever skipped
//   phase B3 k_prune_cleanup  clears the slot marks
// Excess-list slots are not recycled (reference TODO .cu:560-561).
#include "kernels.cuh"
#include "scan.cuh"

namespace dsb {

// findVoxel(..., outHashIdx, outPrevHashIdx), DeviceAgnostic/ITMRepresentationAccess.h:87-131
__device__ __forceinline__ int find_block_prev(const dsb200_hash_entry *hash, int bucketNum, short bx, short by, short bz,
                                               int &prev)
{
    prev = -1;
    int h = hash_index(bx, by, bz, bucketNum - 1);
    while (true) {
        HashEntry e = load_entry_cg(hash, h);
        if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= 0) return h;
        if (e.offset < 1) break;
        prev = h;
        h = bucketNum + e.offset - 1;
    }
    return -1;
}

template <bool SLIDE>
__global__ void __launch_bounds__(256) k_prune_voxels(dsb200_voxel *voxels, const dsb200_hash_entry *hash, int bucketNum,
                                                      const int *__restrict__ ids, const int *__restrict__ countPtr,
                                                      int minAge, int maxWeight, int currentFrame, int *flags,
                                                      FrameState fs, int pruneTiles, unsigned *shadow2)
{
    const int n = *countPtr;
    const int t = threadIdx.x;
    // reset the look-back words / ticket used by k_prune_rank (launched after this kernel)
    for (int i = blockIdx.x * blockDim.x + t; i < pruneTiles; i += gridDim.x * blockDim.x) fs.scanState[2 * fs.scanTiles + i] = 0ull;
    if (blockIdx.x == 0 && t == 0) fs.tickets[2] = 0;
    for (int i = blockIdx.x; i < n; i += gridDim.x) {
        const int slot = __ldg(ids + i);
        int idx = -1;
        bool safe = true;
        if (slot >= 0) {
            HashEntry cur = load_entry_cg(hash, slot);
            if (cur.ptr >= 0) {
                int prev;
                idx = find_block_prev(hash, bucketNum, cur.px, cur.py, cur.pz, prev);
            }
        }
        if (idx < 0) { if (t == 0) flags[i] = 0; continue; }
        if (!SLIDE) {
            // age = currentFrame - allocatedTime; allocatedTime is never written by the reference
            // allocators and is defined as 0 here
            const int allocatedTime = __ldcg(reinterpret_cast<const int *>(hash) + (size_t)idx * 5 + 4);
            if (currentFrame - allocatedTime < minAge) safe = false;
        }
        int allEmpty = 0;
        if (safe) {
            const int ptr = __ldcg(reinterpret_cast<const int *>(hash) + (size_t)idx * 5 + 3);
            uint4 *p = reinterpret_cast<uint4 *>(voxels + (size_t)ptr * kBlock3) + t;
            const unsigned blankLo = 0x00007fffu;
            if (SLIDE) {
                // deleteVoxel: every voxel is reset (pad byte preserved); the block always empties
                uint4 raw = *p;
                uint4 out = make_uint4(blankLo, raw.y & 0xff000000u, blankLo, raw.w & 0xff000000u);
                if (out.x != raw.x || out.y != raw.y || out.z != raw.z || out.w != raw.w) {
                    *p = out;
                    if (shadow2) shadow2[(size_t)ptr * 256 + t] = 0x7fff7fffu;
                }
                allEmpty = 1;
            } else {
                uint4 raw = *p, out = raw;
                unsigned w0 = (raw.x >> 16) & 0xff, w1 = (raw.z >> 16) & 0xff;
                if (w0 > 0 && (int)w0 <= maxWeight) { out.x = blankLo; out.y = raw.y & 0xff000000u; w0 = 0; }
                if (w1 > 0 && (int)w1 <= maxWeight) { out.z = blankLo; out.w = raw.w & 0xff000000u; w1 = 0; }
                if (out.x != raw.x || out.y != raw.y || out.z != raw.z || out.w != raw.w) {
                    *p = out;
                    if (shadow2) shadow2[(size_t)ptr * 256 + t] = (out.x & 0xffffu) | (out.z << 16);
                }
                allEmpty = (w0 == 0 && w1 == 0);
            }
        }
        const int cnt = __syncthreads_count(allEmpty);
        if (t == 0) flags[i] = (safe && cnt == 256) ? 1 : 0;
    }
}

// scratch layout (ints): flags[cap] | compactSlot[cap] | compactBucket[cap]
__global__ void __launch_bounds__(256) k_prune_rank(const dsb200_hash_entry *hash, int bucketNum,
                                                    const int *__restrict__ ids, const int *__restrict__ countPtr,
                                                    const int *__restrict__ flags, int *compactSlot, int *compactBucket,
                                                    int *allocList, unsigned *slotMark, unsigned *bucketLeader,
                                                    FrameState fs)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx[2];
    const int n = *countPtr;
    const int nTiles = (n + 1023) / 1024;
    const int lastFree0 = fs.counters->lastFreeBlock;
    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) sTile = atomicAdd(&fs.tickets[2], 1);
        __syncthreads();
        const int tile = sTile;
        if (tile >= nTiles) break;
        const int i0 = tile * 1024 + threadIdx.x * 4;
        int f[4]; unsigned na = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) { f[j] = (i0 + j < n) ? flags[i0 + j] : 0; na += f[j]; }
        unsigned exA, exB, totA, totB;
        block_scan2(na, 0, sScan, exA, exB, totA, totB);
        if (threadIdx.x < 32) {
            unsigned ta, tb;
            scan_lookback(fs.scanState + 2 * fs.scanTiles, tile, totA, 0, ta, tb);
            if (threadIdx.x == 0) { sEx[0] = ta; }
        }
        __syncthreads();
        unsigned r = sEx[0] + exA;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (!f[j]) continue;
            const int slot = __ldg(ids + i0 + j);
            HashEntry e = load_entry_cg(hash, slot);
            const int b = hash_index(e.px, e.py, e.pz, bucketNum - 1);
            allocList[lastFree0 + 1 + (int)r] = e.ptr;       // deleteBlock .cu:558-559, in list order
            compactSlot[r] = slot; compactBucket[r] = b;
            slotMark[slot] = r + 1u;
            atomicMax(&bucketLeader[b], 0xffffffffu - r);     // smallest rank wins the bucket
            r++;
        }
        if (threadIdx.x == 0 && tile == nTiles - 1) {
            fs.counters->freedBlocks = (int)(sEx[0] + totA);
            fs.counters->totalFreed += (int)(sEx[0] + totA);
            fs.counters->lastFreeBlockNew = lastFree0 + (int)(sEx[0] + totA);
        }
    }
    if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0) { fs.counters->freedBlocks = 0; fs.counters->lastFreeBlockNew = lastFree0; }
}

__global__ void __launch_bounds__(128) k_prune_unlink(dsb200_hash_entry *hash, int bucketNum, const int *compactSlot,
                                                      const int *compactBucket, unsigned *slotMark,
                                                      unsigned *bucketLeader, uint8_t *visType, unsigned *visBits, FrameState fs)
{
    const int total = fs.counters->freedBlocks;
    int *hw = reinterpret_cast<int *>(hash);
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < total; r += gridDim.x * blockDim.x) {
        const int b = compactBucket[r];
        if (bucketLeader[b] != 0xffffffffu - (unsigned)r) continue;
        // this thread owns bucket b: apply its deletions in ascending rank (= list) order
        unsigned done = 0;   // marks <= done are processed
        while (true) {
            // find the chain member with the smallest mark > done
            int bestIdx = -1, bestPrev = -1; unsigned bestMark = 0xffffffffu;
            int h = b, prev = -1;
            while (true) {
                const unsigned m = slotMark[h];
                const int ptr = hw[(size_t)h * 5 + 3], off = hw[(size_t)h * 5 + 2];
                if (m > done && m < bestMark && ptr >= 0) { bestMark = m; bestIdx = h; bestPrev = prev; }
                if (off < 1) break;
                prev = h; h = bucketNum + off - 1;
            }
            if (bestIdx < 0) break;
            done = bestMark;
            const int idx = bestIdx;
            slotMark[idx] = 0;
            if (bestPrev == -1) {
                const int off = hw[(size_t)idx * 5 + 2];
                if (off >= 1) {        // bucket head with a successor: pull the successor up
                    const int next = bucketNum + off - 1;
#pragma unroll
                    for (int k = 0; k < 5; k++) hw[(size_t)idx * 5 + k] = hw[(size_t)next * 5 + k];
                    visType[idx] = visType[next]; visType[next] = 0;
                    if (visBits && visType[idx]) atomicOr(&visBits[idx >> 5], 1u << (idx & 31));
                    slotMark[idx] = slotMark[next]; slotMark[next] = 0;
                    hw[(size_t)next * 5 + 2] = 0; hw[(size_t)next * 5 + 3] = -2;
                } else {               // bucket head, no successor
                    hw[(size_t)idx * 5 + 3] = -2; visType[idx] = 0;
                }
            } else {                   // chain member: bypass
                hw[(size_t)bestPrev * 5 + 2] = hw[(size_t)idx * 5 + 2];
                hw[(size_t)idx * 5 + 2] = 0; hw[(size_t)idx * 5 + 3] = -2;
                visType[bestPrev] = visType[idx]; visType[idx] = 0;
                if (visBits && visType[bestPrev]) atomicOr(&visBits[bestPrev >> 5], 1u << (bestPrev & 31));
            }
        }
        bucketLeader[b] = 0;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) fs.counters->lastFreeBlock = fs.counters->lastFreeBlockNew;
}

__global__ void k_prune_cleanup(const int *compactSlot, unsigned *slotMark, FrameState fs)
{
    const int total = fs.counters->freedBlocks;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < total; r += gridDim.x * blockDim.x) slotMark[compactSlot[r]] = 0;
}

void launch_prune(cudaStream_t st, bool slide, dsb200_voxel *voxels, dsb200_hash_entry *hash, int bucketNum,
                  const int *ids, const int *count, int countUpperBound, int minAge, int maxWeight, int currentFrame,
                  int *allocList, uint8_t *visType, unsigned *visBits, unsigned *slotMark, int *scratch, Counters *counters,
                  const FrameState &fs, short *sdfShadow)
{
    unsigned *shadow2 = reinterpret_cast<unsigned *>(sdfShadow);
    (void)counters;
    const int cap = countUpperBound;
    int *flags = scratch, *compactSlot = scratch + cap, *compactBucket = scratch + 2 * cap;
    unsigned *bucketLeader = slotMark + (size_t)fs.scanTiles * 1024;   // second half of the mark array
    const int pruneTiles = div_up(cap, 1024);
    const int grid = kNumSMs * 8;
    if (slide)
        DSB_LAUNCH(k_prune_voxels<true>, grid, 256, 0, st, voxels, hash, bucketNum, ids, count, minAge, maxWeight,
                   currentFrame, flags, fs, pruneTiles, shadow2);
    else
        DSB_LAUNCH(k_prune_voxels<false>, grid, 256, 0, st, voxels, hash, bucketNum, ids, count, minAge, maxWeight,
                   currentFrame, flags, fs, pruneTiles, shadow2);
    int rankGrid = pruneTiles < kNumSMs * 4 ? pruneTiles : kNumSMs * 4;
    if (rankGrid < 1) rankGrid = 1;
    DSB_LAUNCH(k_prune_rank, rankGrid, 256, 0, st, hash, bucketNum, ids, count, flags, compactSlot, compactBucket,
               allocList, slotMark, bucketLeader, fs);
    DSB_LAUNCH(k_prune_unlink, kNumSMs * 2, 128, 0, st, hash, bucketNum, compactSlot, compactBucket, slotMark,
               bucketLeader, visType, visBits, fs);
    DSB_LAUNCH(k_prune_cleanup, kNumSMs, 256, 0, st, compactSlot, slotMark, fs);
}

} // namespace dsb
