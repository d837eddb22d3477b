// recon_alloc.cu -- AllocateSceneFromDepth for sm_100a: allocation requests, ordered allocation,
// ordered visible-list compaction.  Deterministic and bit-identical to the reference's SERIAL CPU
// engine (CPU/ITMSceneReconstructionEngine_CPU.cpp:137-342), which the reference's own CUDA
// kernels are not (racy last-writer-wins requests, atomicSub free-list pops, per-CTA atomicAdd
// compaction: CUDA/ITMSceneReconstructionEngine_CUDA.cu:1521-1754).
//
//   k_begin_frame      previous visible slots -> type 3; reset scan states
//   k_alloc_requests   one thread / depth pixel: walk the +-mu segment, probe the hash; hits mark
//                      visibility, misses race with atomicMax on an ORDER KEY (pixel, step) so the
//                      winner is the serial loop's last writer
//   k_allocate         ordered scan over slots: rank -> free-list position, write entries
//   k_visible_list     re-test type-3 slots against the frustum, ordered compaction of visible slots
#include "kernels.cuh"
#include "scan.cuh"
#include "visibility.cuh"

namespace dsb {

// ---------------------------------------------------------------- shared ray set-up
struct RaySeg { float px, py, pz, dx, dy, dz; int noSteps; };

__device__ __forceinline__ bool ray_segment(RaySeg &r, int x, int y, float d, const AllocParams &P)
{
    if (d <= 0 || (d - P.mu) < 0 || (d - P.mu) < P.vfMin || (d + P.mu) > P.vfMax) return false;
    float cz = d;
    float cx = cz * (((float)x - P.invProj[2]) * P.invProj[0]);
    float cy = cz * (((float)y - P.invProj[3]) * P.invProj[1]);
    float norm = sqrtf(cx * cx + cy * cy + cz * cz);
    Vec4 t;
    t.x = cx * (1.0f - P.mu / norm); t.y = cy * (1.0f - P.mu / norm); t.z = cz * (1.0f - P.mu / norm); t.w = 1.0f;
    Vec4 a = mul(P.invM, t);
    float sx = a.x * P.oneOverBlock, sy = a.y * P.oneOverBlock, sz = a.z * P.oneOverBlock;
    t.x = cx * (1.0f + P.mu / norm); t.y = cy * (1.0f + P.mu / norm); t.z = cz * (1.0f + P.mu / norm);
    Vec4 b = mul(P.invM, t);
    float ex = b.x * P.oneOverBlock, ey = b.y * P.oneOverBlock, ez = b.z * P.oneOverBlock;
    float dx = ex - sx, dy = ey - sy, dz = ez - sz;
    norm = sqrtf(dx * dx + dy * dy + dz * dz);
    int noSteps = (int)ceilf(2.0f * norm);
    float div = (float)(noSteps - 1);
    r.px = sx; r.py = sy; r.pz = sz;
    r.dx = dx / div; r.dy = dy / div; r.dz = dz / div;
    r.noSteps = noSteps;
    return true;
}

__global__ void __launch_bounds__(256) k_begin_frame(FrameState fs, int hostLastFreeBlock, int hostLastFreeExcess,
                                                     int hostNoVisible, int useHostCounters, const int *visibleIds,
                                                     uint8_t *visType, int nScanWords)
{
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    // every thread derives the previous count the same way (no intra-kernel dependency)
    const int nPrev = useHostCounters ? hostNoVisible : fs.counters->noVisible;
    for (int i = tid; i < nPrev; i += stride) visType[visibleIds[i]] = 3;
    for (int i = tid; i < nScanWords; i += stride) fs.scanState[i] = 0ull;
    if (tid == 0) {
        fs.tickets[0] = 0; fs.tickets[1] = 0;
    }
    __syncthreads();
    if (tid == 0) {
        // written last by the only thread that reads noVisible late; other threads read it above.
        // (all reads of counters->noVisible in this kernel happen before this store only within
        //  this thread; other CTAs may still be reading -> keep noVisible untouched here.)
        if (useHostCounters) { fs.counters->lastFreeBlock = hostLastFreeBlock; fs.counters->lastFreeExcess = hostLastFreeExcess; }
        fs.counters->noVisiblePrev = nPrev;
    }
}

__global__ void __launch_bounds__(256) k_alloc_requests(AllocParams P, const float *__restrict__ depth,
                                                        const dsb200_hash_entry *__restrict__ hash,
                                                        uint8_t *visType, unsigned *reqKey)
{
    const int loc = blockIdx.x * blockDim.x + threadIdx.x;
    if (loc >= P.W * P.H) return;
    const int y = loc / P.W, x = loc - y * P.W;
    RaySeg r;
    if (!ray_segment(r, x, y, __ldg(depth + loc), P)) return;
    float px = r.px, py = r.py, pz = r.pz;
    int lastH = -1;   // consecutive steps usually fall in the same block: skip duplicate probes
    short lbx = 0, lby = 0, lbz = 0; bool lastMiss = false;
    for (int i = 0; i < r.noSteps; i++) {
        short bx = (short)floorf(px), by = (short)floorf(py), bz = (short)floorf(pz);
        if (lastH >= 0 && bx == lbx && by == lby && bz == lbz) {
            // same block as the previous step: a hit re-writes the same value; a miss must still
            // raise the order key to this later step
            if (lastMiss) atomicMax(&reqKey[lastH], (unsigned)loc * 256u + (unsigned)i + 1u);
        } else {
            int h = hash_index(bx, by, bz, P.bucketMask);
            HashEntry e = load_entry(hash, h);
            bool found = false;
            if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= -1) { visType[h] = (e.ptr == -1) ? 2 : 1; found = true; }
            if (!found) {
                if (e.ptr >= -1) {
                    while (e.offset >= 1) {
                        h = P.bucketNum + e.offset - 1;
                        e = load_entry(hash, h);
                        if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= -1) { visType[h] = (e.ptr == -1) ? 2 : 1; found = true; break; }
                    }
                }
                if (!found) atomicMax(&reqKey[h], (unsigned)loc * 256u + (unsigned)i + 1u);
            }
            lastH = h; lbx = bx; lby = by; lbz = bz; lastMiss = !found;
        }
        px += r.dx; py += r.dy; pz += r.dz;
    }
}

// One CTA = one tile of 1024 consecutive slots (4 per thread, 16 B key loads).
__global__ void __launch_bounds__(256) k_allocate(AllocParams P, FrameState fs, const float *__restrict__ depth,
                                                  dsb200_hash_entry *hash, const int *__restrict__ allocList,
                                                  const int *__restrict__ excessList, uint8_t *visType,
                                                  unsigned *reqKey, int total, int doAllocate)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx[2];
    if (threadIdx.x == 0) sTile = atomicAdd(&fs.tickets[0], 1);
    __syncthreads();
    const int tile = sTile;
    const int slot0 = tile * 1024 + threadIdx.x * 4;
    unsigned key[4] = {0, 0, 0, 0};
    if (slot0 + 3 < total) {
        uint4 k = *reinterpret_cast<const uint4 *>(reqKey + slot0);
        key[0] = k.x; key[1] = k.y; key[2] = k.z; key[3] = k.w;
    } else {
        for (int j = 0; j < 4; j++) if (slot0 + j < total) key[j] = reqKey[slot0 + j];
    }
    unsigned na = 0, nb = 0; unsigned char type[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if (key[j]) {
            int h = slot0 + j;
            // the request's type is a function of the slot alone: free bucket head -> 1, else
            // (occupied head with no chain, or chain tail in the excess area) -> 2
            bool ordered = (h < P.bucketNum) && (load_entry_cg(hash, h).ptr < -1);
            type[j] = ordered ? 1 : 2;
            na++; nb += ordered ? 0 : 1;
        }
    }
    unsigned exA, exB, totA, totB;
    block_scan2(na, nb, sScan, exA, exB, totA, totB);
    if (threadIdx.x < 32) {
        unsigned ta, tb;
        scan_lookback(fs.scanState, tile, totA, totB, ta, tb);
        if (threadIdx.x == 0) { sEx[0] = ta; sEx[1] = tb; }
    }
    __syncthreads();
    const int lastFreeBlock0 = fs.counters->lastFreeBlock, lastFreeExcess0 = fs.counters->lastFreeExcess;
    unsigned ra = sEx[0] + exA, rb = sEx[1] + exB;
    if (na) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (!key[j]) continue;
            const int h = slot0 + j;
            if (type[j] == 1) visType[h] = 1;             // new ordered entry is visible (DA/...:391)
            if (doAllocate) {
                // recompute the winning request's block: pixel = key/256, step = key%256 - 1
                const unsigned ord = key[j] - 1u;
                const int loc = (int)(ord >> 8), step = (int)(ord & 255u);
                const int y = loc / P.W, x = loc - y * P.W;
                RaySeg r; ray_segment(r, x, y, __ldg(depth + loc), P);
                float px = r.px, py = r.py, pz = r.pz;
                for (int s = 0; s < step; s++) { px += r.dx; py += r.dy; pz += r.dz; }
                const short bx = (short)floorf(px), by = (short)floorf(py), bz = (short)floorf(pz);
                const int vbaIdx = lastFreeBlock0 - (int)ra;
                if (type[j] == 1) {
                    if (vbaIdx >= 0) store_entry(hash, h, bx, by, bz, 0, allocList[vbaIdx], 0);
                } else {
                    const int exlIdx = lastFreeExcess0 - (int)rb;
                    if (vbaIdx >= 0 && exlIdx >= 0) {
                        const int exlOffset = excessList[exlIdx];
                        reinterpret_cast<int *>(hash)[(size_t)h * 5 + 2] = exlOffset + 1;       // parent.offset
                        store_entry(hash, P.bucketNum + exlOffset, bx, by, bz, 0, allocList[vbaIdx], 0);
                        visType[P.bucketNum + exlOffset] = 1;
                    }
                    rb++;
                }
                ra++;
            }
        }
        // keys are consumed: leave the array zeroed for the next frame (no per-frame memset)
        if (slot0 + 3 < total) *reinterpret_cast<uint4 *>(reqKey + slot0) = make_uint4(0, 0, 0, 0);
        else for (int j = 0; j < 4; j++) if (slot0 + j < total) reqKey[slot0 + j] = 0;
    }
    // the tile holding the last slot knows the grand totals
    // (written to separate fields: other tiles may still be reading lastFreeBlock / lastFreeExcess)
    if (threadIdx.x == 0 && tile == gridDim.x - 1) {
        fs.counters->lastFreeBlockNew = doAllocate ? lastFreeBlock0 - (int)(sEx[0] + totA) : lastFreeBlock0;
        fs.counters->lastFreeExcessNew = doAllocate ? lastFreeExcess0 - (int)(sEx[1] + totB) : lastFreeExcess0;
    }
}

// Ordered visible-list build (CPU .cpp:261-336).  Counter a = visible slots, counter b = visible
// slots that are swapped out (ptr == -1) and get a VBA block back when useSwapping is on.
__global__ void __launch_bounds__(256) k_visible_list(VisParams P, FrameState fs, dsb200_hash_entry *hash,
                                                      const int *__restrict__ allocList, uint8_t *visType,
                                                      int *visibleIds, int visibleCapacity, int *snapshotIds,
                                                      int *snapshotCount, int total)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx[2];
    if (threadIdx.x == 0) sTile = atomicAdd(&fs.tickets[1], 1);
    __syncthreads();
    const int tile = sTile;
    const int slot0 = tile * 1024 + threadIdx.x * 4;
    unsigned char vt[4] = {0, 0, 0, 0};
    if (slot0 + 3 < total) {
        unsigned w = *reinterpret_cast<const unsigned *>(visType + slot0);
        vt[0] = w & 0xff; vt[1] = (w >> 8) & 0xff; vt[2] = (w >> 16) & 0xff; vt[3] = w >> 24;
    } else {
        for (int j = 0; j < 4; j++) if (slot0 + j < total) vt[j] = visType[slot0 + j];
    }
    unsigned na = 0, nb = 0; bool swapIn[4] = {false, false, false, false};
    if (vt[0] | vt[1] | vt[2] | vt[3]) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (vt[j] == 3) {
                HashEntry e = load_entry_cg(hash, slot0 + j);
                bool vis, visEnl;
                block_visibility(vis, visEnl, e.px, e.py, e.pz, P.M, P.proj, P.voxelSize, P.W, P.H, P.useSwapping != 0);
                if (!(P.useSwapping ? visEnl : vis)) { vt[j] = 0; visType[slot0 + j] = 0; }
            }
            if (vt[j] > 0) {
                na++;
                if (P.useSwapping && load_entry_cg(hash, slot0 + j).ptr == -1) { swapIn[j] = true; nb++; }
            }
        }
    }
    unsigned exA, exB, totA, totB;
    block_scan2(na, nb, sScan, exA, exB, totA, totB);
    if (threadIdx.x < 32) {
        unsigned ta, tb;
        scan_lookback(fs.scanState + fs.scanTiles, tile, totA, totB, ta, tb);
        if (threadIdx.x == 0) { sEx[0] = ta; sEx[1] = tb; }
    }
    __syncthreads();
    if (na) {
        unsigned ra = sEx[0] + exA, rb = sEx[1] + exB;
        const int lastFreeBlock0 = fs.counters->lastFreeBlockNew;   // result of k_allocate (previous launch)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (vt[j] == 0) continue;
            if ((int)ra < visibleCapacity) {
                visibleIds[ra] = slot0 + j;
                if (snapshotIds) snapshotIds[ra] = slot0 + j;
            }
            ra++;
            if (swapIn[j]) {
                int vbaIdx = lastFreeBlock0 - (int)rb;
                if (vbaIdx >= 0) reinterpret_cast<int *>(hash)[(size_t)(slot0 + j) * 5 + 3] = allocList[vbaIdx];
                rb++;
            }
        }
    }
    if (threadIdx.x == 0 && tile == gridDim.x - 1) {
        const int n = (int)(sEx[0] + totA);
        fs.counters->noVisible = n;
        if (snapshotCount) *snapshotCount = n < visibleCapacity ? n : visibleCapacity;
        // commit the frame's free-list tops (no tile of this kernel reads these two fields)
        fs.counters->lastFreeBlock = fs.counters->lastFreeBlockNew - (int)(sEx[1] + totB);
        fs.counters->lastFreeExcess = fs.counters->lastFreeExcessNew;
    }
}

// ---------------------------------------------------------------- host launchers
void launch_begin_frame(cudaStream_t st, const FrameState &fs, int hostLastFreeBlock, int hostLastFreeExcess,
                        int hostNoVisible, int useHost, const int *visibleIds, uint8_t *visType, int prevUpperBound)
{
    int work = prevUpperBound > 2 * fs.scanTiles ? prevUpperBound : 2 * fs.scanTiles;
    int grid = div_up(work > 1 ? work : 1, 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    DSB_LAUNCH(k_begin_frame, grid, 256, 0, st, fs, hostLastFreeBlock, hostLastFreeExcess, hostNoVisible, useHost,
               visibleIds, visType, 2 * fs.scanTiles);
}

void launch_alloc_requests(cudaStream_t st, const AllocParams &P, const float *depth, const dsb200_hash_entry *hash,
                           uint8_t *visType, unsigned *reqKey)
{
    DSB_LAUNCH(k_alloc_requests, div_up(P.W * P.H, 256), 256, 0, st, P, depth, hash, visType, reqKey);
}

void launch_allocate(cudaStream_t st, const AllocParams &P, const FrameState &fs, const float *depth,
                     dsb200_hash_entry *hash, const int *allocList, const int *excessList, uint8_t *visType,
                     unsigned *reqKey, int total, int doAllocate)
{
    DSB_LAUNCH(k_allocate, fs.scanTiles, 256, 0, st, P, fs, depth, hash, allocList, excessList, visType, reqKey, total,
               doAllocate);
}

void launch_visible_list(cudaStream_t st, const VisParams &P, const FrameState &fs, dsb200_hash_entry *hash,
                         const int *allocList, uint8_t *visType, int *visibleIds, int visibleCapacity, int *snapshotIds,
                         int *snapshotCount, int total)
{
    DSB_LAUNCH(k_visible_list, fs.scanTiles, 256, 0, st, P, fs, hash, allocList, visType, visibleIds, visibleCapacity,
               snapshotIds, snapshotCount, total);
}

} // namespace dsb
