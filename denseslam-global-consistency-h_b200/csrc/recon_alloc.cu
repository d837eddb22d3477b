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
constexpr unsigned kVerified = 0x80u;   // transient marker on a visibility flag: "type 3, frustum test passed this frame"

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
    for (int i = tid; i < nPrev; i += stride) {
        const int h = visibleIds[i];
        visType[h] = 3;
        atomicOr(&fs.visBits[h >> 5], 1u << (h & 31));   // normally already set by the previous build; makes any caller-supplied list safe
    }
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

// One thread per depth pixel.  A hit writes the visibility flag (and the bitmap bit when the flag was 0: the flag read is
// issued together with the hash-entry load, so it adds no latency); a miss races for the slot with an ORDER KEY and the
// first requester records the request (and whether it is an excess-list request: a function of the slot alone).
__global__ void __launch_bounds__(256) k_alloc_requests(AllocParams P, FrameState fs, const float *__restrict__ depth,
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
        if (P.shardCount > 1 && !shard_holds(bx, by, bz, P.shardRank, P.shardCount, P.shardBrickLog2, P.shardHalo)) {
            // another shard's block: no probe, no request (and it breaks the same-block run, as in the serial loop
            // where the next owned step is simply processed on its own)
            lastH = -1;
        } else if (lastH >= 0 && bx == lbx && by == lby && bz == lbz) {
            // same block as the previous step: a hit re-writes the same value; a miss must still
            // raise the order key to this later step (its request bit was set by the first miss)
            if (lastMiss) atomicMax(&reqKey[lastH], (unsigned)loc * 256u + (unsigned)i + 1u);
        } else {
            int h = hash_index(bx, by, bz, P.bucketMask);
            unsigned char oldVis = visType[h];
            HashEntry e = load_entry(hash, h);
            const bool headOccupied = e.ptr >= -1;
            bool found = e.px == bx && e.py == by && e.pz == bz && headOccupied;
            if (!found && headOccupied) {
                while (e.offset >= 1) {
                    h = P.bucketNum + e.offset - 1;
                    oldVis = visType[h];
                    e = load_entry(hash, h);
                    if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= -1) { found = true; break; }
                }
            }
            if (found) {
                const unsigned char v = (e.ptr == -1) ? 2 : 1;
                if (oldVis != v) visType[h] = v;
                if (oldVis == 0) atomicOr(&fs.visBits[h >> 5], 1u << (h & 31));   // a stale 0 only costs a redundant RED
            } else if (atomicMax(&reqKey[h], (unsigned)loc * 256u + (unsigned)i + 1u) == 0u) {
                atomicOr(&fs.reqBits[h >> 5], 1u << (h & 31));                    // first request for this slot
                if (headOccupied) atomicOr(&fs.reqExBits[h >> 5], 1u << (h & 31));
            }
            lastH = h; lbx = bx; lby = by; lbz = bz; lastMiss = !found;
        }
        px += r.dx; py += r.dy; pz += r.dz;
    }
}

// The ordered passes below scan BITMAPS (1 bit per hash slot) instead of the slot arrays: reqBits / reqExBits mark the slots
// that received an (excess-list) allocation request this frame; visBits is a superset of the slots whose visibility flag is
// non-zero.  One CTA = one tile of 1024 bitmap words = 32768 consecutive slots (4 words per thread, one 16 B load), so the
// 5 M-slot table of the bench configuration is 160 co-resident tiles whose look-back chain resolves in a few round trips,
// and the per-thread first pass is pure popcount (k_allocate) or at most eight independent 16 B flag loads (k_visible_list):
// no thread walks a serial chain of dependent loads before the scan.

// iterate over the set bits of the thread's 4 words in ascending slot order
#define DSB_FOR_EACH_BIT(words, wordBase, slotVar, jVar, bitVar)                                  \
    for (int jVar = 0; jVar < 4; jVar++)                                                          \
        for (unsigned rem_ = (words)[jVar]; rem_;)                                                \
            for (int bitVar = __ffs(rem_) - 1, slotVar = (((wordBase) + jVar) << 5) + bitVar, once_ = 1; once_; once_ = 0, rem_ &= rem_ - 1)

__global__ void __launch_bounds__(256) k_allocate(AllocParams P, FrameState fs, const float *__restrict__ depth,
                                                  dsb200_hash_entry *hash, const int *__restrict__ allocList,
                                                  const int *__restrict__ excessList, uint8_t *visType,
                                                  unsigned *reqKey, int total, int doAllocate, DeepList deep)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx[2];
    if (threadIdx.x == 0) sTile = atomicAdd(&fs.tickets[0], 1);
    __syncthreads();
    const int tile = sTile;
    const int word0 = tile * 1024 + threadIdx.x * 4;
    unsigned bits[4], excess[4];
    {
        const uint4 w = *reinterpret_cast<const uint4 *>(fs.reqBits + word0);   // padded to whole tiles
        const uint4 x = *reinterpret_cast<const uint4 *>(fs.reqExBits + word0);
        bits[0] = w.x; bits[1] = w.y; bits[2] = w.z; bits[3] = w.w;
        excess[0] = x.x; excess[1] = x.y; excess[2] = x.z; excess[3] = x.w;
    }
    const unsigned na = __popc(bits[0]) + __popc(bits[1]) + __popc(bits[2]) + __popc(bits[3]);
    const unsigned nb = __popc(excess[0]) + __popc(excess[1]) + __popc(excess[2]) + __popc(excess[3]);
    unsigned exA, exB, totA, totB;
    block_scan2(na, nb, sScan, exA, exB, totA, totB);
    if (threadIdx.x < 32) {
        unsigned ta, tb;
        scan_lookback(fs.scanState, tile, totA, totB, ta, tb);
        if (threadIdx.x == 0) { sEx[0] = ta; sEx[1] = tb; }
    }
    __syncthreads();
    const int lastFreeBlock0 = fs.counters->lastFreeBlock, lastFreeExcess0 = fs.counters->lastFreeExcess;
    if (na) {
        unsigned ra = sEx[0] + exA, rb = sEx[1] + exB;
        DSB_FOR_EACH_BIT(bits, word0, h, j, bit) {
            const bool isExcess = (excess[j] >> bit) & 1u;
            const unsigned key = reqKey[h];
            reqKey[h] = 0;                                    // keys are consumed: the array stays zeroed between frames
            if (!isExcess) {                                  // new ordered entry is visible (DA/...:391)
                visType[h] = 1;
                atomicOr(&fs.visBits[h >> 5], 1u << (h & 31));
            }
            if (doAllocate) {
                // recompute the winning request's block: pixel = key/256, step = key%256 - 1
                const unsigned ord = key - 1u;
                const int loc = (int)(ord >> 8), step = (int)(ord & 255u);
                const int y = loc / P.W, x = loc - y * P.W;
                RaySeg r; ray_segment(r, x, y, __ldg(depth + loc), P);
                float px = r.px, py = r.py, pz = r.pz;
                for (int s = 0; s < step; s++) { px += r.dx; py += r.dy; pz += r.dz; }
                const short bx = (short)floorf(px), by = (short)floorf(py), bz = (short)floorf(pz);
                const int vbaIdx = lastFreeBlock0 - (int)ra;
                int written = -1, parent = -1;                // deep look-ahead: what this rank wrote
                if (!isExcess) {
                    if (vbaIdx >= 0) written = h;
                } else {
                    const int exlIdx = lastFreeExcess0 - (int)rb;
                    if (vbaIdx >= 0 && exlIdx >= 0) { written = P.bucketNum + excessList[exlIdx]; parent = h; }
                    rb++;
                }
                if (written >= 0) {
                    int *pw = reinterpret_cast<int *>(hash) + (size_t)written * 5, *pp = reinterpret_cast<int *>(hash) + (size_t)h * 5 + 2;
                    int ptr = allocList[vbaIdx];
                    if (deep.slot && (int)ra < deep.capacity) {   // keep the bytes being overwritten (exact undo), publish the entry as PENDING (ra <= lastFreeBlock0 < capacity)
                        int *o = deep.old + (size_t)ra * 6;
                        o[0] = pw[0]; o[1] = pw[1]; o[2] = pw[2]; o[3] = pw[3]; o[4] = pw[4]; o[5] = parent >= 0 ? *pp : 0;
                        ptr = kPendingBase - ptr;
                    }
                    if (parent >= 0) {
                        *pp = written - P.bucketNum + 1;                                         // parent.offset
                        visType[written] = 1;
                        atomicOr(&fs.visBits[written >> 5], 1u << (written & 31));
                    }
                    store_entry(hash, written, bx, by, bz, 0, ptr, 0);
                }
                if (deep.slot && (int)ra < deep.capacity) { deep.slot[ra] = written; deep.parent[ra] = parent; }
                ra++;
            }
        }
        *reinterpret_cast<uint4 *>(fs.reqBits + word0) = make_uint4(0, 0, 0, 0);
        if (nb) *reinterpret_cast<uint4 *>(fs.reqExBits + word0) = make_uint4(0, 0, 0, 0);
    }
    // the tile holding the last slot knows the grand totals
    // (written to separate fields: other tiles may still be reading lastFreeBlock / lastFreeExcess)
    if (threadIdx.x == 0 && tile == gridDim.x - 1) {
        fs.counters->lastFreeBlockNew = doAllocate ? lastFreeBlock0 - (int)(sEx[0] + totA) : lastFreeBlock0;
        fs.counters->lastFreeExcessNew = doAllocate ? lastFreeExcess0 - (int)(sEx[1] + totB) : lastFreeExcess0;
        fs.counters->numNew = doAllocate ? (int)(sEx[0] + totA) : 0;
    }
    (void)total;
}

// Frustum re-test of the slots that were visible last frame and were not re-observed by this frame's depth rays (flag 3;
// CPU .cpp:268-285).  One thread per entry of the previous visible list = perfectly balanced, and the ordered compaction
// that follows only has to look at flags.  Runs BEFORE k_allocate: it touches flag-3 slots only, and a freed slot of the old list
// that this frame re-allocates gets its flag 1 from k_allocate afterwards, whatever the re-test of its stale position said
// (as in the serial engine, where the allocation pass precedes the visibility pass and flag 1 is never re-tested).
__global__ void __launch_bounds__(256) k_recheck_visibility(VisParams P, FrameState fs, const dsb200_hash_entry *__restrict__ hash,
                                                            const int *__restrict__ prevVisibleIds, uint8_t *visType)
{
    const int nPrev = fs.counters->noVisiblePrev;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nPrev; i += gridDim.x * blockDim.x) {
        const int h = __ldg(prevVisibleIds + i);
        if (visType[h] != 3) continue;
        const HashEntry e = load_entry(hash, h);
        bool vis, visEnl;
        block_visibility(vis, visEnl, e.px, e.py, e.pz, P.M, P.proj, P.voxelSize, P.W, P.H, P.useSwapping != 0);
        // verified entries carry a marker bit until k_visible_list emits them; a flag 3 WITHOUT the marker can only come from
        // outside the list (pruning moves flags between slots, .cu:574-579) and is re-tested there
        visType[h] = (P.useSwapping ? visEnl : vis) ? (uint8_t)(3 | kVerified) : (uint8_t)0;
    }
}

// non-zero bytes of the 32 visibility flags of bitmap word w -> mask
__device__ __forceinline__ unsigned flag_mask32(const uint8_t *visType, int w, int total)
{
    const int s0 = w << 5;
    unsigned m = 0;
    if (s0 + 31 < total) {
        const uint4 a = *reinterpret_cast<const uint4 *>(visType + s0), b = *reinterpret_cast<const uint4 *>(visType + s0 + 16);
        const unsigned v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            // byte != 0 -> bit: fold each byte onto its low bit
            unsigned t = v[k]; t |= t >> 4; t |= t >> 2; t |= t >> 1; t &= 0x01010101u;
            m |= ((t & 1u) | ((t >> 7) & 2u) | ((t >> 14) & 4u) | ((t >> 21) & 8u)) << (4 * k);
        }
    } else {
        for (int k = 0; k < 32; k++) if (s0 + k < total && visType[s0 + k]) m |= 1u << k;
    }
    return m;
}

// Ordered visible-list build (CPU .cpp:261-336).  Counter a = visible slots, counter b = visible
// slots that are swapped out (ptr == -1) and get a VBA block back when useSwapping is on.
__global__ void __launch_bounds__(256) k_visible_list(VisParams P, FrameState fs, dsb200_hash_entry *hash,
                                                      const int *__restrict__ allocList, uint8_t *visType,
                                                      int *visibleIds, int visibleCapacity, int *snapshotIds,
                                                      int *snapshotCount, int snapshotCapacity, int total, int defer)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx[2];
    if (threadIdx.x == 0) sTile = atomicAdd(&fs.tickets[1], 1);
    __syncthreads();
    const int tile = sTile;
    const int word0 = tile * 1024 + threadIdx.x * 4;
    unsigned bits[4], keep[4], swapIn[4] = {0, 0, 0, 0};
    {
        const uint4 w = *reinterpret_cast<const uint4 *>(fs.visBits + word0);
        bits[0] = w.x; bits[1] = w.y; bits[2] = w.z; bits[3] = w.w;
    }
    // the bitmap is a superset: confirm against the flags (independent vector loads, only for non-empty words)
#pragma unroll
    for (int j = 0; j < 4; j++) {
        keep[j] = 0;
        if (!bits[j]) continue;
        const int s0 = (word0 + j) << 5;
        if (s0 + 31 < total) {
            uint4 *fp = reinterpret_cast<uint4 *>(visType + s0);
            uint4 a = fp[0], b = fp[1];
            unsigned v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
            unsigned nz = 0, marked = 0, raw3 = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                unsigned t = v[k]; t |= t >> 4; t |= t >> 2; t |= t >> 1; t &= 0x01010101u;
                nz |= ((t & 1u) | ((t >> 7) & 2u) | ((t >> 14) & 4u) | ((t >> 21) & 8u)) << (4 * k);
                unsigned m = (v[k] >> 7) & 0x01010101u;
                marked |= ((m & 1u) | ((m >> 7) & 2u) | ((m >> 14) & 4u) | ((m >> 21) & 8u)) << (4 * k);
                unsigned x = v[k] ^ 0x03030303u; x |= x >> 4; x |= x >> 2; x |= x >> 1; x = ~x & 0x01010101u;   // byte == 3
                raw3 |= ((x & 1u) | ((x >> 7) & 2u) | ((x >> 14) & 4u) | ((x >> 21) & 8u)) << (4 * k);
            }
            keep[j] = bits[j] & nz;
            if (marked) {                                   // strip the markers: the stored flag is the reference's 3
                a.x &= 0x7f7f7f7fu; a.y &= 0x7f7f7f7fu; a.z &= 0x7f7f7f7fu; a.w &= 0x7f7f7f7fu;
                b.x &= 0x7f7f7f7fu; b.y &= 0x7f7f7f7fu; b.z &= 0x7f7f7f7fu; b.w &= 0x7f7f7f7fu;
                fp[0] = a; fp[1] = b;
            }
            for (unsigned rem = keep[j] & raw3; rem; rem &= rem - 1) {       // rare: flag 3 that k_recheck_visibility did not see
                const int bit = __ffs(rem) - 1, h = s0 + bit;
                const HashEntry e = load_entry_cg(hash, h);
                bool vis, visEnl;
                block_visibility(vis, visEnl, e.px, e.py, e.pz, P.M, P.proj, P.voxelSize, P.W, P.H, P.useSwapping != 0);
                if (!(P.useSwapping ? visEnl : vis)) { visType[h] = 0; keep[j] &= ~(1u << bit); }
            }
        } else {
            for (int k = 0; k < 32 && s0 + k < total; k++) {
                if (!((bits[j] >> k) & 1u)) continue;
                unsigned char f = visType[s0 + k];
                if (f == 3) {
                    const HashEntry e = load_entry_cg(hash, s0 + k);
                    bool vis, visEnl;
                    block_visibility(vis, visEnl, e.px, e.py, e.pz, P.M, P.proj, P.voxelSize, P.W, P.H, P.useSwapping != 0);
                    if (!(P.useSwapping ? visEnl : vis)) f = 0;
                }
                if (f & kVerified) f &= 0x7f;
                visType[s0 + k] = f;
                if (f) keep[j] |= 1u << k;
            }
        }
    }
    if (bits[0] != keep[0] || bits[1] != keep[1] || bits[2] != keep[2] || bits[3] != keep[3])
        *reinterpret_cast<uint4 *>(fs.visBits + word0) = make_uint4(keep[0], keep[1], keep[2], keep[3]);
    unsigned na = __popc(keep[0]) + __popc(keep[1]) + __popc(keep[2]) + __popc(keep[3]), nb = 0;
    if (P.useSwapping && na) {
        DSB_FOR_EACH_BIT(keep, word0, h, j, bit) {
            if (load_entry_cg(hash, h).ptr == -1) { swapIn[j] |= 1u << bit; nb++; }
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
        DSB_FOR_EACH_BIT(keep, word0, h, j, bit) {
            if ((int)ra < visibleCapacity) {
                visibleIds[ra] = h;
                if (snapshotIds && (int)ra < snapshotCapacity) snapshotIds[ra] = h;
            }
            ra++;
            if ((swapIn[j] >> bit) & 1u) {
                int vbaIdx = lastFreeBlock0 - (int)rb;
                if (vbaIdx >= 0) reinterpret_cast<int *>(hash)[(size_t)h * 5 + 3] = allocList[vbaIdx];
                rb++;
            }
        }
    }
    if (threadIdx.x == 0 && tile == gridDim.x - 1) {
        const int n = (int)(sEx[0] + totA);
        if (defer) { fs.counters->noVisibleNext = n; return; }     // deep look-ahead: k_commit_allocation publishes the frame
        fs.counters->noVisible = n;
        if (snapshotCount) {
            const int m = n < visibleCapacity ? n : visibleCapacity;
            *snapshotCount = m < snapshotCapacity ? m : snapshotCapacity;
            if (m > snapshotCapacity) fs.counters->snapshotOverflow = 1;
        }
        // commit the frame's free-list tops (no tile of this kernel reads these two fields)
        fs.counters->lastFreeBlock = fs.counters->lastFreeBlockNew - (int)(sEx[1] + totB);
        fs.counters->lastFreeExcess = fs.counters->lastFreeExcessNew;
    }
}

// visBits <- (visType != 0): run when an engine meets a render state it has not been tracking
__global__ void __launch_bounds__(256) k_rebuild_visbits(const uint8_t *__restrict__ visType, unsigned *visBits, int total, int words)
{
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < words; w += gridDim.x * blockDim.x)
        visBits[w] = ((w << 5) < total) ? flag_mask32(visType, w, total) : 0u;
}

// Undo of k_begin_frame + k_alloc_requests for a look-ahead that is being ABANDONED (the scene was modified behind the
// engine's back, or the next allocation came with another view / pose): request keys and request bitmaps back to all-zero,
// visibility flags back to "non-zero exactly for the members of the current visible list" (1, or 2 for a swapped-out entry;
// the 1 / 3 distinction of the last build is not recoverable and no consumer depends on it: the next allocation turns every
// member into 3 first, CPU .cpp:185-188).
__global__ void __launch_bounds__(256) k_abandon_requests_clear(FrameState fs, uint8_t *visType, unsigned *reqKey, int total)
{
    const int words = fs.bitTiles * 1024;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < words; w += gridDim.x * blockDim.x) {
        const unsigned rq = fs.reqBits[w];
        for (unsigned rem = rq; rem; rem &= rem - 1) { const int h = (w << 5) + __ffs(rem) - 1; if (h < total) reqKey[h] = 0u; }
        if (rq) { fs.reqBits[w] = 0u; fs.reqExBits[w] = 0u; }
        const unsigned vb = fs.visBits[w];
        for (unsigned rem = vb; rem; rem &= rem - 1) { const int h = (w << 5) + __ffs(rem) - 1; if (h < total) visType[h] = 0; }
        if (vb) fs.visBits[w] = 0u;
    }
}
__global__ void __launch_bounds__(256) k_abandon_requests_restore(FrameState fs, const dsb200_hash_entry *__restrict__ hash,
                                                                  const int *__restrict__ visibleIds, uint8_t *visType)
{
    const int n = fs.counters->noVisible;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int h = visibleIds[i];
        visType[h] = (load_entry(hash, h).ptr == -1) ? 2 : 1;
        atomicOr(&fs.visBits[h >> 5], 1u << (h & 31));
    }
}

// DEEP LOOK-AHEAD.  With the request pass, the ordered allocation and the visible-list build of frame k+1 can run beside frame
// k's raycast too, provided the raycast cannot see them: k_allocate (deep) publishes a new entry with ptr = kPendingBase - block
// (<= -3: every reader of the table tests ptr >= 0 or ptr >= -1, so a pending entry reads as "no block here", which is the state
// frame k's raycast must see; a torn read mixes two negative ptr values), keeps the 20 bytes it overwrote and the parent's old
// offset, and k_visible_list (defer) builds the list into the engine's second buffer and leaves every counter of frame k alone.
// The consuming allocation call then runs ONE small kernel: pending ptr -> ptr, list -> the render state (+ history snapshot),
// counters committed.  An abandoned look-ahead restores the saved bytes: the table is bit for bit what it was.
__global__ void __launch_bounds__(256) k_commit_allocation(FrameState fs, dsb200_hash_entry *hash, DeepList deep, const int *__restrict__ nextIds,
                                                           int *visibleIds, int visibleCapacity, int *snapshotIds, int *snapshotCount,
                                                           int snapshotCapacity)
{
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    const int n = fs.counters->noVisibleNext;
    int m = fs.counters->numNew; if (m > deep.capacity) m = deep.capacity;
    for (int i = tid; i < m; i += stride) {
        const int sl = deep.slot[i];
        if (sl >= 0) { int *p = reinterpret_cast<int *>(hash) + (size_t)sl * 5 + 3; *p = kPendingBase - *p; }
    }
    const int nc = n < visibleCapacity ? n : visibleCapacity;
    for (int i = tid; i < nc; i += stride) {
        const int id = nextIds[i];
        visibleIds[i] = id;
        if (snapshotIds && i < snapshotCapacity) snapshotIds[i] = id;
    }
    if (tid == 0) {          // no thread of this kernel reads the fields written here
        fs.counters->noVisible = n;
        if (snapshotCount) {
            *snapshotCount = nc < snapshotCapacity ? nc : snapshotCapacity;
            if (nc > snapshotCapacity) fs.counters->snapshotOverflow = 1;
        }
        fs.counters->lastFreeBlock = fs.counters->lastFreeBlockNew;
        fs.counters->lastFreeExcess = fs.counters->lastFreeExcessNew;
    }
}

__global__ void __launch_bounds__(256) k_abandon_allocation(FrameState fs, dsb200_hash_entry *hash, DeepList deep)
{
    int m = fs.counters->numNew; if (m > deep.capacity) m = deep.capacity;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const int sl = deep.slot[i];
        if (sl < 0) continue;
        const int *o = deep.old + (size_t)i * 6;
        int *p = reinterpret_cast<int *>(hash) + (size_t)sl * 5;
        p[0] = o[0]; p[1] = o[1]; p[2] = o[2]; p[3] = o[3]; p[4] = o[4];
        const int par = deep.parent[i];
        if (par >= 0) reinterpret_cast<int *>(hash)[(size_t)par * 5 + 2] = o[5];
    }
}

// entries a deep look-ahead has published but not committed (checkpointing refuses to write them)
__global__ void __launch_bounds__(256) k_count_pending(const dsb200_hash_entry *__restrict__ hash, int total, int *count)
{
    int n = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
        n += __ldg(reinterpret_cast<const int *>(hash) + (size_t)i * 5 + 3) <= kPendingBase ? 1 : 0;
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(count, n);
}
void launch_count_pending(cudaStream_t st, const dsb200_hash_entry *hash, int total, int *count)
{
    DSB_LAUNCH(k_count_pending, kNumSMs * 4, 256, 0, st, hash, total, count);
}

// ---------------------------------------------------------------- host launchers
void launch_commit_allocation(cudaStream_t st, const FrameState &fs, dsb200_hash_entry *hash, const DeepList &deep, const int *nextIds,
                              int *visibleIds, int visibleCapacity, int *snapshotIds, int *snapshotCount, int snapshotCapacity)
{
    DSB_LAUNCH(k_commit_allocation, kNumSMs, 256, 0, st, fs, hash, deep, nextIds, visibleIds, visibleCapacity, snapshotIds, snapshotCount,
               snapshotCapacity);
}
void launch_abandon_allocation(cudaStream_t st, const FrameState &fs, dsb200_hash_entry *hash, const DeepList &deep)
{
    DSB_LAUNCH(k_abandon_allocation, kNumSMs, 256, 0, st, fs, hash, deep);
}

void launch_abandon_requests(cudaStream_t st, const FrameState &fs, const dsb200_hash_entry *hash, const int *visibleIds,
                             uint8_t *visType, unsigned *reqKey, int total)
{
    DSB_LAUNCH(k_abandon_requests_clear, kNumSMs * 4, 256, 0, st, fs, visType, reqKey, total);
    DSB_LAUNCH(k_abandon_requests_restore, kNumSMs * 2, 256, 0, st, fs, hash, visibleIds, visType);
}

void launch_begin_frame(cudaStream_t st, const FrameState &fs, int hostLastFreeBlock, int hostLastFreeExcess,
                        int hostNoVisible, int useHost, const int *visibleIds, uint8_t *visType, int prevUpperBound)
{
    int work = prevUpperBound > 2 * fs.scanTiles ? prevUpperBound : 2 * fs.scanTiles;
    int grid = div_up(work > 1 ? work : 1, 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    DSB_LAUNCH(k_begin_frame, grid, 256, 0, st, fs, hostLastFreeBlock, hostLastFreeExcess, hostNoVisible, useHost,
               visibleIds, visType, 2 * fs.scanTiles);
}

void launch_alloc_requests(cudaStream_t st, const AllocParams &P, const FrameState &fs, const float *depth,
                           const dsb200_hash_entry *hash, uint8_t *visType, unsigned *reqKey)
{
    DSB_LAUNCH(k_alloc_requests, div_up(P.W * P.H, 256), 256, 0, st, P, fs, depth, hash, visType, reqKey);
}

void launch_allocate(cudaStream_t st, const AllocParams &P, const FrameState &fs, const float *depth,
                     dsb200_hash_entry *hash, const int *allocList, const int *excessList, uint8_t *visType,
                     unsigned *reqKey, int total, int doAllocate, const DeepList *deep)
{
    DeepList d{}; if (deep) d = *deep;
    DSB_LAUNCH(k_allocate, fs.bitTiles, 256, 0, st, P, fs, depth, hash, allocList, excessList, visType, reqKey, total,
               doAllocate, d);
}

void launch_visible_list(cudaStream_t st, const VisParams &P, const FrameState &fs, dsb200_hash_entry *hash,
                         const int *allocList, uint8_t *visType, int *visibleIds, int visibleCapacity, int *snapshotIds,
                         int *snapshotCount, int snapshotCapacity, int total, int defer)
{
    DSB_LAUNCH(k_visible_list, fs.bitTiles, 256, 0, st, P, fs, hash, allocList, visType, visibleIds, visibleCapacity,
               snapshotIds, snapshotCount, snapshotCapacity, total, defer);
}

// must run after launch_alloc_requests and before launch_allocate / launch_visible_list: visibleIds still holds the PREVIOUS list
void launch_recheck_visibility(cudaStream_t st, const VisParams &P, const FrameState &fs, const dsb200_hash_entry *hash,
                               const int *prevVisibleIds, uint8_t *visType)
{
    DSB_LAUNCH(k_recheck_visibility, kNumSMs * 2, 256, 0, st, P, fs, hash, prevVisibleIds, visType);
}

void launch_rebuild_visbits(cudaStream_t st, const FrameState &fs, const uint8_t *visType, int total)
{
    DSB_LAUNCH(k_rebuild_visbits, kNumSMs * 4, 256, 0, st, visType, fs.visBits, total, fs.bitTiles * 1024);
}

} // namespace dsb
