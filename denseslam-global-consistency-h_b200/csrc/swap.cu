// swap.cu -- host <-> device voxel-block swapping: ITMSwappingEngine on ITMGlobalCache.
//
//   reference                                                          here
//   ITMGlobalCache (Objects/ITMGlobalCache.h:18-129)                   dsb200_swap: device swap states, pinned host store
//   swap-state pass of AllocateSceneFromDepth (CUDA .cu:1730-1733)     k_swap_mark_visible
//   LoadFromGlobalMemory / IntegrateGlobalIntoLocal                    k_swap_count<0> -> k_swap_scan -> k_swap_emit<0>,
//     (CPU/ITMSwappingEngine_CPU.cpp:22-131, CUDA .cu:263-306)           k_swap_integrate reading the stored blocks in place
//   SaveToGlobalMemory, both overloads (CPU .cpp:133-297,              k_swap_count<1> -> k_swap_scan -> k_swap_emit<1>,
//     CUDA .cu:308-390)                                                  k_swap_save writing the blocks into the store
//
// The reference's CUDA engine compacts the candidate entries with an atomic per CTA, so which 4096 entries move when more
// are eligible depends on the schedule; its CPU engine takes the first 4096 in ascending entry order and pushes freed blocks on
// the free list in that order.  Here the candidates are ranked in entry order (per-tile counts, one scan over the tiles, ranked
// emit), which makes the result the CPU engine's, bit for bit.
//
// The host store.  The reference keeps total_entries x 4 KB of malloc'd memory indexed by entry and moves blocks through a pinned
// staging buffer: device -> staging -> memcpy into the store, whose pages are touched for the first time on that memcpy (measured
// here with that layout: 1.4 ms per 4096-block swap-in, 7.8 ms per swap-out call).  Here the store is a compact arena
// of PINNED 64 MB chunks with one slot index per entry; pinned memory is addressable from the device, so k_swap_save writes each
// block straight into its slot and k_swap_integrate reads it from there, 16 bytes per thread over PCIe: no staging buffer, no host
// copy, no page fault on the call path (a chunk is allocated once per 16384 new blocks).
#include "kernels.cuh"
#include "scan.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace dsb {
int fail_with(int code, const char *msg);

namespace {

constexpr int kSlotsPerThread = 8, kTileSlots = 256 * kSlotsPerThread;

// MODE 0: entries to load (state 1).  MODE 1: entries to save (state 2, block in the volume, not visible when flags are given).
template <int MODE>
__device__ __forceinline__ bool swap_candidate(int h, const uint8_t *__restrict__ states, const dsb200_hash_entry *__restrict__ hash,
                                               const uint8_t *__restrict__ visType)
{
    if (MODE == 0) return states[h] == 1;
    if (states[h] != 2) return false;
    if (reinterpret_cast<const int *>(hash)[(size_t)h * 5 + 3] < 0) return false;
    return visType == nullptr || visType[h] == 0;
}

__global__ void __launch_bounds__(256) k_swap_mark_visible(uint8_t *states, const uint8_t *__restrict__ visType, int total)
{
    for (int h = blockIdx.x * blockDim.x + threadIdx.x; h < total; h += gridDim.x * blockDim.x)
        if (visType[h] > 0 && states[h] != 2) states[h] = 1;
}

template <int MODE>
__global__ void __launch_bounds__(256) k_swap_count(const uint8_t *__restrict__ states, const dsb200_hash_entry *__restrict__ hash,
                                                    const uint8_t *__restrict__ visType, int total, int *__restrict__ tileCount)
{
    __shared__ unsigned sScan[16];
    const int h0 = blockIdx.x * kTileSlots + threadIdx.x * kSlotsPerThread;
    unsigned n = 0;
    for (int k = 0; k < kSlotsPerThread; k++)
        if (h0 + k < total && swap_candidate<MODE>(h0 + k, states, hash, visType)) n++;
    unsigned exA, exB, totA, totB;
    block_scan2(n, 0u, sScan, exA, exB, totA, totB);
    if (threadIdx.x == 0) tileCount[blockIdx.x] = (int)totA;
}

// exclusive scan of the tile counts by one CTA; result[0] = number of candidates
__global__ void __launch_bounds__(1024) k_swap_scan(const int *__restrict__ tileCount, int nTiles, int *__restrict__ tileOffset,
                                                    int *__restrict__ result)
{
    __shared__ int sWarp[32];
    __shared__ int sCarry;
    if (threadIdx.x == 0) sCarry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nTiles; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = (i < nTiles) ? tileCount[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) sWarp[warp] = inc;
        __syncthreads();
        int before = 0, all = 0;
        for (int w = 0; w < 32; w++) { const int s = sWarp[w]; if (w < warp) before += s; all += s; }
        const int carry = sCarry;
        if (i < nTiles) tileOffset[i] = carry + before + inc - v;
        __syncthreads();
        if (threadIdx.x == 0) sCarry = carry + all;
        __syncthreads();
    }
    if (threadIdx.x == 0) result[0] = sCarry;
}

template <int MODE>
__global__ void __launch_bounds__(256) k_swap_emit(const uint8_t *__restrict__ states, const dsb200_hash_entry *__restrict__ hash,
                                                   const uint8_t *__restrict__ visType, int total, const int *__restrict__ tileOffset,
                                                   int cap, int *__restrict__ needed)
{
    __shared__ unsigned sScan[16];
    const int tile0 = tileOffset[blockIdx.x];
    if (tile0 >= cap) return;                                  // uniform per CTA
    const int h0 = blockIdx.x * kTileSlots + threadIdx.x * kSlotsPerThread;
    unsigned flags = 0;
    for (int k = 0; k < kSlotsPerThread; k++)
        if (h0 + k < total && swap_candidate<MODE>(h0 + k, states, hash, visType)) flags |= 1u << k;
    unsigned exA, exB, totA, totB;
    block_scan2((unsigned)__popc(flags), 0u, sScan, exA, exB, totA, totB);
    int rank = tile0 + (int)exA;
    for (int k = 0; k < kSlotsPerThread; k++)
        if ((flags >> k) & 1u) { if (rank < cap) needed[rank] = h0 + k; rank++; }
}

__device__ __forceinline__ unsigned char swap_to_uchar(float v)
{   // Vector3::toUChar: (int)ROUND(x) then CLAMP(0, 255)  (ORUtils/Vector.h:245-247, MathUtils.h:24)
    const int i = (int)((v < 0) ? (v - 0.5f) : (v + 0.5f));
    const int lo = (255 < i) ? 255 : i;
    return (unsigned char)((0 < lo) ? lo : 0);
}

// combineVoxelDepthInformation + combineVoxelColorInformation, DeviceAgnostic/ITMSwappingEngine.h:7-45, on the 8-byte record
// {sdf:16, w_depth:8, r:8 | g:8, b:8, w_color:8, pad:8}
__device__ __forceinline__ uint2 combine_voxel(uint2 src, uint2 dst, int maxW)
{
    {
        int newW = (int)((dst.x >> 16) & 0xffu);
        const int oldW = (int)((src.x >> 16) & 0xffu);
        if (oldW != 0) {
            float newF = (float)(short)(dst.x & 0xffffu) / 32767.0f;
            const float oldF = (float)(short)(src.x & 0xffffu) / 32767.0f;
            newF = (float)oldW * oldF + (float)newW * newF;
            newW = oldW + newW;
            newF /= (float)newW;
            newW = min(newW, maxW);
            const short sdf = (short)(int)(newF * 32767.0f);
            dst.x = (dst.x & 0xff000000u) | ((unsigned)(newW & 0xff) << 16) | (unsigned)(unsigned short)sdf;
        }
    }
    {
        int newW = (int)((dst.y >> 16) & 0xffu);
        const int oldW = (int)((src.y >> 16) & 0xffu);
        if (oldW != 0) {
            float nr = (float)(dst.x >> 24) / 255.0f, ng = (float)(dst.y & 0xffu) / 255.0f, nb = (float)((dst.y >> 8) & 0xffu) / 255.0f;
            const float orr = (float)(src.x >> 24) / 255.0f, og = (float)(src.y & 0xffu) / 255.0f, ob = (float)((src.y >> 8) & 0xffu) / 255.0f;
            nr = orr * (float)oldW + nr * (float)newW;
            ng = og * (float)oldW + ng * (float)newW;
            nb = ob * (float)oldW + nb * (float)newW;
            newW = oldW + newW;
            nr /= (float)newW; ng /= (float)newW; nb /= (float)newW;
            newW = min(newW, maxW);
            const unsigned r = swap_to_uchar(nr * 255.0f), g = swap_to_uchar(ng * 255.0f), b = swap_to_uchar(nb * 255.0f);
            dst.x = (dst.x & 0x00ffffffu) | (r << 24);
            dst.y = (dst.y & 0xff000000u) | ((unsigned)(newW & 0xff) << 16) | (b << 8) | g;
        }
    }
    return dst;
}

// one CTA per transferred entry, two voxels (16 B) per thread; src[i] = the entry's stored block in pinned host memory, or null
__global__ void __launch_bounds__(256) k_swap_integrate(int n, const int *__restrict__ needed, const uint4 *const *__restrict__ src,
                                                        const dsb200_hash_entry *__restrict__ hash, dsb200_voxel *voxels, uint8_t *states,
                                                        int maxW)
{
    const int i = blockIdx.x;
    if (i >= n) return;
    const int entry = needed[i];
    const int ptr = reinterpret_cast<const int *>(hash)[(size_t)entry * 5 + 3];
    const uint4 *from = src[i];
    if (from != nullptr && ptr >= 0) {
        uint4 *dst = reinterpret_cast<uint4 *>(voxels + (size_t)ptr * kBlock3);
        const uint4 a = from[threadIdx.x];
        uint4 b = dst[threadIdx.x];
        const uint2 lo = combine_voxel(make_uint2(a.x, a.y), make_uint2(b.x, b.y), maxW);
        const uint2 hi = combine_voxel(make_uint2(a.z, a.w), make_uint2(b.z, b.w), maxW);
        dst[threadIdx.x] = make_uint4(lo.x, lo.y, hi.x, hi.y);
    }
    if (threadIdx.x == 0) states[entry] = 2;
}

// one CTA per transferred entry: block -> its slot of the host store, state 0, and while the free list has room: push, ptr = -1,
// block reset
__global__ void __launch_bounds__(256) k_swap_save(int n, const int *__restrict__ needed, dsb200_hash_entry *hash, dsb200_voxel *voxels,
                                                   uint4 *const *__restrict__ dstBlocks, uint8_t *states, int *allocList, int lastFree0,
                                                   int roomLimit)
{
    const int i = blockIdx.x;
    if (i >= n) return;
    const int entry = needed[i];
    int *ptrField = reinterpret_cast<int *>(hash) + (size_t)entry * 5 + 3;
    const int ptr = *ptrField;
    __syncthreads();                                           // every thread has read ptr before thread 0 rewrites it
    uint4 *loc = reinterpret_cast<uint4 *>(voxels + (size_t)ptr * kBlock3);
    dstBlocks[i][threadIdx.x] = loc[threadIdx.x];              // 256 threads x 16 B = one 4 KB block, written over PCIe
    const int vbaIdx = lastFree0 + i;                          // the serial loop's noAllocatedVoxelEntries at this entry
    const bool room = vbaIdx < roomLimit;
    if (room) loc[threadIdx.x] = make_uint4(0x00007fffu, 0u, 0x00007fffu, 0u);   // two TVoxel(): sdf 32767, everything else 0
    if (threadIdx.x == 0) {
        states[entry] = 0;
        if (room) { allocList[vbaIdx + 1] = ptr; *ptrField = -1; }
    }
}

} // namespace
} // namespace dsb

using namespace dsb;

struct dsb200_swap {
    int total = 0, transfer = 0, nTiles = 0;
    cudaStream_t stream = nullptr;
    // "global memory": flags + slot per entry, blocks in pinned chunks
    uint8_t *hasStored = nullptr; int *slotOf = nullptr; int slotsUsed = 0;
    int budgetBlocks = 0;          // most blocks the (pinned) host store may hold; 0 = as many as the table has entries
    std::vector<char *> chunks;
    // device
    uint8_t *states = nullptr; int *needed = nullptr, *tileCount = nullptr, *tileOffset = nullptr, *result = nullptr;
    uint4 **blockPtrs = nullptr;
    // pinned
    int *neededHost = nullptr, *resultHost = nullptr; uint4 **blockPtrsHost = nullptr;
    // the reconstruction engine working on the same scene (dsb200_swap_attach_recon): synchronised before and notified after every
    // call here that reads or changes the free list, so a stale host mirror of an asynchronous engine can never be used
    dsb200_recon *recon = nullptr;
};

// before a call that reads scene->last_free_block / changes the volume: bring the host mirrors up to date
static int swap_enter(dsb200_swap *s, dsb200_scene *scene) { return s->recon ? dsb200_recon_refresh_mirrors(s->recon, scene, nullptr) : DSB200_OK; }
// after it: the engine's device counters, look-ahead and sdf shadow describe the old volume
static int swap_leave(dsb200_swap *s, const dsb200_scene *scene) { return s->recon ? dsb200_recon_notify_volume_swapped(s->recon, scene) : DSB200_OK; }

static const int kChunkBlocks = 16384;                          // 64 MB of pinned memory per chunk
static const size_t kBlockBytes = sizeof(dsb200_voxel) * DSB200_BLOCK_SIZE3;

#define SWCU(call)                                                                                   \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            static thread_local char buf_[256];                                                      \
            snprintf(buf_, sizeof(buf_), "%s: %s", #call, cudaGetErrorString(e_));                   \
            return fail_with(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? DSB200_ERR_NO_DEVICE : DSB200_ERR_CUDA, buf_); \
        }                                                                                            \
    } while (0)

static char *slot_address(const dsb200_swap *s, int slot) { return s->chunks[slot / kChunkBlocks] + (size_t)(slot % kChunkBlocks) * kBlockBytes; }

static bool add_chunk(dsb200_swap *s)
{
    char *c = nullptr;
    if (cudaHostAlloc(&c, (size_t)kChunkBlocks * kBlockBytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return false; }
    s->chunks.push_back(c);
    return true;
}

// the entry's slot in the store, assigned (and a chunk allocated) on first use; null on allocation failure
static char *slot_for(dsb200_swap *s, int entry)
{
    if (s->slotOf[entry] < 0) {
        if (s->slotsUsed == (int)s->chunks.size() * kChunkBlocks && !add_chunk(s)) return nullptr;
        s->slotOf[entry] = s->slotsUsed++;
    }
    return slot_address(s, s->slotOf[entry]);
}

// rank the candidates in entry order; *n = how many will move (<= transfer), their ids in s->needed and s->neededHost
template <int MODE>
static int swap_select(dsb200_swap *s, const dsb200_scene *scene, const uint8_t *visType, int *n)
{
    DSB_LAUNCH(k_swap_count<MODE>, s->nTiles, 256, 0, s->stream, s->states, scene->hash, visType, s->total, s->tileCount);
    DSB_LAUNCH(k_swap_scan, 1, 1024, 0, s->stream, s->tileCount, s->nTiles, s->tileOffset, s->result);
    DSB_LAUNCH(k_swap_emit<MODE>, s->nTiles, 256, 0, s->stream, s->states, scene->hash, visType, s->total, s->tileOffset, s->transfer,
               s->needed);
    SWCU(cudaGetLastError());
    SWCU(cudaMemcpyAsync(s->resultHost, s->result, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    SWCU(cudaStreamSynchronize(s->stream));
    *n = s->resultHost[0] < s->transfer ? s->resultHost[0] : s->transfer;
    if (*n > 0) {
        SWCU(cudaMemcpyAsync(s->neededHost, s->needed, sizeof(int) * (size_t)*n, cudaMemcpyDeviceToHost, s->stream));
        SWCU(cudaStreamSynchronize(s->stream));
    }
    return DSB200_OK;
}

extern "C" {

int dsb200_swap_destroy(dsb200_swap *s)
{
    if (!s) return DSB200_OK;
    free(s->hasStored); free(s->slotOf);
    for (char *c : s->chunks) cudaFreeHost(c);
    cudaFree(s->states); cudaFree(s->needed); cudaFree(s->tileCount); cudaFree(s->tileOffset); cudaFree(s->result); cudaFree(s->blockPtrs);
    cudaFreeHost(s->neededHost); cudaFreeHost(s->resultHost); cudaFreeHost(s->blockPtrsHost);
    delete s;
    return DSB200_OK;
}

int dsb200_swap_create(dsb200_swap **out, int32_t total_entries, int32_t transfer_blocks, void *stream)
{
    if (!out || total_entries <= 0 || transfer_blocks <= 0) return fail_with(DSB200_ERR_INVALID, "swap_create: bad arguments");
    if (dsb200_device_count() == 0) return fail_with(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    dsb200_swap *s = new dsb200_swap();
    s->total = total_entries; s->transfer = transfer_blocks; s->stream = (cudaStream_t)stream;
    s->nTiles = div_up(total_entries, kTileSlots);
    s->hasStored = (uint8_t *)calloc((size_t)total_entries, 1);
    s->slotOf = (int *)malloc(sizeof(int) * (size_t)total_entries);
    if (!s->hasStored || !s->slotOf) { dsb200_swap_destroy(s); return fail_with(DSB200_ERR_INVALID, "swap_create: host allocation failed"); }
    for (int i = 0; i < total_entries; i++) s->slotOf[i] = -1;
    cudaError_t e = cudaSuccess;
    if (e == cudaSuccess) e = cudaMalloc(&s->states, (size_t)total_entries);
    if (e == cudaSuccess) e = cudaMemset(s->states, 0, (size_t)total_entries);
    if (e == cudaSuccess) e = cudaMalloc(&s->needed, sizeof(int) * (size_t)transfer_blocks);
    if (e == cudaSuccess) e = cudaMalloc(&s->tileCount, sizeof(int) * (size_t)s->nTiles);
    if (e == cudaSuccess) e = cudaMalloc(&s->tileOffset, sizeof(int) * (size_t)s->nTiles);
    if (e == cudaSuccess) e = cudaMalloc(&s->result, sizeof(int) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&s->blockPtrs, sizeof(uint4 *) * (size_t)transfer_blocks);
    if (e == cudaSuccess) e = cudaMallocHost(&s->neededHost, sizeof(int) * (size_t)transfer_blocks);
    if (e == cudaSuccess) e = cudaMallocHost(&s->resultHost, sizeof(int) * 4);
    if (e == cudaSuccess) e = cudaMallocHost(&s->blockPtrsHost, sizeof(uint4 *) * (size_t)transfer_blocks);
    if (e != cudaSuccess) { dsb200_swap_destroy(s); SWCU(e); }
    // the first chunk of the store now, so that pinning 64 MB (tens of milliseconds) is not part of the first swap-out
    if (!add_chunk(s)) { dsb200_swap_destroy(s); return fail_with(DSB200_ERR_CUDA, "swap_create: could not allocate the first pinned chunk of the host store"); }
    *out = s;
    return DSB200_OK;
}

int dsb200_swap_cache_of(dsb200_swap *s, dsb200_swap_cache *out)
{
    if (!s || !out) return fail_with(DSB200_ERR_INVALID, "swap_cache_of: null argument");
    out->states = s->states; out->has_stored = s->hasStored; out->stored = nullptr;     // blocks: dsb200_swap_stored_block
    out->total_entries = s->total; out->transfer_blocks = s->transfer;
    return DSB200_OK;
}

const dsb200_voxel *dsb200_swap_stored_block(const dsb200_swap *s, int32_t entry)
{
    if (!s || entry < 0 || entry >= s->total || !s->hasStored[entry] || s->slotOf[entry] < 0) return nullptr;
    return reinterpret_cast<const dsb200_voxel *>(slot_address(s, s->slotOf[entry]));
}

int dsb200_swap_mark_visible(dsb200_swap *s, const dsb200_render_state *rs)
{
    if (!s || !rs || !rs->visible_type) return fail_with(DSB200_ERR_INVALID, "swap_mark_visible: null argument");
    int grid = div_up(s->total, 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    DSB_LAUNCH(k_swap_mark_visible, grid, 256, 0, s->stream, s->states, rs->visible_type, s->total);
    SWCU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_swap_integrate_global_into_local(dsb200_swap *s, dsb200_scene *scene, int32_t *moved)
{
    if (!s || !scene) return fail_with(DSB200_ERR_INVALID, "swap_integrate_global_into_local: null argument");
    if (scene->bucket_num + scene->excess_size != s->total) return fail_with(DSB200_ERR_INVALID, "swap: the cache was sized for another hash table");
    int n = 0;
    int rc = swap_enter(s, scene);
    if (rc) return rc;
    rc = swap_select<0>(s, scene, nullptr, &n);
    if (rc) return rc;
    if (moved) *moved = n;
    if (n == 0) return DSB200_OK;
    // LoadFromGlobalMemory (CPU .cpp:52-68) without the copy: the kernel reads the stored blocks where they lie
    for (int i = 0; i < n; i++) {
        const int entry = s->neededHost[i];
        s->blockPtrsHost[i] = (s->hasStored[entry] && s->slotOf[entry] >= 0) ? reinterpret_cast<uint4 *>(slot_address(s, s->slotOf[entry])) : nullptr;
    }
    SWCU(cudaMemcpyAsync(s->blockPtrs, s->blockPtrsHost, sizeof(uint4 *) * (size_t)n, cudaMemcpyHostToDevice, s->stream));
    DSB_LAUNCH(k_swap_integrate, n, 256, 0, s->stream, n, s->needed, s->blockPtrs, scene->hash, scene->voxels, s->states, scene->params.max_w);
    SWCU(cudaGetLastError());
    SWCU(cudaStreamSynchronize(s->stream));
    return swap_leave(s, scene);
}

int dsb200_swap_save_to_global_memory(dsb200_swap *s, dsb200_scene *scene, const dsb200_render_state *rs, int32_t *moved)
{
    if (!s || !scene) return fail_with(DSB200_ERR_INVALID, "swap_save_to_global_memory: null argument");
    if (scene->bucket_num + scene->excess_size != s->total) return fail_with(DSB200_ERR_INVALID, "swap: the cache was sized for another hash table");
    if (rs && !rs->visible_type) return fail_with(DSB200_ERR_INVALID, "swap_save_to_global_memory: render state without visibility flags");
    int n = 0;
    int rc = swap_enter(s, scene);                                 // scene->last_free_block below must be the engine's current value
    if (rc) return rc;
    rc = swap_select<1>(s, scene, rs ? rs->visible_type : nullptr, &n);
    if (rc) return rc;
    if (moved) *moved = n;
    if (n == 0) return DSB200_OK;
    // Destinations first (SetStoredData, Objects/ITMGlobalCache.h:30-34); nothing has been changed yet.  The store is pinned
    // memory: when its budget is used up or the next chunk cannot be pinned, the call moves the candidates that still fit (a
    // shorter transfer, as if the transfer limit were lower) and fails cleanly only when none does.
    int fit = n;
    for (int i = 0; i < n; i++) {
        const int entry = s->neededHost[i];
        char *slot = nullptr;
        if (s->slotOf[entry] >= 0 || s->budgetBlocks <= 0 || s->slotsUsed < s->budgetBlocks) slot = slot_for(s, entry);
        if (!slot) { fit = i; break; }
        s->blockPtrsHost[i] = reinterpret_cast<uint4 *>(slot);
    }
    if (fit == 0) {
        if (moved) *moved = 0;
        return fail_with(DSB200_ERR_CUDA, "swap_save_to_global_memory: the host block store is full (dsb200_swap_set_store_budget, or no more "
                                          "memory could be pinned); nothing was moved");
    }
    n = fit;
    if (moved) *moved = n;
    SWCU(cudaMemcpyAsync(s->blockPtrs, s->blockPtrsHost, sizeof(uint4 *) * (size_t)n, cudaMemcpyHostToDevice, s->stream));
    // the reference lets the free list grow while vbaIdx < SDF_BUCKET_NUM - 1 (CPU .cpp:178); the list itself holds num_blocks
    // entries, so the second bound only differs where the reference would write past its end
    const int lastFree0 = scene->last_free_block;
    const int roomLimit = (scene->bucket_num - 1 < scene->num_blocks - 1) ? scene->bucket_num - 1 : scene->num_blocks - 1;
    DSB_LAUNCH(k_swap_save, n, 256, 0, s->stream, n, s->needed, scene->hash, scene->voxels, s->blockPtrs, s->states, scene->alloc_list,
               lastFree0, roomLimit);
    SWCU(cudaGetLastError());
    SWCU(cudaStreamSynchronize(s->stream));                    // the blocks are in the store when the call returns
    for (int i = 0; i < n; i++) s->hasStored[s->neededHost[i]] = 1;
    int pushed = roomLimit - lastFree0;
    if (pushed < 0) pushed = 0;
    if (pushed > n) pushed = n;
    scene->last_free_block = lastFree0 + pushed;
    return swap_leave(s, scene);
}

int dsb200_swap_set_store_budget(dsb200_swap *s, int32_t max_blocks)
{
    if (!s || max_blocks < 0) return fail_with(DSB200_ERR_INVALID, "swap_set_store_budget: bad argument");
    s->budgetBlocks = max_blocks;
    return DSB200_OK;
}

int dsb200_swap_attach_recon(dsb200_swap *s, dsb200_recon *recon)
{
    if (!s) return fail_with(DSB200_ERR_INVALID, "swap_attach_recon: null cache");
    s->recon = recon;
    return DSB200_OK;
}

int dsb200_swap_save_file(dsb200_swap *s, const char *path)
{
    if (!s || !path) return fail_with(DSB200_ERR_INVALID, "swap_save_file: null argument");
    FILE *f = fopen(path, "wb");
    if (!f) return fail_with(DSB200_ERR_INVALID, "swap_save_file: could not open the file for writing");
    std::vector<char> zeros(kBlockBytes, 0);                    // a block that was never stored is all-zero in the reference's file
    bool ok = fwrite(s->hasStored, 1, (size_t)s->total, f) == (size_t)s->total;
    for (int i = 0; ok && i < s->total; i++)
        ok = fwrite(s->hasStored[i] ? slot_address(s, s->slotOf[i]) : zeros.data(), kBlockBytes, 1, f) == 1;
    ok = (fclose(f) == 0) && ok;
    return ok ? DSB200_OK : fail_with(DSB200_ERR_INVALID, "swap_save_file: short write");
}

int dsb200_swap_load_file(dsb200_swap *s, const char *path)
{
    if (!s || !path) return fail_with(DSB200_ERR_INVALID, "swap_load_file: null argument");
    FILE *f = fopen(path, "rb");
    if (!f) return fail_with(DSB200_ERR_INVALID, "swap_load_file: could not open the file");
    // The file is `total` flags followed by `total` blocks (ITMGlobalCache's layout): check its size up front -- a seek past
    // the end of a truncated file succeeds -- and read the flags into a temporary: a flag is committed to the live cache only
    // once its block has been read, so a failed load never leaves "stored" entries without a slot.
    const long long want = (long long)s->total * (1 + (long long)kBlockBytes);
    bool ok = fseek(f, 0, SEEK_END) == 0 && (long long)ftell(f) == want && fseek(f, 0, SEEK_SET) == 0;
    if (!ok) { fclose(f); return fail_with(DSB200_ERR_INVALID, "swap_load_file: file size differs from total * (1 + 4096) (other capacities, or truncated)"); }
    std::vector<unsigned char> flags((size_t)s->total);
    ok = fread(flags.data(), 1, (size_t)s->total, f) == (size_t)s->total;
    int failedAt = -1;
    for (int i = 0; ok && i < s->total; i++) {
        if (flags[i]) {
            char *slot = slot_for(s, i);
            ok = slot != nullptr && fread(slot, kBlockBytes, 1, f) == 1;
            if (ok) s->hasStored[i] = 1; else failedAt = i;
        } else {
            ok = fseek(f, (long)kBlockBytes, SEEK_CUR) == 0;
            if (ok) s->hasStored[i] = 0;
        }
    }
    fclose(f);
    if (!ok) {
        if (failedAt >= 0) s->hasStored[failedAt] = 0;       // its slot (if any) holds a partial block: not stored
        for (int i = 0; i < s->total; i++) if (s->slotOf[i] < 0) s->hasStored[i] = 0;
        return fail_with(DSB200_ERR_INVALID, "swap_load_file: read failed; entries read so far are kept, the rest are not stored");
    }
    return DSB200_OK;
}

} // extern "C"
