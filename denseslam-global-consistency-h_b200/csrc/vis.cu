// vis.cu -- ITMVisualisationEngine for sm_100a: FindVisibleBlocks, CreateExpectedDepths,
// raycast (castRay), ICP point/normal maps, free-view shading.
//
// Reference math: DeviceAgnostic/ITMVisualisationEngine.h:36-196,243-439 and
// DeviceAgnostic/ITMRepresentationAccess.h:176-427; reference kernels:
// CUDA/ITMVisualisationEngine_CUDA.cu:655-942.  Differences in structure (not in results):
//  * visible-block compaction is ORDERED (ascending slot) like the CPU engine, via the look-back scan;
//  * expected depths: no rendering-block list and no CAS-loop float atomics -- every visible block
//    projects its 8 corners and min/max-es its bounding box straight into the range image with
//    native integer atomics on the float bit patterns (all depths are >= 0.05 > 0, so integer
//    order == float order); the reference's rendering-block cap is honoured by an exact serial
//    fallback that only runs when the cap would be hit;
//  * nothing returns to the host between stages (counts stay on the device).
#include <cstdlib>
#include <cstring>
#include "kernels.cuh"
#include "scan.cuh"
#include "visibility.cuh"
#include "exactdiv.cuh"

namespace dsb {


constexpr float kFarAway = 999999.9f, kVeryClose = 0.05f;
constexpr int kSubsample = 8;

__global__ void k_reset_scan(FrameState fs, int region, int tiles, int ticket)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < tiles; i += gridDim.x * blockDim.x) fs.scanState[region + i] = 0ull;
    if (blockIdx.x == 0 && threadIdx.x == 0) fs.tickets[ticket] = 0;
}

// FindVisibleBlocks (CPU/ITMVisualisationEngine_CPU.cpp:40-78): all allocated slots inside the frustum, ascending
__global__ void __launch_bounds__(256) k_find_visible(VisParams P, FrameState fs, const dsb200_hash_entry *__restrict__ hash,
                                                      int *visibleIds, int visibleCapacity, int total)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx[2];
    if (threadIdx.x == 0) sTile = atomicAdd(&fs.tickets[1], 1);
    __syncthreads();
    const int tile = sTile;
    const int slot0 = tile * 1024 + threadIdx.x * 4;
    bool v[4]; unsigned na = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        v[j] = false;
        if (slot0 + j < total) {
            HashEntry e = load_entry(hash, slot0 + j);
            if (e.ptr >= 0) {
                bool vis, visEnl;
                block_visibility(vis, visEnl, e.px, e.py, e.pz, P.M, P.proj, P.voxelSize, P.W, P.H, false);
                v[j] = vis;
            }
        }
        na += v[j] ? 1 : 0;
    }
    unsigned exA, exB, totA, totB;
    block_scan2(na, 0, sScan, exA, exB, totA, totB);
    if (threadIdx.x < 32) {
        unsigned ta, tb;
        scan_lookback(fs.scanState + fs.scanTiles, tile, totA, 0, ta, tb);
        if (threadIdx.x == 0) sEx[0] = ta;
    }
    __syncthreads();
    unsigned r = sEx[0] + exA;
#pragma unroll
    for (int j = 0; j < 4; j++) if (v[j]) { if ((int)r < visibleCapacity) visibleIds[r] = slot0 + j; r++; }
    if (threadIdx.x == 0 && tile == gridDim.x - 1) fs.counters->noVisible = (int)(sEx[0] + totA);
}

__global__ void k_count_visible(const dsb200_hash_entry *__restrict__ hash, const int *__restrict__ ids,
                                const int *__restrict__ countPtr, int minId, int maxId, int *out)
{
    const int n = *countPtr;
    int c = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int ptr = load_entry(hash, ids[i]).ptr;
        c += (ptr >= minId && ptr <= maxId) ? 1 : 0;
    }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// ---------------------------------------------------------------- expected depths
// ProjectSingleBlock, DeviceAgnostic/ITMVisualisationEngine.h:36-85
__device__ __forceinline__ bool project_block(short bx, short by, short bz, const ProjParams &P, int ul[2], int lr[2], float zr[2])
{
    ul[0] = P.W / kSubsample; ul[1] = P.H / kSubsample; lr[0] = -1; lr[1] = -1; zr[0] = kFarAway; zr[1] = kVeryClose;
#pragma unroll
    for (int corner = 0; corner < 8; ++corner) {
        short tx = bx + ((corner & 1) ? 1 : 0), ty = by + ((corner & 2) ? 1 : 0), tz = bz + ((corner & 4) ? 1 : 0);
        Vec4 p = {(float)tx * (float)kBlock * P.voxelSize, (float)ty * (float)kBlock * P.voxelSize,
                  (float)tz * (float)kBlock * P.voxelSize, 1.0f};
        p = mul(P.M, p);
        if ((double)p.z < 1e-6) continue;           // double literal in the reference
        float px = (P.proj[0] * p.x / p.z + P.proj[2]) / (float)kSubsample;
        float py = (P.proj[1] * p.y / p.z + P.proj[3]) / (float)kSubsample;
        if (ul[0] > floorf(px)) ul[0] = (int)floorf(px);
        if (lr[0] < ceilf(px)) lr[0] = (int)ceilf(px);
        if (ul[1] > floorf(py)) ul[1] = (int)floorf(py);
        if (lr[1] < ceilf(py)) lr[1] = (int)ceilf(py);
        if (zr[0] > p.z) zr[0] = p.z;
        if (zr[1] < p.z) zr[1] = p.z;
    }
    if (ul[0] < 0) ul[0] = 0;
    if (ul[1] < 0) ul[1] = 0;
    if (lr[0] >= P.W) lr[0] = P.W - 1;
    if (lr[1] >= P.H) lr[1] = P.H - 1;
    if (ul[0] > lr[0]) return false;
    if (ul[1] > lr[1]) return false;
    if (zr[0] < kVeryClose) zr[0] = kVeryClose;
    if (zr[1] < kVeryClose) return false;
    return true;
}

__global__ void __launch_bounds__(256) k_range_init(float2 *range, int n, Counters *counters, int *orderCtl)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) range[i] = make_float2(kFarAway, kVeryClose);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        counters->renderBlocks = 0;
        if (orderCtl) { orderCtl[0] = 0; orderCtl[1] = 0; orderCtl[2] = 0; orderCtl[3] = 0; }      // k_tile_order's class counters
    }
}

__device__ __forceinline__ void range_minmax(float *range, int W, int x, int y, float zmin, float zmax)
{
    int *px = reinterpret_cast<int *>(range) + 2 * ((size_t)x + (size_t)y * W);
    // positive floats: integer order == float order; skip the atomic when it cannot change the value
    if (__float_as_int(zmin) < __ldcg(px)) atomicMin(px, __float_as_int(zmin));
    if (__float_as_int(zmax) > __ldcg(px + 1)) atomicMax(px + 1, __float_as_int(zmax));
}

// One WARP per visible block: lanes 0-7 project one corner each (same arithmetic as project_block; min/max are
// order-independent, so the warp reduction gives the identical bounding box and depth range), then all 32 lanes spread
// over the bounding-box cells.  A thread-per-block version serialises the (up to ~30x30 cell) boxes of near blocks.
__global__ void __launch_bounds__(256) k_project_fill(ProjParams P, const dsb200_hash_entry *__restrict__ hash,
                                                      const int *__restrict__ ids, const int *__restrict__ countPtr,
                                                      float *range, Counters *counters)
{
    const int n = *countPtr;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nWarps = (gridDim.x * blockDim.x) >> 5;
    int tiles = 0;
    for (int i = warp; i < n; i += nWarps) {
        HashEntry e = load_entry(hash, __ldg(ids + i));
        if (e.ptr < 0) continue;
        // corner `lane & 7` (lanes 8-31 duplicate corners: harmless for min/max)
        const int corner = lane & 7;
        int ulx = P.W / kSubsample, uly = P.H / kSubsample, lrx = -1, lry = -1;
        float zmin = kFarAway, zmax = kVeryClose;
        {
            short tx = e.px + ((corner & 1) ? 1 : 0), ty = e.py + ((corner & 2) ? 1 : 0), tz = e.pz + ((corner & 4) ? 1 : 0);
            Vec4 p = {(float)tx * (float)kBlock * P.voxelSize, (float)ty * (float)kBlock * P.voxelSize,
                      (float)tz * (float)kBlock * P.voxelSize, 1.0f};
            p = mul(P.M, p);
            if (!((double)p.z < 1e-6)) {
                float px = (P.proj[0] * p.x / p.z + P.proj[2]) / (float)kSubsample;
                float py = (P.proj[1] * p.y / p.z + P.proj[3]) / (float)kSubsample;
                if (ulx > floorf(px)) ulx = (int)floorf(px);
                if (lrx < ceilf(px)) lrx = (int)ceilf(px);
                if (uly > floorf(py)) uly = (int)floorf(py);
                if (lry < ceilf(py)) lry = (int)ceilf(py);
                if (zmin > p.z) zmin = p.z;
                if (zmax < p.z) zmax = p.z;
            }
        }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            ulx = min(ulx, __shfl_xor_sync(0xffffffffu, ulx, o)); uly = min(uly, __shfl_xor_sync(0xffffffffu, uly, o));
            lrx = max(lrx, __shfl_xor_sync(0xffffffffu, lrx, o)); lry = max(lry, __shfl_xor_sync(0xffffffffu, lry, o));
            zmin = fminf(zmin, __shfl_xor_sync(0xffffffffu, zmin, o)); zmax = fmaxf(zmax, __shfl_xor_sync(0xffffffffu, zmax, o));
        }
        if (ulx < 0) ulx = 0;
        if (uly < 0) uly = 0;
        if (lrx >= P.W) lrx = P.W - 1;
        if (lry >= P.H) lry = P.H - 1;
        if (ulx > lrx || uly > lry) continue;
        if (zmin < kVeryClose) zmin = kVeryClose;
        if (zmax < kVeryClose) continue;
        const int bw = lrx - ulx + 1, bh = lry - uly + 1;
        if (lane == 0) tiles += (int)ceilf((float)bw / 16.0f) * (int)ceilf((float)bh / 16.0f);
        for (int c = lane; c < bw * bh; c += 32) {
            const int cy = c / bw;
            range_minmax(range, P.W, ulx + (c - cy * bw), uly + cy, zmin, zmax);
        }
    }
    for (int o = 16; o > 0; o >>= 1) tiles += __shfl_xor_sync(0xffffffffu, tiles, o);
    if (lane == 0 && tiles) atomicAdd(&counters->renderBlocks, tiles);
}

// Exact fallback for the reference's MAX_RENDERING_BLOCKS rule (CPU .cpp:158: a block whose tiles
// would push the running count to >= cap is skipped, later smaller blocks may still fit).  If the
// grand total is below the cap no block can have been skipped and k_project_fill's result stands.
__global__ void __launch_bounds__(1024) k_range_cap_fallback(ProjParams P, const dsb200_hash_entry *__restrict__ hash,
                                                             const int *__restrict__ ids, const int *__restrict__ countPtr,
                                                             float *range, Counters *counters)
{
    if (counters->renderBlocks < P.maxRenderBlocks) return;
    __shared__ int sNum;
    const int n = *countPtr;
    for (int i = threadIdx.x; i < P.W * P.H; i += blockDim.x) { range[2 * i] = kFarAway; range[2 * i + 1] = kVeryClose; }
    if (threadIdx.x == 0) sNum = 0;
    __syncthreads();
    for (int i = 0; i < n; i++) {
        HashEntry e = load_entry(hash, ids[i]);
        int ul[2], lr[2]; float zr[2];
        bool ok = e.ptr >= 0 && project_block(e.px, e.py, e.pz, P, ul, lr, zr);
        int req = 0;
        if (ok) {
            req = (int)ceilf((float)(lr[0] - ul[0] + 1) / 16.0f) * (int)ceilf((float)(lr[1] - ul[1] + 1) / 16.0f);
            if (sNum + req >= P.maxRenderBlocks) ok = false;
        }
        __syncthreads();
        if (ok) {
            if (threadIdx.x == 0) sNum += req;
            const int bw = lr[0] - ul[0] + 1, cells = bw * (lr[1] - ul[1] + 1);
            for (int c = threadIdx.x; c < cells; c += blockDim.x) {
                float *px = range + 2 * ((size_t)(ul[0] + c % bw) + (size_t)(ul[1] + c / bw) * P.W);
                if (px[0] > zr[0]) px[0] = zr[0];
                if (px[1] < zr[1]) px[1] = zr[1];
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) counters->renderBlocks = sNum;
}

// LONG TILES FIRST.  A ray marches from the near to the far expected depth of its 8x8-pixel range cell; on a street the cells at
// the horizon span tens of metres of mostly free space (tools/ray_trip_stats.py: up to 152 loop trips, against a median of 8) and
// a warp that draws such a tile late finishes long after everything else -- the kernel lasts as long as "start of the longest tile
// + its march".  The range image is known before the raycast starts, so the tickets are dealt from a list that holds the tiles
// whose depth span is worth at least `minTrips` 8-voxel steps at the front (they all start in the first wave) and the others at the
// back.  Warp-aggregated atomics; k_range_init zeroes the counters.  Only the ORDER of the tickets changes.
// Four classes by span (>= L, >= L/2, >= L/4, rest; L = `spanLong`), one list each: order = {4 counters, 4 lists of n entries}.
constexpr int kOrderClasses = 4;
__global__ void __launch_bounds__(256) k_tile_order(const float2 *__restrict__ range, int W, int H, float spanLong, int *order)
{
    const int tilesX = (W + 7) >> 3, tilesY = (H + 3) >> 2, n = tilesX * tilesY;
    const int t = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
    int cls = -1;
    if (t < n) {
        const int ty = t / tilesX, tx = t - ty * tilesX;
        const float2 mm = __ldg(range + tx + ((ty * 4) / kSubsample) * W);
        const float span = mm.y - mm.x;
        cls = span >= spanLong ? 0 : (span >= 0.5f * spanLong ? 1 : (span >= 0.25f * spanLong ? 2 : 3));
    }
#pragma unroll
    for (int c = 0; c < kOrderClasses; c++) {
        const unsigned m = __ballot_sync(0xffffffffu, cls == c);
        if (!m) continue;
        int base = 0;
        if (lane == 0) base = atomicAdd(&order[c], __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (cls == c) order[kOrderClasses + (size_t)c * n + base + __popc(m & ((1u << lane) - 1u))] = t;
    }
}

// ---------------------------------------------------------------- raycast
struct BlockCache { int bx, by, bz; const short *blk; };     // blk: first sdf of the block (element stride Tables::kStride shorts)

// Where a block's hash table and voxel array live.  One scene: the two pointers.  Sharded scene (dsb200_raycast_sharded):
// the tables of the shard that OWNS the block, which may be peer-GPU memory mapped with CUDA IPC -- the probe and the
// voxel read then travel over NVLink, and the march is the single-scene march probe for probe.
// The sdf samples come either from the voxel array itself (8-byte ITMVoxel records: stride 4 shorts) or from the
// reconstruction engine's SDF SHADOW (one short per voxel, stride 1; capi.cu): a quarter of the bytes, so the visible set
// (36 MB at the bench workload instead of 143 MB) stays L2-resident between integration and raycast.
template <int STRIDE>
struct SceneTablesT {
    static constexpr int kStride = STRIDE;
    const short *sdf; const dsb200_hash_entry *hash;
    __device__ __forceinline__ void resolve(int, int, int, const dsb200_hash_entry *&h, const short *&v) const { h = hash; v = sdf; }
};
typedef SceneTablesT<4> SceneTables;      // straight from the voxel array
typedef SceneTablesT<1> ShadowTables;     // from the sdf shadow
__device__ __host__ __forceinline__ SceneTables scene_tables(const dsb200_voxel *voxels, const dsb200_hash_entry *hash)
{
    SceneTables T; T.sdf = reinterpret_cast<const short *>(voxels); T.hash = hash; return T;
}
struct ShardTables {
    static constexpr int kStride = 4;
    int n, brickLog2;
    const dsb200_hash_entry *hash[kMaxShards];
    const dsb200_voxel *voxels[kMaxShards];
    __device__ __forceinline__ void resolve(int bx, int by, int bz, const dsb200_hash_entry *&h, const short *&v) const
    {
        const int o = shard_owner(bx, by, bz, n, brickLog2);
        h = hash[o]; v = reinterpret_cast<const short *>(voxels[o]);
    }
};

// readVoxel + pointToVoxelBlockPos (DeviceAgnostic/ITMRepresentationAccess.h:14-19,176-210); returns sdf as short
template <class Tables>
__device__ __forceinline__ short read_sdf(const Tables &T, int bucketNum, int bucketMask, int px, int py, int pz, bool &found,
                                          BlockCache &c)
{
    // floor division by 8 (arithmetic shift == the reference's ((p<0)?p-7:p)/8)
    const int bx = px >> 3, by = py >> 3, bz = pz >> 3;
    const int lin = (px & 7) + ((py & 7) << 3) + ((pz & 7) << 6);
    if (bx == c.bx && by == c.by && bz == c.bz) { found = true; return __ldg(c.blk + lin * Tables::kStride); }
    const dsb200_hash_entry *hash; const short *voxels;
    T.resolve(bx, by, bz, hash, voxels);
    int h = hash_index(bx, by, bz, bucketMask);
    while (true) {
        HashEntry e = load_entry(hash, h);
        if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= 0) {
            found = true; c.bx = bx; c.by = by; c.bz = bz; c.blk = voxels + (size_t)e.ptr * (kBlock3 * Tables::kStride);
            return __ldg(c.blk + lin * Tables::kStride);
        }
        if (e.offset < 1) break;
        h = bucketNum + e.offset - 1;
    }
    found = false;
    return 32767;
}

// x / 32767.0f, correctly rounded, in three instructions instead of the compiler's generic division sequence: one Markstein step
// with the correctly rounded reciprocal is exact for this divisor (verified exhaustively over a full binade of numerators,
// tests/test_exactdiv.py / oracle/exactdiv_check.c; 0 and every scaling by a power of two follow)
__device__ __forceinline__ float div_32767(float x) { return div_step1(x, 32767.0f, 1.0f / 32767.0f); }

__device__ __forceinline__ float round_half_away(float x) { return (x < 0) ? (x - 0.5f) : (x + 0.5f); }

// findVoxel's block part: hash probe -> first voxel of the block, or nullptr (ITMRepresentationAccess.h:176-210)
template <class Tables>
__device__ __forceinline__ const short *resolve_block(const Tables &T, int bucketNum, int bucketMask, int bx, int by, int bz)
{
    const dsb200_hash_entry *hash; const short *voxels;
    T.resolve(bx, by, bz, hash, voxels);
    int h = hash_index(bx, by, bz, bucketMask);
    while (true) {
        HashEntry e = load_entry(hash, h);
        if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= 0) return voxels + (size_t)e.ptr * (kBlock3 * Tables::kStride);
        if (e.offset < 1) return nullptr;
        h = bucketNum + e.offset - 1;
    }
}

// readFromSDF_float_interpolated (:251-276).  The reference reads the 8 corners one after the other through its 1-entry
// IndexCache; next to a block face the corners alternate between two (up to eight) blocks and every read re-probes the
// hash table -- on the GPU that is up to 16 dependent memory round trips for one sample (measured: ~6 us per march step).
// Here the up-to-8 distinct blocks of the 2x2x2 neighbourhood are resolved once each (only the ones the sample really
// crosses into), then the 8 voxels are loaded together.  (Resolving the 1, 3 or 7 neighbours in one loop over the needed set,
// so that the lanes of a warp probe together, measured 3.4x SLOWER on B200 -- profiles/README.md r02d -- and was dropped.)  Same voxels, same arithmetic, same result; a missing block reads as
// the reference's default voxel (sdf 32767).
template <class Tables>
__device__ __forceinline__ float sdf_trilinear(const Tables &T, int bucketNum, int bucketMask, float x, float y, float z, bool &found,
                                               BlockCache &cache)
{
    const float fx = floorf(x), fy = floorf(y), fz = floorf(z);
    const float cx = x - fx, cy = y - fy, cz = z - fz;
    const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    const int bx = ix >> 3, by = iy >> 3, bz = iz >> 3;
    const int lx = ix & 7, ly = iy & 7, lz = iz & 7;
    const bool ex = lx == 7, ey = ly == 7, ez = lz == 7;        // the +1 corner leaves the block along this axis
    const short *p000;
    if (bx == cache.bx && by == cache.by && bz == cache.bz) p000 = cache.blk;
    else {
        p000 = resolve_block(T, bucketNum, bucketMask, bx, by, bz);
        if (p000) { cache.bx = bx; cache.by = by; cache.bz = bz; cache.blk = p000; }
    }
    // The neighbour blocks the sample crosses into: every crossing axis doubles them, i.e. 0, 1, 3 or 7 blocks, each resolved at its
    // own conditional call site (the single-loop form measured 3.4x slower, see above).
    const short *p100 = ex ? resolve_block(T, bucketNum, bucketMask, bx + 1, by, bz) : p000;
    const short *p010 = ey ? resolve_block(T, bucketNum, bucketMask, bx, by + 1, bz) : p000;
    const short *p001 = ez ? resolve_block(T, bucketNum, bucketMask, bx, by, bz + 1) : p000;
    const short *p110 = (ex && ey) ? resolve_block(T, bucketNum, bucketMask, bx + 1, by + 1, bz) : (ex ? p100 : p010);
    const short *p101 = (ex && ez) ? resolve_block(T, bucketNum, bucketMask, bx + 1, by, bz + 1) : (ex ? p100 : p001);
    const short *p011 = (ey && ez) ? resolve_block(T, bucketNum, bucketMask, bx, by + 1, bz + 1) : (ey ? p010 : p001);
    const short *p111 = (ex && ey && ez) ? resolve_block(T, bucketNum, bucketMask, bx + 1, by + 1, bz + 1)
                                                : ((ex && ey) ? p110 : ((ex && ez) ? p101 : ((ey && ez) ? p011 : (ex ? p100 : (ey ? p010 : p001)))));
    const int x1 = (lx + 1) & 7, y1 = ((ly + 1) & 7) << 3, z1 = ((lz + 1) & 7) << 6, y0 = ly << 3, z0 = lz << 6;
#define DSB_CORNER(P, L) ((P) ? (float)__ldg((P) + (L) * Tables::kStride) : 32767.0f)
    const float v000 = DSB_CORNER(p000, lx + y0 + z0), v100 = DSB_CORNER(p100, x1 + y0 + z0);
    const float v010 = DSB_CORNER(p010, lx + y1 + z0), v110 = DSB_CORNER(p110, x1 + y1 + z0);
    const float v001 = DSB_CORNER(p001, lx + y0 + z1), v101 = DSB_CORNER(p101, x1 + y0 + z1);
    const float v011 = DSB_CORNER(p011, lx + y1 + z1), v111 = DSB_CORNER(p111, x1 + y1 + z1);
#undef DSB_CORNER
    float res1 = (1.0f - cx) * v000 + cx * v100;
    res1 = (1.0f - cy) * res1 + cy * ((1.0f - cx) * v010 + cx * v110);
    float res2 = (1.0f - cx) * v001 + cx * v101;
    res2 = (1.0f - cy) * res2 + cy * ((1.0f - cx) * v011 + cx * v111);
    found = true;
    return div_32767((1.0f - cz) * res1 + cz * res2);
}

// castRay, DeviceAgnostic/ITMVisualisationEngine.h:124-196, in three pieces so that a march can be suspended and resumed (by
// another thread, in another kernel) without changing a single operation of it:
//   ray_setup   :124-158  start point, unit direction, ray length at the near / far expected depth
//   ray_step    :161-178  ONE pass of the march loop body
//   ray_finish  :180-195  the two refinement steps and the output
struct RayState {
    float rx, ry, rz, total, dx, dy, dz, totalMax;
#ifdef DSB_RAY_TRACE
    int trips;
#endif
};
#ifdef DSB_RAY_TRACE
// diagnostic build only (make EXTRA=-DDSB_RAY_TRACE): per raycast tile {start ns, end ns, longest ray's loop trips, SM}
__device__ unsigned long long *g_rayTrace = nullptr;
extern "C" int dsb200_debug_ray_trace(unsigned long long *deviceBuffer)
{
    return cudaMemcpyToSymbol(g_rayTrace, &deviceBuffer, sizeof(deviceBuffer)) == cudaSuccess ? 0 : -1;
}
__device__ __forceinline__ unsigned long long ray_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif
enum RayStep { kStepUnit = 0, kStepJump = 1, kStepHit = 2 };     // unit: stepLength == 1.0f exactly; jump: anything else

__device__ __forceinline__ void ray_setup(RayState &s, int x, int y, const RayParams &P, float zmin, float zmax)
{
    Vec4 pc;
    pc.z = zmin;
    pc.x = pc.z * (((float)x - P.invProj[2]) * P.invProj[0]);
    pc.y = pc.z * (((float)y - P.invProj[3]) * P.invProj[1]);
    pc.w = 1.0f;
    s.total = sqrtf(pc.x * pc.x + pc.y * pc.y + pc.z * pc.z) * P.oneOverVoxel;
    Vec4 t = mul(P.invM, pc);
    const float sx = t.x * P.oneOverVoxel, sy = t.y * P.oneOverVoxel, sz = t.z * P.oneOverVoxel;
    pc.z = zmax;
    pc.x = pc.z * (((float)x - P.invProj[2]) * P.invProj[0]);
    pc.y = pc.z * (((float)y - P.invProj[3]) * P.invProj[1]);
    pc.w = 1.0f;
    s.totalMax = sqrtf(pc.x * pc.x + pc.y * pc.y + pc.z * pc.z) * P.oneOverVoxel;
    t = mul(P.invM, pc);
    float dx = t.x * P.oneOverVoxel - sx, dy = t.y * P.oneOverVoxel - sy, dz = t.z * P.oneOverVoxel - sz;
    const float dn = 1.0f / sqrtf(dx * dx + dy * dy + dz * dz);
    s.dx = dx * dn; s.dy = dy * dn; s.dz = dz * dn;
    s.rx = sx; s.ry = sy; s.rz = sz;
}

// one loop body at position (rx, ry, rz): the sample (sdf, as the loop leaves it) and what the loop does next
template <class Tables>
__device__ __forceinline__ RayStep ray_step(float rx, float ry, float rz, const Tables &T, const RayParams &P, float stepScale,
                                            BlockCache &cache, float &sdf, float &stepLength)
{
    bool found;
    sdf = div_32767((float)read_sdf(T, P.bucketNum, P.bucketMask, (int)round_half_away(rx), (int)round_half_away(ry),
                                    (int)round_half_away(rz), found, cache));
    if (!found) { stepLength = (float)kBlock; return kStepJump; }
    if ((sdf <= 0.1f) && (sdf >= -0.5f)) sdf = sdf_trilinear(T, P.bucketNum, P.bucketMask, rx, ry, rz, found, cache);
    if (!(sdf > 0.0f)) return kStepHit;
    stepLength = (sdf * stepScale < 1.0f) ? 1.0f : sdf * stepScale;
    return stepLength == 1.0f ? kStepUnit : kStepJump;
}

template <class Tables>
__device__ __forceinline__ float4 ray_finish(float rx, float ry, float rz, float dx, float dy, float dz, float sdf, const Tables &T,
                                             const RayParams &P, float stepScale, BlockCache &cache)
{
    bool ptFound;
    if (sdf <= 0.0f) {
        float stepLength = sdf * stepScale;
        rx += stepLength * dx; ry += stepLength * dy; rz += stepLength * dz;
        sdf = sdf_trilinear(T, P.bucketNum, P.bucketMask, rx, ry, rz, ptFound, cache);
        stepLength = sdf * stepScale;
        rx += stepLength * dx; ry += stepLength * dy; rz += stepLength * dz;
    } else ptFound = false;
    return make_float4(rx, ry, rz, ptFound ? 1.0f : 0.0f);
}

// the march; returns false when `maxSteps` loop bodies have run and the ray is still marching (state left at the loop top).
// The loop keeps the shape of the reference's (one condition, one break): with a `return` inside it the compiler no longer
// reconverges the lanes of a warp per iteration and the kernel runs 3.5x slower (measured, profiles/README.md r02e).
template <class Tables>
__device__ __forceinline__ bool cast_ray_bounded(float4 &out, RayState &s, const Tables &T, const RayParams &P, int maxSteps)
{
    const int bucketNum = P.bucketNum, bucketMask = P.bucketMask;
    const float stepScale = P.mu * P.oneOverVoxel;
    BlockCache cache = {0x7fffffff, 0x7fffffff, 0x7fffffff, nullptr};
    float rx = s.rx, ry = s.ry, rz = s.rz, totalLength = s.total;
    const float dx = s.dx, dy = s.dy, dz = s.dz, totalLengthMax = s.totalMax;
    float sdf = 1.0f, stepLength;
    bool found, stopped = false;
    int steps = maxSteps;                 // counts down; a negative cap never reaches 0
    while (totalLength < totalLengthMax && steps != 0) {
        steps--;
        sdf = div_32767((float)read_sdf(T, bucketNum, bucketMask, (int)round_half_away(rx), (int)round_half_away(ry),
                                        (int)round_half_away(rz), found, cache));
        if (!found) stepLength = (float)kBlock;
        else {
            if ((sdf <= 0.1f) && (sdf >= -0.5f)) sdf = sdf_trilinear(T, bucketNum, bucketMask, rx, ry, rz, found, cache);
            if (!(sdf > 0.0f)) { stopped = true; break; }
            stepLength = (sdf * stepScale < 1.0f) ? 1.0f : sdf * stepScale;
        }
        rx += stepLength * dx; ry += stepLength * dy; rz += stepLength * dz;
        totalLength += stepLength;
    }
    const bool suspended = !stopped && totalLength < totalLengthMax;     // left the loop on the step cap
    s.rx = rx; s.ry = ry; s.rz = rz; s.total = totalLength;
#ifdef DSB_RAY_TRACE
    s.trips = maxSteps - steps;
#endif
    if (!suspended) out = ray_finish(rx, ry, rz, dx, dy, dz, sdf, T, P, stepScale, cache);
    return !suspended;
}

template <class Tables>
__device__ __forceinline__ bool cast_ray(float4 &out, int x, int y, const Tables &T, const RayParams &P, float zmin, float zmax)
{
    RayState s;
    ray_setup(s, x, y, P, zmin, zmax);
    cast_ray_bounded(out, s, T, P, -1);
    return out.w > 0.0f;
}

// GROUP MARCH: G lanes (2 .. 32) work on ONE ray and evaluate its next G loop bodies TOGETHER.  Lane j assumes that the j bodies
// before it all stepped by S, the step of the last committed body (8 voxels at the start) -- which is what the long rays do:
// tools/ray_trip_stats.py on the C2 street counts, for the rays with >= 64 loop trips, runs of 26 consecutive 8-voxel steps (no
// block) and of 12 consecutive 4-voxel steps (allocated, empty: sdf = 1) on average.  Lane j derives its position by the very
// additions the sequential loop would perform (rx += stepLength * dx with stepLength == S), samples the volume, and the group
// commits the bodies up to and including the first one that did NOT step by exactly S (another step length, the hit, or the end of
// the range); later lanes' samples are discarded.  Lane 0's body is never speculative, so a round always makes progress, and
// every committed body is the sequential loop's body on the sequential loop's operands: the result is bit-identical to cast_ray
// (tests/test_gpu_parity.py::test_two_pass_raycast_is_bit_identical and every frame test, since the fused frame uses it).
// What changes is the dependent chain: up to G bodies per memory round trip instead of one.
template <class Tables, int G>
__device__ __forceinline__ void cast_ray_group(float4 *dst, float rx, float ry, float rz, float total, float dx, float dy, float dz,
                                               float totalMax, const Tables &T, const RayParams &P, unsigned gmask, int j, BlockCache &cache)
{
    const float stepScale = P.mu * P.oneOverVoxel;
    float S = (float)kBlock;           // the group's hypothesis: the next bodies all step by S (first guess: free space)
    while (true) {
        float px = rx, py = ry, pz = rz, t = total;
        const float sx = S * dx, sy = S * dy, sz = S * dz;
        for (int k = 0; k < j; k++) { px += sx; py += sy; pz += sz; t += S; }
        int kind;                      // RayStep, or 3 = the loop condition fails here (no hit)
        float sdf = 1.0f, stepLength = 1.0f;
        if (t < totalMax) kind = ray_step(px, py, pz, T, P, stepScale, cache, sdf, stepLength);
        else kind = 3;
        // lane j's body is the sequential loop's body iff every body before it stepped by exactly S
        const bool asAssumed = (kind == kStepUnit || kind == kStepJump) && stepLength == S;
        const unsigned stop = __ballot_sync(gmask, !asAssumed) & gmask;
        const int first = stop ? (__ffs(stop) - 1) & (G - 1) : G - 1;             // the body the group commits last
        const int fk = __shfl_sync(gmask, kind, first, G);
        if (fk == kStepHit || fk == 3) {
            if (j == first) *dst = ray_finish(px, py, pz, dx, dy, dz, sdf, T, P, stepScale, cache);
            break;
        }
        // a step (as assumed or not): the committed state after it, and the next hypothesis
        px += stepLength * dx; py += stepLength * dy; pz += stepLength * dz; t += stepLength;
        rx = __shfl_sync(gmask, px, first, G); ry = __shfl_sync(gmask, py, first, G); rz = __shfl_sync(gmask, pz, first, G);
        total = __shfl_sync(gmask, t, first, G);
        S = __shfl_sync(gmask, stepLength, first, G);
    }
}

// GenericRaycast (CPU .cpp:296-333; CUDA .cu:807-819).  Ray lengths differ by orders of magnitude across the image (sky
// pixels do not march at all, grazing road pixels cross hundreds of voxels), so a static pixel-tile -> CTA map leaves
// most of the machine idle behind a few long tiles (41 % achieved occupancy in profiles/r01a).  Here every WARP pulls
// the next 8x4 pixel tile from a device ticket; the last warp to run dry re-arms the ticket pair for the next launch.
// `bottomUp` hands the tiles out from the last image row upwards (the long grazing rays of a forward-looking camera sit
// at the bottom: they start first and the short ones fill the tail).
// Two variants were measured on B200 and dropped (profiles/README.md, r01d): a persistent lane-refilling march (idle lanes
// take new pixels while the others keep marching) was 1.6x slower, and issuing the 8 interpolation loads speculatively
// together with the nearest-voxel load (one round trip per step instead of two) was 1.14x slower; a unit-step look-ahead
// (samples of the next 3 one-voxel steps loaded together, committed one by one) was 2x slower at 80 registers.
// A per-tile globaltimer trace (profiles/README.md, r01f) shows why the kernel is hard to shorten: the tickets run out after
// a third of the kernel's duration and the rest is the tail of ~20 tiles of grazing rays that START in the first 12 us and
// run for 100-130 us each (>100 dependent one-voxel steps).
// A two-pass variant (first pass bounded to S steps per ray, unfinished rays queued with their state and finished by a second
// launch with 1 or 4 rays per warp; results identical) was slower too: S = 16 / 32 / 64 with one ray per warp 0.224 / 0.181 /
// 0.168 ms, S = 32 with four rays per warp 0.161 ms, against 0.144 ms -- so the stragglers are not slowed by divergence
// inside their warps; each of their steps simply is several dependent L2 round trips.
// An adaptive one-round-trip step (after an interpolated step, load the tri-linear cell straight away and take the nearest
// voxel, which is always one of its eight corners, out of it; results identical) measured 0.163 ms (55 registers instead of 48).
// Sharded scene: tileStride shards deal the tiles round-robin (tile = ticket * tileStride + tileFirst) and `out` may be the
// root shard's image (peer memory: the 16-byte stores of a tile row are one 128-byte NVLink write).
// Rays still marching after `maxSteps` loop bodies are SUSPENDED: their state goes to a queue and k_raycast_finish completes
// them (the tail described above: a few hundred grazing rays creeping one voxel per step for > 100 dependent steps).

template <class Tables>
__global__ void __launch_bounds__(128) k_raycast(RayParams P, const __grid_constant__ Tables T,
                                                 const float2 *__restrict__ range, float4 *out, int *ticket, int tileFirst,
                                                 int tileStride, int bottomUp, RayQueue q, int maxSteps, const int *order)
{
    const int lane = threadIdx.x & 31;
    const int tilesX = (P.W + 7) >> 3, tilesY = (P.H + 3) >> 2, nTiles = tilesX * tilesY;
    // k_tile_order's lists: class 0 (the longest depth spans) .. class 3
    int c0 = 0, c1 = 0, c2 = 0;
    if (order) { c0 = order[0]; c1 = order[1]; c2 = order[2]; }
    while (true) {
        int tile = 0;
        if (lane == 0) tile = atomicAdd(&ticket[0], 1);
        tile = __shfl_sync(0xffffffffu, tile, 0) * tileStride + tileFirst;
        if (tile >= nTiles) break;
        if (order) {
            int c = 0, u = tile;
            if (u >= c0) { u -= c0; c = 1; if (u >= c1) { u -= c1; c = 2; if (u >= c2) { u -= c2; c = 3; } } }
            tile = __ldg(order + kOrderClasses + (size_t)c * nTiles + u);
        } else if (bottomUp) tile = nTiles - 1 - tile;
        const int ty = tile / tilesX, tx = tile - ty * tilesX;
        const int x = tx * 8 + (lane & 7), y = ty * 4 + (lane >> 3);
        bool suspended = false;
        RayState s;
#ifdef DSB_RAY_TRACE
        const unsigned long long traceT0 = ray_now();
        s.trips = 0;
#endif
        if (x < P.W && y < P.H) {
            const int loc2 = (int)floorf((float)x / (float)kSubsample) + (int)floorf((float)y / (float)kSubsample) * P.W;
            const float2 mm = __ldg(range + loc2);
            float4 r;
            ray_setup(s, x, y, P, mm.x, mm.y);
            if (cast_ray_bounded(r, s, T, P, maxSteps)) out[x + y * P.W] = r;
            else suspended = true;
        }
#ifdef DSB_RAY_TRACE
        {
            const int tr = __reduce_max_sync(0xffffffffu, s.trips);
            if (lane == 0 && g_rayTrace) {
                unsigned smid; asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
                unsigned long long *w = g_rayTrace + (size_t)tile * 4;
                w[0] = traceT0; w[1] = ray_now(); w[2] = (unsigned long long)tr; w[3] = smid;
            }
        }
#endif
        const unsigned m = __ballot_sync(0xffffffffu, suspended);
        if (m) {
            int base = 0;
            if (lane == 0) base = atomicAdd(&q.ctl[0], __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (suspended) {
                const int i = base + __popc(m & ((1u << lane) - 1u));      // capacity = W * H: cannot overflow
                q.pixel[i] = x + y * P.W;
                q.posLen[i] = make_float4(s.rx, s.ry, s.rz, s.total);
                q.dirMax[i] = make_float4(s.dx, s.dy, s.dz, s.totalMax);
            }
        }
    }
    if (lane == 0) {
        const int nWarps = gridDim.x * (blockDim.x >> 5);
        if (atomicAdd(&ticket[1], 1) == nWarps - 1) { ticket[1] = 0; ticket[0] = 0; }   // all tickets were handed out: re-arm
    }
}

// Finishing pass of the two-pass schedule (ray_steps > 0; OFF by default): the rays the bounded first pass suspended are marched
// by groups of G lanes (cast_ray_group).  Measured on B200 (profiles/README.md r02f, r02z): with the "repeat the last step"
// hypothesis a round commits ~4 bodies but costs ~2.2 us against ~1.1 us for a sequential body, and the first pass keeps a tail of
// its own: 0.141-0.151 ms against 0.139 ms for the single pass inside a C2 frame (the unit-step hypothesis of r02f: 0.229 ms).
template <class Tables, int G>
__global__ void __launch_bounds__(128) k_raycast_finish(RayParams P, const __grid_constant__ Tables T, float4 *out, RayQueue q)
{
    const int lane = threadIdx.x & 31, j = lane & (G - 1);
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (lane & ~(G - 1)));
    const int n = q.ctl[0];
    BlockCache cache = {0x7fffffff, 0x7fffffff, 0x7fffffff, nullptr};
    while (true) {
        int i = 0;
        if (j == 0) i = atomicAdd(&q.ctl[1], 1);
        i = __shfl_sync(gmask, i, 0, G);
        if (i >= n) break;
        const float4 a = q.posLen[i], b = q.dirMax[i];
        float rx = a.x, ry = a.y, rz = a.z, total = a.w;
        const float dx = b.x, dy = b.y, dz = b.z, totalMax = b.w;
        cast_ray_group<Tables, G>(out + q.pixel[i], rx, ry, rz, total, dx, dy, dz, totalMax, T, P, gmask, j, cache);
    }
    __syncwarp();
    if (lane == 0) {
        const int nWarps = gridDim.x * (blockDim.x >> 5);
        if (atomicAdd(&q.ctl[2], 1) == nWarps - 1) { q.ctl[0] = 0; q.ctl[1] = 0; q.ctl[2] = 0; }   // queue drained: re-arm for the next raycast
    }
}

// ---------------------------------------------------------------- ICP maps (processPixelICP<true>, DA/...:391-439)
__device__ __forceinline__ void normal_from_image(bool &found, int x, int y, const float4 *__restrict__ pts, const RayParams &P,
                                                  float n[3], float &angle)
{   // computeNormalAndAngle<true>, DA/ITMVisualisationEngine.h:243-312
    if (!found) return;
    if (y <= 2 || y >= P.H - 3 || x <= 2 || x >= P.W - 3) { found = false; return; }
    float4 xp = __ldg(pts + (x + 2) + y * P.W), yp = __ldg(pts + x + (y + 2) * P.W);
    float4 xm = __ldg(pts + (x - 2) + y * P.W), ym = __ldg(pts + x + (y - 2) * P.W);
    float dxx = 0, dxy = 0, dxz = 0, dyx = 0, dyy = 0, dyz = 0;
    bool doPlus1 = false;
    if (xp.w <= 0 || yp.w <= 0 || xm.w <= 0 || ym.w <= 0) doPlus1 = true;
    else {
        dxx = xp.x - xm.x; dxy = xp.y - xm.y; dxz = xp.z - xm.z;
        dyx = yp.x - ym.x; dyy = yp.y - ym.y; dyz = yp.z - ym.z;
        float a = dxx * dxx + dxy * dxy + dxz * dxz, b = dyx * dyx + dyy * dyy + dyz * dyz;
        float lengthDiff = (a < b) ? b : a;
        if (lengthDiff * P.voxelSize * P.voxelSize > (0.15f * 0.15f)) doPlus1 = true;
    }
    if (doPlus1) {
        xp = __ldg(pts + (x + 1) + y * P.W); yp = __ldg(pts + x + (y + 1) * P.W);
        xm = __ldg(pts + (x - 1) + y * P.W); ym = __ldg(pts + x + (y - 1) * P.W);
        dxx = xp.x - xm.x; dxy = xp.y - xm.y; dxz = xp.z - xm.z;
        dyx = yp.x - ym.x; dyy = yp.y - ym.y; dyz = yp.z - ym.z;
        if (xp.w <= 0 || yp.w <= 0 || xm.w <= 0 || ym.w <= 0) { found = false; return; }
    }
    n[0] = -(dxy * dyz - dxz * dyy);
    n[1] = -(dxz * dyx - dxx * dyz);
    n[2] = -(dxx * dyy - dxy * dyx);
    float sc = 1.0f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    n[0] *= sc; n[1] *= sc; n[2] *= sc;
    angle = n[0] * P.light[0] + n[1] * P.light[1] + n[2] * P.light[2];
    if (!(angle > 0.0f)) found = false;
}

__global__ void __launch_bounds__(256) k_icp_maps(RayParams P, const float4 *__restrict__ pts, float4 *points, float4 *normals,
                                                  uchar4 *rendering)
{
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= P.W || y >= P.H) return;
    const int loc = x + y * P.W;
    const float4 pt = __ldg(pts + loc);
    bool found = pt.w > 0.0f;
    float n[3], angle = 0.0f;
    normal_from_image(found, x, y, pts, P, n, angle);
    if (found) {
        unsigned char g = (unsigned char)((0.8f * angle + 0.2f) * 255.0f);
        rendering[loc] = make_uchar4(g, g, g, g);
        points[loc] = make_float4(pt.x * P.voxelSize, pt.y * P.voxelSize, pt.z * P.voxelSize, 1.0f);
        normals[loc] = make_float4(n[0], n[1], n[2], 0.0f);
    } else {
        points[loc] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        normals[loc] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        rendering[loc] = make_uchar4(0, 0, 0, 0);
    }
}


// ---------------------------------------------------------------- sharded scene: nearest-hit composite
// Every shard raycasts the blocks it owns; the surface a camera ray sees is the hit closest to the camera among the
// shards.  One thread per pixel loads the n candidate points (16 B each; peers' images are read in place over NVLink
// when the pointers are IPC mappings) and keeps the one with the smallest squared distance to the camera centre
// (ties: lowest shard index).  Replaces an all-reduce(min) on packed keys + a second collective for the winner's point.
struct PeerImages { const float4 *p[kMaxShards]; };

__global__ void __launch_bounds__(256) k_composite_nearest(PeerImages imgs, int n, float cx, float cy, float cz, int begin,
                                                           int end, float4 *out)
{
    for (int i = begin + blockIdx.x * blockDim.x + threadIdx.x; i < end; i += gridDim.x * blockDim.x) {
        float4 best = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float bestD = 0.0f;
        bool have = false;
#pragma unroll 4
        for (int r = 0; r < n; r++) {
            const float4 q = __ldcg(imgs.p[r] + i);        // produced by another kernel / another GPU: bypass L1
            if (q.w > 0.0f) {
                const float dx = q.x - cx, dy = q.y - cy, dz = q.z - cz;
                const float d = dx * dx + dy * dy + dz * dz;
                if (!have || d < bestD) { best = q; bestD = d; have = true; }
            }
        }
        out[i] = best;
    }
}

void launch_composite_nearest(cudaStream_t st, int n, const float *const *results, float cx, float cy, float cz, int W,
                              int rowBegin, int rowEnd, float *out)
{
    PeerImages imgs;
    for (int r = 0; r < kMaxShards; r++) imgs.p[r] = reinterpret_cast<const float4 *>(results[r < n ? r : 0]);
    const int begin = rowBegin * W, end = rowEnd * W;
    int grid = div_up(end - begin > 0 ? end - begin : 1, 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    DSB_LAUNCH(k_composite_nearest, grid, 256, 0, st, imgs, n, cx, cy, cz, begin, end, reinterpret_cast<float4 *>(out));
}

// ---------------------------------------------------------------- ForwardRender (approximate raycast)
// CPU/ITMVisualisationEngine_CPU.cpp:464-540, CUDA .cu:486-547: re-project the previous raycast result into the new
// pose, raycast only the holes, shade.  The reference scatters with plain stores (CPU: serial, last pixel in scan order
// wins; CUDA: a race); here the winner is made explicit -- atomicMax of the source pixel index per target pixel, then
// a gather -- which is the serial result.  Points that project to NaN (z == 0) are skipped (the reference indexes with
// an undefined int there).
__global__ void __launch_bounds__(256) k_fwd_project(RayParams P, float4 projParams, const float4 *__restrict__ pts, int *owner)
{
    const int loc = blockIdx.x * blockDim.x + threadIdx.x;
    if (loc >= P.W * P.H) return;
    const float4 px = __ldg(pts + loc);
    Vec4 p = {px.x * P.voxelSize, px.y * P.voxelSize, px.z * P.voxelSize, 1.0f};   // forwardProjectPixel: pixel.w forced to 1
    p = mul(P.M, p);
    const float u = projParams.x * p.x / p.z + projParams.z, v = projParams.y * p.y / p.z + projParams.w;
    if (!(u >= 0 && u <= (float)(P.W - 1) && v >= 0 && v <= (float)(P.H - 1))) return;
    atomicMax(owner + (int)(u + 0.5f) + (int)(v + 0.5f) * P.W, loc + 1);
}

__global__ void __launch_bounds__(128) k_fwd_fill(RayParams P, const dsb200_voxel *__restrict__ voxels,
                                                  const dsb200_hash_entry *__restrict__ hash, const float2 *__restrict__ range,
                                                  const float4 *__restrict__ pts, const float *__restrict__ depth,
                                                  const int *__restrict__ owner, float4 *fwd, int *noMissing)
{
    const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 8 + (threadIdx.x >> 4);
    if (x >= P.W || y >= P.H) return;
    const int loc = x + y * P.W;
    const int o = owner[loc];
    float4 f = (o > 0) ? __ldg(pts + (o - 1)) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const int loc2 = (int)floorf((float)x / (float)kSubsample) + (int)floorf((float)y / (float)kSubsample) * P.W;
    const float2 mm = __ldg(range + loc2);
    // findMissingPoints_device, CUDA .cu:873-905 / CPU .cpp:497-515
    if ((f.w <= 0) && ((f.x == 0 && f.y == 0 && f.z == 0) || (__ldg(depth + loc) >= 0)) && (mm.x < mm.y)) {
        atomicAdd(noMissing, 1);
        const SceneTables T = scene_tables(voxels, hash);
        cast_ray(f, x, y, T, P, mm.x, mm.y);
    }
    fwd[loc] = f;
}

// processPixelForwardRender<true>, DeviceAgnostic/ITMVisualisationEngine.h:441-456
__global__ void __launch_bounds__(256) k_fwd_shade(RayParams P, const float4 *__restrict__ fwd, uchar4 *rendering)
{
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= P.W || y >= P.H) return;
    const int loc = x + y * P.W;
    bool found = __ldg(fwd + loc).w > 0.0f;
    float n[3], angle = 0.0f;
    normal_from_image(found, x, y, fwd, P, n, angle);
    const unsigned char g = found ? (unsigned char)((0.8f * angle + 0.2f) * 255.0f) : 0;
    rendering[loc] = make_uchar4(g, g, g, g);
}

void launch_forward_render(cudaStream_t st, const RayParams &P, const float *proj, const dsb200_voxel *voxels,
                           const dsb200_hash_entry *hash, const float *rangeImage, const float *raycastResult,
                           const float *depth, float *fwd, int *owner, int *noMissing, uint8_t *rendering)
{
    const int n = P.W * P.H;
    cudaMemsetAsync(owner, 0, (size_t)n * sizeof(int), st);
    cudaMemsetAsync(noMissing, 0, sizeof(int), st);
    DSB_LAUNCH(k_fwd_project, div_up(n, 256), 256, 0, st, P, make_float4(proj[0], proj[1], proj[2], proj[3]),
               reinterpret_cast<const float4 *>(raycastResult), owner);
    DSB_LAUNCH(k_fwd_fill, dim3(div_up(P.W, 16), div_up(P.H, 8)), 128, 0, st, P, voxels, hash, reinterpret_cast<const float2 *>(rangeImage),
               reinterpret_cast<const float4 *>(raycastResult), depth, owner, reinterpret_cast<float4 *>(fwd), noMissing);
    DSB_LAUNCH(k_fwd_shade, dim3(div_up(P.W, 32), div_up(P.H, 8)), 256, 0, st, P, reinterpret_cast<const float4 *>(fwd),
               reinterpret_cast<uchar4 *>(rendering));
}

// ---------------------------------------------------------------- free-view shading (RenderImage)
#define DSB_RV(DX, DY, DZ) ((float)read_sdf(T, bucketNum, bucketMask, X + (DX), Y + (DY), Z + (DZ), f, cache))

// computeSingleNormalFromSDF, DeviceAgnostic/ITMRepresentationAccess.h:314-427
__device__ void normal_from_sdf(const dsb200_voxel *__restrict__ voxels, const dsb200_hash_entry *__restrict__ hash,
                                int bucketNum, int bucketMask, float px, float py, float pz, float ret[3])
{
    const SceneTables T = scene_tables(voxels, hash);
    bool f; BlockCache cache = {0x7fffffff, 0x7fffffff, 0x7fffffff, nullptr};
    float flx = floorf(px), fly = floorf(py), flz = floorf(pz);
    float cx = px - flx, cy = py - fly, cz = pz - flz;
    int X = (int)flx, Y = (int)fly, Z = (int)flz;
    float nx = 1.0f - cx, ny = 1.0f - cy, nz = 1.0f - cz;
    float fr0 = DSB_RV(0, 0, 0), fr1 = DSB_RV(1, 0, 0), fr2 = DSB_RV(0, 1, 0), fr3 = DSB_RV(1, 1, 0);
    float bk0 = DSB_RV(0, 0, 1), bk1 = DSB_RV(1, 0, 1), bk2 = DSB_RV(0, 1, 1), bk3 = DSB_RV(1, 1, 1);
    float t0, t1, t2, t3, p1, p2, v1;
    p1 = fr0 * ny * nz + fr2 * cy * nz + bk0 * ny * cz + bk2 * cy * cz;
    t0 = DSB_RV(-1, 0, 0); t1 = DSB_RV(-1, 1, 0); t2 = DSB_RV(-1, 0, 1); t3 = DSB_RV(-1, 1, 1);
    p2 = t0 * ny * nz + t1 * cy * nz + t2 * ny * cz + t3 * cy * cz;
    v1 = p1 * cx + p2 * nx;
    p1 = fr1 * ny * nz + fr3 * cy * nz + bk1 * ny * cz + bk3 * cy * cz;
    t0 = DSB_RV(2, 0, 0); t1 = DSB_RV(2, 1, 0); t2 = DSB_RV(2, 0, 1); t3 = DSB_RV(2, 1, 1);
    p2 = t0 * ny * nz + t1 * cy * nz + t2 * ny * cz + t3 * cy * cz;
    ret[0] = (p1 * nx + p2 * cx - v1) / 32767.0f;
    p1 = fr0 * nx * nz + fr1 * cx * nz + bk0 * nx * cz + bk1 * cx * cz;
    t0 = DSB_RV(0, -1, 0); t1 = DSB_RV(1, -1, 0); t2 = DSB_RV(0, -1, 1); t3 = DSB_RV(1, -1, 1);
    p2 = t0 * nx * nz + t1 * cx * nz + t2 * nx * cz + t3 * cx * cz;
    v1 = p1 * cy + p2 * ny;
    p1 = fr2 * nx * nz + fr3 * cx * nz + bk2 * nx * cz + bk3 * cx * cz;
    t0 = DSB_RV(0, 2, 0); t1 = DSB_RV(1, 2, 0); t2 = DSB_RV(0, 2, 1); t3 = DSB_RV(1, 2, 1);
    p2 = t0 * nx * nz + t1 * cx * nz + t2 * nx * cz + t3 * cx * cz;
    ret[1] = (p1 * ny + p2 * cy - v1) / 32767.0f;
    p1 = fr0 * nx * ny + fr1 * cx * ny + fr2 * nx * cy + fr3 * cx * cy;
    t0 = DSB_RV(0, 0, -1); t1 = DSB_RV(1, 0, -1); t2 = DSB_RV(0, 1, -1); t3 = DSB_RV(1, 1, -1);
    p2 = t0 * nx * ny + t1 * cx * ny + t2 * nx * cy + t3 * cx * cy;
    v1 = p1 * cz + p2 * nz;
    p1 = bk0 * nx * ny + bk1 * cx * ny + bk2 * nx * cy + bk3 * cx * cy;
    t0 = DSB_RV(0, 0, 2); t1 = DSB_RV(1, 0, 2); t2 = DSB_RV(0, 1, 2); t3 = DSB_RV(1, 1, 2);
    p2 = t0 * nx * ny + t1 * cx * ny + t2 * nx * cy + t3 * cx * cy;
    ret[2] = (p1 * nz + p2 * cz - v1) / 32767.0f;
}

// full voxel read for colour rendering
__device__ __forceinline__ uchar4 read_clr(const dsb200_voxel *__restrict__ voxels, const dsb200_hash_entry *__restrict__ hash,
                                           int bucketNum, int bucketMask, int px, int py, int pz, BlockCache &c)
{
    bool found;
    const SceneTables T = scene_tables(voxels, hash);
    read_sdf(T, bucketNum, bucketMask, px, py, pz, found, c);   // positions the cache
    if (!found) return make_uchar4(0, 0, 0, 0);
    const int lin = (px & 7) + ((py & 7) << 3) + ((pz & 7) << 6);
    const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(c.blk) + lin);
    return make_uchar4(raw.x >> 24, raw.y & 0xff, (raw.y >> 8) & 0xff, 0);
}

// RenderImage shading pass: CPU/ITMVisualisationEngine_CPU.cpp:339-405, CUDA .cu:405-435 (depth falls through to grey)
__global__ void __launch_bounds__(128) k_shade(RayParams P, int type, const dsb200_voxel *__restrict__ voxels,
                                               const dsb200_hash_entry *__restrict__ hash, const float4 *__restrict__ pts,
                                               uchar4 *out, float *outDepth)
{
    const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 8 + (threadIdx.x >> 4);
    if (x >= P.W || y >= P.H) return;
    const int loc = x + y * P.W;
    const float4 pt = __ldg(pts + loc);
    bool found = pt.w > 0;
    if (type == DSB200_RENDER_DEPTH_MAP && outDepth) {
        // processPixelColourDepth / drawPixelDepth, DeviceAgnostic/ITMVisualisationEngine.h:340-357,527-540
        float d = 0.0f;
        if (found) {
            Vec4 ph = {pt.x * P.voxelSize, pt.y * P.voxelSize, pt.z * P.voxelSize, 1.0f};
            Vec4 pc = mul(P.M, ph);
            d = pc.z / pc.w;
        }
        outDepth[loc] = d;
    }
    float n[3] = {0, 0, 0}, angle = 0.0f;
    if (found) {   // computeNormalAndAngle<TVoxel,TIndex>, DA/...:220-234
        normal_from_sdf(voxels, hash, P.bucketNum, P.bucketMask, pt.x, pt.y, pt.z, n);
        float sc = 1.0f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        n[0] *= sc; n[1] *= sc; n[2] *= sc;
        angle = n[0] * P.light[0] + n[1] * P.light[1] + n[2] * P.light[2];
        if (!(angle > 0.0f)) found = false;
    }
    if (!found) { out[loc] = make_uchar4(0, 0, 0, 0); return; }
    if (type == DSB200_RENDER_COLOUR_FROM_VOLUME) {
        // readFromSDF_color4u_interpolated, DeviceAgnostic/ITMRepresentationAccess.h:278-312
        BlockCache c = {0x7fffffff, 0x7fffffff, 0x7fffffff, nullptr};
        float flx = floorf(pt.x), fly = floorf(pt.y), flz = floorf(pt.z);
        float cx = pt.x - flx, cy = pt.y - fly, cz = pt.z - flz;
        int X = (int)flx, Y = (int)fly, Z = (int)flz;
        float r0 = 0.0f, r1 = 0.0f, r2 = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int ox = i & 1, oy = (i >> 1) & 1, oz = i >> 2;
            uchar4 v = read_clr(voxels, hash, P.bucketNum, P.bucketMask, X + ox, Y + oy, Z + oz, c);
            float w = (ox ? cx : (1.0f - cx)) * (oy ? cy : (1.0f - cy)) * (oz ? cz : (1.0f - cz));
            r0 += w * (float)v.x; r1 += w * (float)v.y; r2 += w * (float)v.z;
        }
        out[loc] = make_uchar4((unsigned char)((r0 / 255.0f) * 255.0f), (unsigned char)((r1 / 255.0f) * 255.0f),
                               (unsigned char)((r2 / 255.0f) * 255.0f), 255);
    } else if (type == DSB200_RENDER_COLOUR_FROM_NORMAL) {
        uchar4 o = out[loc];   // drawPixelNormal leaves alpha untouched
        o.x = (unsigned char)((0.3f + (-n[0] + 1.0f) * 0.35f) * 255.0f);
        o.y = (unsigned char)((0.3f + (-n[1] + 1.0f) * 0.35f) * 255.0f);
        o.z = (unsigned char)((0.3f + (-n[2] + 1.0f) * 0.35f) * 255.0f);
        out[loc] = o;
    } else {
        unsigned char g = (unsigned char)((0.8f * angle + 0.2f) * 255.0f);
        out[loc] = make_uchar4(g, g, g, g);
    }
}

// ---------------------------------------------------------------- view building
// convertDepthAffineToFloat (DeviceAgnostic/ITMViewBuilder.h:22-28) + CvToItm's BGR -> RGBA (DS/InfiniTamDriver.cpp:84-104)
// in one pass: 4 pixels per thread (one 8-byte depth load, three 4-byte colour loads, 16-byte stores).
__global__ void __launch_bounds__(256) k_view_from_raw(const short *__restrict__ raw, const uint8_t *__restrict__ bgr, float scale,
                                                       float offset, float *depth, uchar4 *rgb, int n)
{
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q * 4 < n; q += gridDim.x * blockDim.x) {
        const int i = q * 4;
        if (i + 3 < n) {
            const short4 d = *reinterpret_cast<const short4 *>(raw + i);
            float4 o;
            o.x = ((d.x <= 0) || (d.x > 32000)) ? -1.0f : (float)d.x * scale + offset;
            o.y = ((d.y <= 0) || (d.y > 32000)) ? -1.0f : (float)d.y * scale + offset;
            o.z = ((d.z <= 0) || (d.z > 32000)) ? -1.0f : (float)d.z * scale + offset;
            o.w = ((d.w <= 0) || (d.w > 32000)) ? -1.0f : (float)d.w * scale + offset;
            *reinterpret_cast<float4 *>(depth + i) = o;
            if (bgr) {
                const unsigned *c = reinterpret_cast<const unsigned *>(bgr + (size_t)i * 3);     // 12 bytes = 4 pixels, 4-byte aligned
                const unsigned a = __ldg(c), b = __ldg(c + 1), e = __ldg(c + 2);
                // bytes: a = B0 G0 R0 B1 | b = G1 R1 B2 G2 | e = R2 B3 G3 R3
                uint4 px;
                px.x = ((a >> 16) & 0xffu) | (a & 0xff00u) | ((a & 0xffu) << 16) | 0xff000000u;
                px.y = ((b >> 8) & 0xffu) | ((b & 0xffu) << 8) | ((a >> 24) << 16) | 0xff000000u;
                px.z = (e & 0xffu) | ((b >> 24) << 8) | (((b >> 16) & 0xffu) << 16) | 0xff000000u;
                px.w = (e >> 24) | (((e >> 16) & 0xffu) << 8) | (((e >> 8) & 0xffu) << 16) | 0xff000000u;
                *reinterpret_cast<uint4 *>(rgb + i) = px;
            }
        } else {
            for (int k = i; k < n; k++) {
                const short d = raw[k];
                depth[k] = ((d <= 0) || (d > 32000)) ? -1.0f : (float)d * scale + offset;
                if (bgr) rgb[k] = make_uchar4(bgr[3 * k + 2], bgr[3 * k + 1], bgr[3 * k], 255);
            }
        }
    }
}

void launch_view_from_raw(cudaStream_t st, const int16_t *raw, const uint8_t *bgr, float scale, float offset, float *depth,
                          uint8_t *rgb, int n)
{
    int grid = div_up(div_up(n, 4), 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    DSB_LAUNCH(k_view_from_raw, grid, 256, 0, st, raw, bgr, scale, offset, depth, reinterpret_cast<uchar4 *>(rgb), n);
}

// ---------------------------------------------------------------- host launchers
void launch_find_visible(cudaStream_t st, const VisParams &P, const FrameState &fs, const dsb200_hash_entry *hash,
                         int *visibleIds, int visibleCapacity, int total)
{
    DSB_LAUNCH(k_reset_scan, 8, 256, 0, st, fs, fs.scanTiles, fs.scanTiles, 1);
    DSB_LAUNCH(k_find_visible, fs.scanTiles, 256, 0, st, P, fs, hash, visibleIds, visibleCapacity, total);
}

void launch_count_visible(cudaStream_t st, const dsb200_hash_entry *hash, const int *ids, const int *count, int minId,
                          int maxId, int *out)
{
    cudaMemsetAsync(out, 0, sizeof(int), st);
    DSB_LAUNCH(k_count_visible, kNumSMs, 256, 0, st, hash, ids, count, minId, maxId, out);
}

void launch_expected_depths(cudaStream_t st, const ProjParams &P, const dsb200_hash_entry *hash, const int *ids,
                            const int *count, float *rangeImage, int *tileOrder, Counters *counters)
{
    const bool lpt = tileOrder && tunables().ray_lpt > 0;
    DSB_LAUNCH(k_range_init, kNumSMs * 4, 256, 0, st, reinterpret_cast<float2 *>(rangeImage), P.W * P.H, counters, lpt ? tileOrder : nullptr);
    DSB_LAUNCH(k_project_fill, kNumSMs * 4, 256, 0, st, P, hash, ids, count, rangeImage, counters);
    DSB_LAUNCH(k_range_cap_fallback, 1, 1024, 0, st, P, hash, ids, count, rangeImage, counters);
    if (lpt) {
        const int n = div_up(P.W, 8) * div_up(P.H, 4);
        // a span of ray_lpt 8-voxel steps, in metres (the range image holds camera-space depths)
        DSB_LAUNCH(k_tile_order, div_up(n, 256), 256, 0, st, reinterpret_cast<const float2 *>(rangeImage), P.W, P.H,
                   (float)tunables().ray_lpt * 8.0f * P.voxelSize, tileOrder);
    }
}

// tuning knobs (kernels.cuh: Tunables)
static int env_int(const char *name, int dflt) { const char *v = getenv(name); return v ? atoi(v) : dflt; }
Tunables &tunables()
{
    static Tunables t = {env_int("DSB200_RAY_STEPS", 0), env_int("DSB200_RAY_GROUP", 8), env_int("DSB200_RAY_CTAS", 12),
                         env_int("DSB200_RAY_FINISH_CTAS", 4), env_int("DSB200_RAY_BOTTOM_UP", 0), env_int("DSB200_FUSE_KERNEL", 4), env_int("DSB200_FUSE_CTAS", 10),
                         env_int("DSB200_FUSE_MINB", 5), env_int("DSB200_DEEP_LOOKAHEAD", 1), env_int("DSB200_RAY_LPT", 48)};
    return t;
}
bool set_tunable(const char *name, int value)
{
    Tunables &t = tunables();
    const struct { const char *n; int *p; } table[] = {{"ray_steps", &t.ray_steps}, {"ray_group", &t.ray_group}, {"ray_ctas", &t.ray_ctas},
        {"ray_finish_ctas", &t.ray_finish_ctas}, {"ray_bottom_up", &t.ray_bottom_up}, {"fuse_kernel", &t.fuse_kernel}, {"fuse_ctas", &t.fuse_ctas}, {"fuse_minb", &t.fuse_minb}, {"deep_lookahead", &t.deep_lookahead}, {"ray_lpt", &t.ray_lpt}};
    for (const auto &e : table) if (!strcmp(e.n, name)) { *e.p = value; return true; }
    return false;
}
static int ray_bottom_up() { return tunables().ray_bottom_up != 0; }
static int ray_ctas_per_sm() { const int v = tunables().ray_ctas; return v < 1 ? 1 : v; }

// bounded first pass + finishing pass (k_raycast_finish); DSB200_RAY_STEPS = 0 (the default: the two-pass schedule measured
// SLOWER inside a frame, 0.229-0.276 ms against 0.158 ms at C2, profiles/README.md r02f) runs the single unbounded pass,
// DSB200_RAY_GROUP = lanes per suspended ray in the finishing pass (8, 16 or 32)
static int ray_max_steps() { return tunables().ray_steps; }
static int ray_group() { return tunables().ray_group; }

template <class Tables>
static void launch_raycast_two_pass(cudaStream_t st, const RayParams &P, const Tables &T, const float2 *range, float4 *out,
                                    int *ticketPair, int tileFirst, int tileStride, int grid, const RayQueue *queue, const int *order = nullptr)
{
    RayQueue q{};
    int maxSteps = -1;
    if (queue && queue->ctl && queue->capacity >= P.W * P.H && ray_max_steps() > 0) { q = *queue; maxSteps = ray_max_steps(); }
    DSB_LAUNCH(k_raycast<Tables>, grid, 128, 0, st, P, T, range, out, ticketPair, tileFirst, tileStride, ray_bottom_up(), q, maxSteps,
               (tileStride == 1 && tunables().ray_lpt > 0) ? order : nullptr);
    if (maxSteps < 0) return;
    const int g2 = kNumSMs * (tunables().ray_finish_ctas < 1 ? 1 : tunables().ray_finish_ctas);
    switch (ray_group()) {
    case 32: DSB_LAUNCH((k_raycast_finish<Tables, 32>), g2, 128, 0, st, P, T, out, q); break;
    case 16: DSB_LAUNCH((k_raycast_finish<Tables, 16>), g2, 128, 0, st, P, T, out, q); break;
    case 4: DSB_LAUNCH((k_raycast_finish<Tables, 4>), g2, 128, 0, st, P, T, out, q); break;
    case 2: DSB_LAUNCH((k_raycast_finish<Tables, 2>), g2, 128, 0, st, P, T, out, q); break;
    default: DSB_LAUNCH((k_raycast_finish<Tables, 8>), g2, 128, 0, st, P, T, out, q); break;
    }
}

void launch_raycast(cudaStream_t st, const RayParams &P, const dsb200_voxel *voxels, const dsb200_hash_entry *hash,
                    const float *rangeImage, float *raycastResult, int *ticketPair, const short *sdfShadow, const RayQueue *queue,
                    const int *tileOrder)
{
    const int nTiles = div_up(P.W, 8) * div_up(P.H, 4);
    int grid = div_up(nTiles, 4);
    if (grid > kNumSMs * ray_ctas_per_sm()) grid = kNumSMs * ray_ctas_per_sm();
    if (sdfShadow) {
        ShadowTables T; T.sdf = sdfShadow; T.hash = hash;
        launch_raycast_two_pass(st, P, T, reinterpret_cast<const float2 *>(rangeImage), reinterpret_cast<float4 *>(raycastResult),
                                ticketPair, 0, 1, grid, queue, tileOrder);
        return;
    }
    const SceneTables T = scene_tables(voxels, hash);
    launch_raycast_two_pass(st, P, T, reinterpret_cast<const float2 *>(rangeImage), reinterpret_cast<float4 *>(raycastResult),
                            ticketPair, 0, 1, grid, queue, tileOrder);
}

// Expected-depth image of the whole sharded scene = element-wise (min of mins, max of maxes) of the shards' images: what
// CreateExpectedDepths computes over the union of the visible lists (min / max are order-independent).
__global__ void __launch_bounds__(256) k_range_merge(PeerImages imgs, int n, int count, float2 *out)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        float2 m = make_float2(kFarAway, kVeryClose);
#pragma unroll 4
        for (int r = 0; r < n; r++) {
            const float2 q = __ldcg(reinterpret_cast<const float2 *>(imgs.p[r]) + i);
            m.x = fminf(m.x, q.x); m.y = fmaxf(m.y, q.y);
        }
        out[i] = m;
    }
}

void launch_raycast_sharded(cudaStream_t st, const RayParams &P, int n, int rank, int brickLog2,
                            const dsb200_hash_entry *const *hash, const dsb200_voxel *const *voxels,
                            const float *const *rangeImages, float *mergedRange, float *raycastResult, int *ticketPair,
                            const RayQueue *queue)
{
    // only rows [0, H/8] of the full-stride range image are ever read
    PeerImages imgs;
    for (int r = 0; r < kMaxShards; r++) imgs.p[r] = reinterpret_cast<const float4 *>(rangeImages[r < n ? r : 0]);
    const int count = P.W * (P.H / kSubsample + 1);
    DSB_LAUNCH(k_range_merge, div_up(count, 256) < kNumSMs * 4 ? div_up(count, 256) : kNumSMs * 4, 256, 0, st, imgs, n, count,
               reinterpret_cast<float2 *>(mergedRange));
    ShardTables T;
    T.n = n; T.brickLog2 = brickLog2;
    for (int r = 0; r < kMaxShards; r++) { T.hash[r] = hash[r < n ? r : 0]; T.voxels[r] = voxels[r < n ? r : 0]; }
    const int nTiles = div_up(div_up(P.W, 8) * div_up(P.H, 4), n);
    int grid = div_up(nTiles, 4);
    if (grid > kNumSMs * ray_ctas_per_sm()) grid = kNumSMs * ray_ctas_per_sm();
    launch_raycast_two_pass(st, P, T, reinterpret_cast<const float2 *>(mergedRange), reinterpret_cast<float4 *>(raycastResult),
                            ticketPair, rank, n, grid, queue);
}

void launch_icp_maps(cudaStream_t st, const RayParams &P, const float *raycastResult, float *points, float *normals,
                     uint8_t *rendering)
{
    dim3 grid(div_up(P.W, 32), div_up(P.H, 8));
    DSB_LAUNCH(k_icp_maps, grid, 256, 0, st, P, reinterpret_cast<const float4 *>(raycastResult),
               reinterpret_cast<float4 *>(points), reinterpret_cast<float4 *>(normals), reinterpret_cast<uchar4 *>(rendering));
}

void launch_shade(cudaStream_t st, const RayParams &P, int renderType, const dsb200_voxel *voxels,
                  const dsb200_hash_entry *hash, const float *raycastResult, uint8_t *outRGBA, float *outDepth)
{
    dim3 grid(div_up(P.W, 16), div_up(P.H, 8));
    DSB_LAUNCH(k_shade, grid, 128, 0, st, P, renderType, voxels, hash, reinterpret_cast<const float4 *>(raycastResult),
               reinterpret_cast<uchar4 *>(outRGBA), outDepth);
}

} // namespace dsb
