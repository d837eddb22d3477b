// mesh.cu -- ITMMeshingEngine<ITMVoxel,ITMVoxelBlockHash>::MeshScene for sm_100a.
//
// Reference: CPU/ITMMeshingEngine_CPU.cpp:20-83 (serial scan over the hash table, one marching-cubes cell per voxel, triangles
// appended in scan order) and CUDA/ITMMeshingEngine_CUDA.cu:60-195 (one 512-thread CTA per VBA block, triangles appended with
// atomicAdd: order = schedule).  Math: DeviceAgnostic/ITMMeshingEngine.h:40-160 (findPointNeighbors, buildVertList, sdfInterp)
// and DeviceAgnostic/ITMRepresentationAccess.h:699-736 (vertex colours).
// Here the output is the CPU engine's, triangle for triangle, bit for bit:
//   k_mesh_list    ordered compaction of the allocated hash slots (ascending slot, scan.cuh)
//   k_mesh_count   one CTA per allocated block: marching-cubes case of its 512 cells -> triangles of the block
//   k_mesh_offsets ordered exclusive scan of the per-block counts (+ grand total)
//   k_mesh_emit    one CTA per block again: in-block scan in cell order (z, y, x), vertices + colours written straight to their
//                  final position; the CPU engine's cap rule (count stops at max - 1, the last slot keeps being overwritten)
//                  is reproduced by letting only the very last triangle write the overflow slot.
#include <cstring>
#include "kernels.cuh"
#include "scan.cuh"
#include "mc_tables.h"

namespace dsb {

int fail_with(int code, const char *msg);

__constant__ unsigned long long cMcTriangles[256];

struct VoxelCache { int bx, by, bz; const uint2 *blk; };

// readVoxel (DeviceAgnostic/ITMRepresentationAccess.h:176-210): the 8-byte voxel record, or the default voxel when the block is missing
__device__ __forceinline__ uint2 read_voxel(const dsb200_voxel *__restrict__ voxels, const dsb200_hash_entry *__restrict__ hash,
                                            int bucketNum, int bucketMask, int px, int py, int pz, bool &found, VoxelCache &c)
{
    const int bx = px >> 3, by = py >> 3, bz = pz >> 3;
    const int lin = (px & 7) + ((py & 7) << 3) + ((pz & 7) << 6);
    if (bx == c.bx && by == c.by && bz == c.bz) { found = true; return __ldg(c.blk + lin); }
    int h = hash_index(bx, by, bz, bucketMask);
    while (true) {
        const HashEntry e = load_entry(hash, h);
        if (e.px == bx && e.py == by && e.pz == bz && e.ptr >= 0) {
            found = true; c.bx = bx; c.by = by; c.bz = bz;
            c.blk = reinterpret_cast<const uint2 *>(voxels + (size_t)e.ptr * kBlock3);
            return __ldg(c.blk + lin);
        }
        if (e.offset < 1) break;
        h = bucketNum + e.offset - 1;
    }
    found = false;
    return make_uint2(0x00007fffu, 0u);      // sdf 32767, everything else 0
}

// corners in the order of findPointNeighbors, edges in the order of buildVertList
__device__ __constant__ int cCorner[8][3] = {{0, 0, 0}, {1, 0, 0}, {1, 1, 0}, {0, 1, 0}, {0, 0, 1}, {1, 0, 1}, {1, 1, 1}, {0, 1, 1}};
__device__ __constant__ int cEdge[12][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}, {4, 5}, {5, 6}, {6, 7}, {7, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};

// marching-cubes case of the cell whose low corner is voxel (gx, gy, gz): -1 = no triangles; sdf[] = the 8 corner values
__device__ __forceinline__ int cell_case(const dsb200_voxel *__restrict__ voxels, const dsb200_hash_entry *__restrict__ hash,
                                         int bucketNum, int bucketMask, int gx, int gy, int gz, float sdf[8], VoxelCache &c)
{
    int cube = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        bool found;
        const uint2 v = read_voxel(voxels, hash, bucketNum, bucketMask, gx + cCorner[i][0], gy + cCorner[i][1], gz + cCorner[i][2], found, c);
        const short s = (short)(v.x & 0xffffu);
        sdf[i] = (float)s / 32767.0f;
        if (!found || s == 32767) return -1;                  // sdf == 1.0f  <=>  the short is 32767
        if (sdf[i] < 0) cube |= 1 << i;
    }
    return (cMcTriangles[cube] & 0xFull) == 0xFull ? -1 : cube;   // edgeTable[cube] == 0  <=>  the case has no triangle
}

__device__ __forceinline__ int case_triangles(int cube)
{
    if (cube < 0) return 0;
    const unsigned long long w = cMcTriangles[cube];
    int n = 0;
#pragma unroll
    for (int k = 0; k < 5; k++) n += (((w >> (12 * k)) & 0xFull) != 0xFull) ? 1 : 0;
    return n;
}

// ---- k_mesh_list: allocated slots, ascending
__global__ void __launch_bounds__(256) k_mesh_list(const dsb200_hash_entry *__restrict__ hash, int total, unsigned long long *scanState,
                                                   int *ticket, int *list, int *count)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx;
    if (threadIdx.x == 0) sTile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = sTile, slot0 = tile * 1024 + threadIdx.x * 4;
    bool live[4]; unsigned na = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        live[j] = slot0 + j < total && load_entry(hash, slot0 + j).ptr >= 0;
        na += live[j] ? 1 : 0;
    }
    unsigned exA, exB, totA, totB;
    block_scan2(na, 0, sScan, exA, exB, totA, totB);
    if (threadIdx.x < 32) {
        unsigned ta, tb;
        scan_lookback(scanState, tile, totA, 0, ta, tb);
        if (threadIdx.x == 0) sEx = ta;
    }
    __syncthreads();
    unsigned r = sEx + exA;
#pragma unroll
    for (int j = 0; j < 4; j++) if (live[j]) list[r++] = slot0 + j;
    if (threadIdx.x == 0 && tile == gridDim.x - 1) *count = (int)(sEx + totA);
}

// ---- k_mesh_count / k_mesh_emit share the per-block classification: thread t owns cells t and t + 256 (cell = x + 8 y + 64 z)
template <bool EMIT>
__global__ void __launch_bounds__(256) k_mesh_blocks(const dsb200_voxel *__restrict__ voxels, const dsb200_hash_entry *__restrict__ hash,
                                                     int bucketNum, const int *__restrict__ list, const int *__restrict__ countPtr,
                                                     unsigned *blockTri /* counts (EMIT = false) / exclusive offsets (EMIT = true) */,
                                                     const unsigned *__restrict__ totalPtr, float factor, dsb200_triangle *tris,
                                                     unsigned maxTriangles)
{
    __shared__ unsigned sScan[16];
    const int n = *countPtr, bucketMask = bucketNum - 1;
    for (int b = blockIdx.x; b < n; b += gridDim.x) {
        const HashEntry e = load_entry(hash, list[b]);
        VoxelCache c = {0x7fffffff, 0x7fffffff, 0x7fffffff, nullptr};
        int cube[2]; float sdf[2][8]; unsigned cnt[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int cell = threadIdx.x + h * 256;
            cube[h] = cell_case(voxels, hash, bucketNum, bucketMask, e.px * kBlock + (cell & 7), e.py * kBlock + ((cell >> 3) & 7),
                                e.pz * kBlock + (cell >> 6), sdf[h], c);
            cnt[h] = (unsigned)case_triangles(cube[h]);
        }
        // cells 0..255 (first halves) precede cells 256..511 (second halves): two scans in one pass
        unsigned ex0, ex1, tot0, tot1;
        block_scan2(cnt[0], cnt[1], sScan, ex0, ex1, tot0, tot1);
        if (!EMIT) {
            if (threadIdx.x == 0) blockTri[b] = tot0 + tot1;
            continue;
        }
        const unsigned total = *totalPtr;
        const unsigned base = blockTri[b];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (cube[h] < 0) continue;
            const int cell = threadIdx.x + h * 256;
            const int gx = e.px * kBlock + (cell & 7), gy = e.py * kBlock + ((cell >> 3) & 7), gz = e.pz * kBlock + (cell >> 6);
            unsigned g = base + (h ? tot0 + ex1 : ex0);
            const unsigned long long w = cMcTriangles[cube[h]];
            for (int i = 0; i < 15 && ((w >> (4 * i)) & 0xFull) != 0xFull; i += 3, g++) {
                unsigned slot = g;
                if (g >= maxTriangles - 1u) {               // capped: only the last triangle of all survives, in the last slot
                    if (g != total - 1u) continue;
                    slot = maxTriangles - 1u;
                }
                float out[18];
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const int ed = (int)((w >> (4 * (i + k))) & 0xFull);
                    const int ca = cEdge[ed][0], cb = cEdge[ed][1];
                    const float ax = (float)(gx + cCorner[ca][0]), ay = (float)(gy + cCorner[ca][1]), az = (float)(gz + cCorner[ca][2]);
                    const float bx = (float)(gx + cCorner[cb][0]), by = (float)(gy + cCorner[cb][1]), bz = (float)(gz + cCorner[cb][2]);
                    const float va = sdf[h][ca], vb = sdf[h][cb];
                    float px, py, pz;                        // sdfInterp
                    if (fabsf(0.0f - va) < 0.00001f) { px = ax; py = ay; pz = az; }
                    else if (fabsf(0.0f - vb) < 0.00001f) { px = bx; py = by; pz = bz; }
                    else if (fabsf(va - vb) < 0.00001f) { px = ax; py = ay; pz = az; }
                    else {
                        const float t = (0.0f - va) / (vb - va);
                        px = ax + t * (bx - ax); py = ay + t * (by - ay); pz = az + t * (bz - az);
                    }
                    out[3 * k] = px * factor; out[3 * k + 1] = py * factor; out[3 * k + 2] = pz * factor;
                    // readFromSDF_color4u_interpolated_noalpha at the vertex (voxel units)
                    const float flx = floorf(px), fly = floorf(py), flz = floorf(pz);
                    const float cx = px - flx, cy = py - fly, cz = pz - flz;
                    const int X = (int)flx, Y = (int)fly, Z = (int)flz;
                    float r0 = 0.0f, r1 = 0.0f, r2 = 0.0f;
#pragma unroll
                    for (int q = 0; q < 8; q++) {
                        const int ox = q & 1, oy = (q >> 1) & 1, oz = q >> 2;
                        bool f;
                        const uint2 v = read_voxel(voxels, hash, bucketNum, bucketMask, X + ox, Y + oy, Z + oz, f, c);
                        const float wgt = (ox ? cx : (1.0f - cx)) * (oy ? cy : (1.0f - cy)) * (oz ? cz : (1.0f - cz));
                        r0 += wgt * (float)(v.x >> 24); r1 += wgt * (float)(v.y & 0xffu); r2 += wgt * (float)((v.y >> 8) & 0xffu);
                    }
                    out[9 + 3 * k] = r0 / 255.0f; out[9 + 3 * k + 1] = r1 / 255.0f; out[9 + 3 * k + 2] = r2 / 255.0f;
                }
                float *dst = reinterpret_cast<float *>(tris + slot);
#pragma unroll
                for (int q = 0; q < 18; q++) dst[q] = out[q];
            }
        }
    }
}

// ---- k_mesh_offsets: ordered exclusive scan of the per-block counts, in place; *total = sum
__global__ void __launch_bounds__(256) k_mesh_offsets(unsigned *blockTri, const int *__restrict__ countPtr, unsigned long long *scanState,
                                                      int *ticket, unsigned *total)
{
    __shared__ unsigned sScan[16];
    __shared__ int sTile;
    __shared__ unsigned sEx;
    const int n = *countPtr, tiles = (n + 1023) / 1024;
    if (threadIdx.x == 0) sTile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = sTile;
    if (tile >= tiles) { if (tile == 0 && threadIdx.x == 0) *total = 0; return; }
    const int i0 = tile * 1024 + threadIdx.x * 4;
    unsigned v[4], s = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) { v[j] = (i0 + j < n) ? blockTri[i0 + j] : 0u; s += v[j]; }
    unsigned exA, exB, totA, totB;
    block_scan2(s, 0, sScan, exA, exB, totA, totB);
    if (threadIdx.x < 32) {
        unsigned ta, tb;
        scan_lookback(scanState, tile, totA, 0, ta, tb);
        if (threadIdx.x == 0) sEx = ta;
    }
    __syncthreads();
    unsigned r = sEx + exA;
#pragma unroll
    for (int j = 0; j < 4; j++) { if (i0 + j < n) blockTri[i0 + j] = r; r += v[j]; }
    if (threadIdx.x == 0 && tile == tiles - 1) *total = sEx + totA;
}

} // namespace dsb

using namespace dsb;

struct dsb200_mesher {
    cudaStream_t stream = nullptr;
    int numBlocks = 0;
    int *list = nullptr, *ctl = nullptr;        // ctl: [0] list count, [1] ticket of k_mesh_list, [2] ticket of k_mesh_offsets, [3] total triangles
    unsigned *blockTri = nullptr;
    unsigned long long *scanState = nullptr; int scanTiles = 0;
    int *hostCtl = nullptr;                     // pinned copy of ctl
};

#define MCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_with(DSB200_ERR_CUDA, cudaGetErrorString(e_)); } while (0)

extern "C" {

int dsb200_mesher_create(dsb200_mesher **out, int32_t num_blocks, void *stream)
{
    if (!out || num_blocks <= 0) return fail_with(DSB200_ERR_INVALID, "mesher_create: bad arguments");
    if (dsb200_device_count() == 0) return fail_with(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    dsb200_mesher *m = new dsb200_mesher();
    m->stream = (cudaStream_t)stream; m->numBlocks = num_blocks;
    MCU(cudaMalloc(&m->list, (size_t)num_blocks * sizeof(int)));
    MCU(cudaMalloc(&m->blockTri, (size_t)num_blocks * sizeof(unsigned)));
    MCU(cudaMalloc(&m->ctl, 4 * sizeof(int)));
    MCU(cudaMallocHost(&m->hostCtl, 4 * sizeof(int)));
    MCU(cudaMemcpyToSymbol(cMcTriangles, kMcTrianglesHost, sizeof(kMcTrianglesHost)));
    *out = m;
    return DSB200_OK;
}

int dsb200_mesher_destroy(dsb200_mesher *m)
{
    if (!m) return DSB200_OK;
    cudaStreamSynchronize(m->stream);
    cudaFree(m->list); cudaFree(m->blockTri); cudaFree(m->ctl); cudaFree(m->scanState); cudaFreeHost(m->hostCtl);
    delete m;
    return DSB200_OK;
}

int dsb200_mesh_scene(dsb200_mesher *m, const dsb200_scene *scene, dsb200_triangle *triangles, uint32_t max_triangles,
                      uint32_t *no_triangles)
{
    if (!m || !scene || !triangles || !no_triangles || max_triangles < 2) return fail_with(DSB200_ERR_INVALID, "mesh_scene: bad arguments");
    if (scene->num_blocks > m->numBlocks) return fail_with(DSB200_ERR_INVALID, "mesh_scene: the scene holds more blocks than the mesher was created for");
    const int total = scene->bucket_num + scene->excess_size;
    const int tilesA = div_up(total, 1024), tilesB = div_up(m->numBlocks, 1024);
    const int tiles = tilesA > tilesB ? tilesA : tilesB;
    if (tiles > m->scanTiles) {
        cudaFree(m->scanState); m->scanState = nullptr; m->scanTiles = 0;
        MCU(cudaMalloc(&m->scanState, (size_t)tiles * sizeof(unsigned long long)));
        m->scanTiles = tiles;
    }
    cudaStream_t st = m->stream;
    MCU(cudaMemsetAsync(m->ctl, 0, 4 * sizeof(int), st));
    MCU(cudaMemsetAsync(m->scanState, 0, (size_t)tiles * sizeof(unsigned long long), st));
    MCU(cudaMemsetAsync(triangles, 0, (size_t)max_triangles * sizeof(dsb200_triangle), st));      // mesh->triangles->Clear()
    DSB_LAUNCH(k_mesh_list, tilesA, 256, 0, st, scene->hash, total, m->scanState, m->ctl + 1, m->list, m->ctl);
    DSB_LAUNCH((k_mesh_blocks<false>), kNumSMs * 8, 256, 0, st, scene->voxels, scene->hash, scene->bucket_num, m->list, m->ctl, m->blockTri,
               nullptr, 0.0f, nullptr, 0u);
    MCU(cudaMemsetAsync(m->scanState, 0, (size_t)tiles * sizeof(unsigned long long), st));
    DSB_LAUNCH(k_mesh_offsets, tilesB, 256, 0, st, m->blockTri, m->ctl, m->scanState, m->ctl + 2, reinterpret_cast<unsigned *>(m->ctl + 3));
    DSB_LAUNCH((k_mesh_blocks<true>), kNumSMs * 8, 256, 0, st, scene->voxels, scene->hash, scene->bucket_num, m->list, m->ctl, m->blockTri,
               reinterpret_cast<const unsigned *>(m->ctl + 3), scene->params.voxel_size, triangles, max_triangles);
    MCU(cudaGetLastError());
    MCU(cudaMemcpyAsync(m->hostCtl, m->ctl, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    MCU(cudaStreamSynchronize(st));
    const uint32_t total_tris = (uint32_t)m->hostCtl[3];
    *no_triangles = total_tris < max_triangles - 1u ? total_tris : max_triangles - 1u;
    return DSB200_OK;
}

} // extern "C"
