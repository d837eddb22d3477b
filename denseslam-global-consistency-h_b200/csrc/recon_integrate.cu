// recon_integrate.cu -- IntegrateIntoScene / DeIntegrateIntoScene / ResetScene for sm_100a.
//
// Reference: one 512-thread CTA per visible block, each thread one strided 8-byte voxel RMW
// (CUDA/ITMSceneReconstructionEngine_CUDA.cu:1409-1442, 1483-1517); math in
// DeviceAgnostic/ITMSceneReconstructionEngine.h:17-281.
// Here: a persistent grid walks the visible list (count read on the device, so no host round
// trip between allocation and integration); one 4 KiB voxel block = 256 threads x one 16-byte
// vector load/store (two voxels per thread, fully coalesced), stores only for lanes whose voxels
// changed.  Arithmetic order is the reference's (no FMA contraction, IEEE division).
#include "kernels.cuh"

namespace dsb {

struct VoxelRegs { short sdf; unsigned char w, r, g, b, wc; };

__device__ __forceinline__ VoxelRegs unpack(unsigned lo, unsigned hi)
{
    VoxelRegs v;
    v.sdf = (short)(lo & 0xffff); v.w = (lo >> 16) & 0xff; v.r = lo >> 24;
    v.g = hi & 0xff; v.b = (hi >> 8) & 0xff; v.wc = (hi >> 16) & 0xff;
    return v;
}
__device__ __forceinline__ void pack(const VoxelRegs &v, unsigned &lo, unsigned &hi)
{
    lo = (unsigned)(unsigned short)v.sdf | ((unsigned)v.w << 16) | ((unsigned)v.r << 24);
    hi = (hi & 0xff000000u) | (unsigned)v.g | ((unsigned)v.b << 8) | ((unsigned)v.wc << 16);
}
__device__ __forceinline__ void reset(VoxelRegs &v) { v.sdf = 32767; v.w = 0; v.r = v.g = v.b = 0; v.wc = 0; }

__device__ __forceinline__ unsigned char to_uchar(float v)
{   // Vector3::toUChar: (int)ROUND(x) then CLAMP(0,255)  (ORUtils/Vector.h:245-247, MathUtils.h:24)
    int i = (int)((v < 0) ? (v - 0.5f) : (v + 0.5f));
    int lo = (255 < i) ? 255 : i;
    return (unsigned char)((0 < lo) ? lo : 0);
}

template <bool DE>
__device__ __forceinline__ void fuse_voxel(VoxelRegs &vox, float ptx, float pty, float ptz, const FuseParams &P,
                                           const float *__restrict__ depth, const uchar4 *__restrict__ rgb)
{
    if (P.stopMaxW && vox.w == P.maxW) return;
    if (P.approximate && vox.w != 0) return;
    // computeUpdatedVoxelDepthInfo / computeDeUpdatedVoxelDepthInfo
    Vec4 pt = {ptx, pty, ptz, 1.0f};
    Vec4 pc = mul(P.M_d, pt);
    if (pc.z <= 0) return;
    float u = P.proj_d[0] * pc.x / pc.z + P.proj_d[2];
    float v = P.proj_d[1] * pc.y / pc.z + P.proj_d[3];
    if ((u < 1) || (u > P.W - 2) || (v < 1) || (v > P.H - 2)) return;
    float dm = __ldg(depth + (int)(u + 0.5f) + (int)(v + 0.5f) * P.W);
    if (dm <= 0.0f) return;
    float eta = dm - pc.z;
    if (eta < -P.mu) return;
    {
        float oldF = (float)vox.sdf / 32767.0f; int oldW = vox.w;
        float q = eta / P.mu;
        float newF = (1.0f < q) ? 1.0f : q;
        int newW;
        if (P.depthWeighting) {
            newW = (int)((double)((float)P.maxDistance / dm) + 0.5);   // "+ 0.5" is a double add in the reference
            if (newW < 1) newW = 1;
            if (newW > P.maxNewW) newW = P.maxNewW;
        } else newW = 1;
        if (!DE) { newF = oldW * oldF + newW * newF; newW = oldW + newW; }
        else     { newF = oldW * oldF - newW * newF; newW = oldW - newW; }
        if (DE && newW == 0) { reset(vox); }      // defined behaviour for the reference's 0/0
        else {
            newF /= newW;
            newW = (newW < P.maxW) ? newW : P.maxW;
            vox.sdf = (short)(int)(newF * 32767.0f);
            vox.w = (unsigned char)newW;
            if (DE && vox.w == 0) reset(vox);
        }
    }
    if ((eta > P.mu) || (fabsf(eta / P.mu) > 0.25f)) return;
    // computeUpdatedVoxelColorInfo / computeDeUpdatedVoxelColorInfo
    float oldWc = (float)vox.wc;
    float oc0 = (float)vox.r / 255.0f, oc1 = (float)vox.g / 255.0f, oc2 = (float)vox.b / 255.0f;
    Vec4 pr = mul(P.M_rgb, pt);
    float ur = P.proj_rgb[0] * pr.x / pr.z + P.proj_rgb[2];
    float vr = P.proj_rgb[1] * pr.y / pr.z + P.proj_rgb[3];
    if ((ur < 1) || (ur > P.Wrgb - 2) || (vr < 1) || (vr > P.Hrgb - 2)) return;
    // interpolateBilinear (DeviceAgnostic/ITMPixelUtils.h:11-39)
    int ix = (int)floorf(ur), iy = (int)floorf(vr);
    float dx = ur - (float)ix, dy = vr - (float)iy;
    uchar4 a = __ldg(rgb + ix + iy * P.Wrgb), b = make_uchar4(0, 0, 0, 0), c = b, d = b;
    if (dx != 0) b = __ldg(rgb + (ix + 1) + iy * P.Wrgb);
    if (dy != 0) c = __ldg(rgb + ix + (iy + 1) * P.Wrgb);
    if (dx != 0 && dy != 0) d = __ldg(rgb + (ix + 1) + (iy + 1) * P.Wrgb);
    float m0 = ((float)a.x * (1.0f - dx) * (1.0f - dy) + (float)b.x * dx * (1.0f - dy) + (float)c.x * (1.0f - dx) * dy + (float)d.x * dx * dy);
    float m1 = ((float)a.y * (1.0f - dx) * (1.0f - dy) + (float)b.y * dx * (1.0f - dy) + (float)c.y * (1.0f - dx) * dy + (float)d.y * dx * dy);
    float m2 = ((float)a.z * (1.0f - dx) * (1.0f - dy) + (float)b.z * dx * (1.0f - dy) + (float)c.z * (1.0f - dx) * dy + (float)d.z * dx * dy);
    m0 = m0 / 255.0f; m1 = m1 / 255.0f; m2 = m2 / 255.0f;
    float newWc = 1;
    float n0, n1, n2;
    if (!DE) { n0 = oc0 * oldWc + m0 * newWc; n1 = oc1 * oldWc + m1 * newWc; n2 = oc2 * oldWc + m2 * newWc; newWc = oldWc + newWc; }
    else     { n0 = oc0 * oldWc - m0 * newWc; n1 = oc1 * oldWc - m1 * newWc; n2 = oc2 * oldWc - m2 * newWc; newWc = oldWc - newWc; }
    if (DE && newWc == 0.0f) { vox.r = vox.g = vox.b = 0; vox.wc = 0; return; }   // defined behaviour for x/0
    n0 /= newWc; n1 /= newWc; n2 /= newWc;
    float mw = (float)(unsigned char)P.maxW;
    newWc = (newWc < mw) ? newWc : mw;
    vox.r = to_uchar(n0 * 255.0f); vox.g = to_uchar(n1 * 255.0f); vox.b = to_uchar(n2 * 255.0f);
    vox.wc = (unsigned char)(int)newWc;
}

template <bool DE>
__global__ void __launch_bounds__(256) k_fuse(FuseParams P, dsb200_voxel *voxels,
                                              const dsb200_hash_entry *__restrict__ hash,
                                              const int *__restrict__ ids, const int *__restrict__ countPtr,
                                              const float *__restrict__ depth, const uchar4 *__restrict__ rgb)
{
    const int n = *countPtr;
    const int t = threadIdx.x;
    const int lx = (2 * t) & 7, ly = (t >> 2) & 7, lz = t >> 5;     // voxel 2t: x even, same y/z for 2t+1
    for (int i = blockIdx.x; i < n; i += gridDim.x) {
        const int slot = __ldg(ids + i);
        const HashEntry e = load_entry(hash, slot);
        if (e.ptr < 0) continue;
        uint4 *p = reinterpret_cast<uint4 *>(voxels + (size_t)e.ptr * kBlock3) + t;
        uint4 raw = *p;
        const float pty = (float)(e.py * kBlock + ly) * P.voxelSize, ptz = (float)(e.pz * kBlock + lz) * P.voxelSize;
        const int gx = e.px * kBlock + lx;
        VoxelRegs v0 = unpack(raw.x, raw.y), v1 = unpack(raw.z, raw.w);
        fuse_voxel<DE>(v0, (float)gx * P.voxelSize, pty, ptz, P, depth, rgb);
        fuse_voxel<DE>(v1, (float)(gx + 1) * P.voxelSize, pty, ptz, P, depth, rgb);
        uint4 out = raw;
        pack(v0, out.x, out.y); pack(v1, out.z, out.w);
        if (out.x != raw.x || out.y != raw.y || out.z != raw.z || out.w != raw.w) *p = out;
    }
}

void launch_fuse(cudaStream_t st, const FuseParams &P, bool deintegrate, dsb200_voxel *voxels,
                 const dsb200_hash_entry *hash, const int *ids, const int *count, const float *depth, const uint8_t *rgb)
{
    const int grid = kNumSMs * 8;
    if (deintegrate)
        DSB_LAUNCH(k_fuse<true>, grid, 256, 0, st, P, voxels, hash, ids, count, depth, reinterpret_cast<const uchar4 *>(rgb));
    else
        DSB_LAUNCH(k_fuse<false>, grid, 256, 0, st, P, voxels, hash, ids, count, depth, reinterpret_cast<const uchar4 *>(rgb));
}

// ResetScene (CPU .cpp:25-45; CUDA .cu:111-148): pure store bandwidth, 16-byte stores.
__global__ void __launch_bounds__(256) k_reset_scene(uint4 *voxels2, size_t numVoxelPairs, int *hashWords, int total,
                                                     int *allocList, int numBlocks, int *excessList, int excessSize,
                                                     uint8_t *visType, unsigned *reqKey, Counters *counters)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    const uint4 blank = make_uint4(0x00007fffu, 0u, 0x00007fffu, 0u);   // {sdf 32767, w 0, clr 0, wc 0} x2
    for (size_t i = tid; i < numVoxelPairs; i += stride) voxels2[i] = blank;
    for (size_t i = tid; i < (size_t)total; i += stride) {
        int *p = hashWords + i * 5;
        p[0] = 0; p[1] = 0; p[2] = 0; p[3] = -2; p[4] = 0;
        if (visType) visType[i] = 0;
        reqKey[i] = 0;
    }
    for (size_t i = tid; i < (size_t)numBlocks; i += stride) allocList[i] = (int)i;
    for (size_t i = tid; i < (size_t)excessSize; i += stride) excessList[i] = (int)i;
    if (tid == 0) {
        counters->lastFreeBlock = numBlocks - 1; counters->lastFreeExcess = excessSize - 1;
        counters->lastFreeBlockNew = numBlocks - 1; counters->lastFreeExcessNew = excessSize - 1;
        counters->noVisible = 0; counters->noVisiblePrev = 0; counters->renderBlocks = 0; counters->freedBlocks = 0;
    }
}

void launch_reset_scene(cudaStream_t st, dsb200_voxel *voxels, size_t numVoxels, dsb200_hash_entry *hash, int total,
                        int *allocList, int numBlocks, int *excessList, int excessSize, uint8_t *visTypeOrNull,
                        unsigned *reqKey, Counters *counters)
{
    DSB_LAUNCH(k_reset_scene, kNumSMs * 16, 256, 0, st, reinterpret_cast<uint4 *>(voxels), numVoxels / 2,
               reinterpret_cast<int *>(hash), total, allocList, numBlocks, excessList, excessSize, visTypeOrNull, reqKey,
               counters);
}

} // namespace dsb
