// recon_integrate.cu -- IntegrateIntoScene / DeIntegrateIntoScene / ResetScene for sm_100a.
//
// Reference: one 512-thread CTA per visible block, each thread one strided 8-byte voxel RMW
// (CUDA/ITMSceneReconstructionEngine_CUDA.cu:1409-1442, 1483-1517); math in
// DeviceAgnostic/ITMSceneReconstructionEngine.h:17-281.
// Here: a persistent grid walks the visible list (count read on the device, so no host round
// trip between allocation and integration); one 4 KiB voxel block = 256 threads x one 16-byte
// vector load/store (two voxels per thread, fully coalesced), stores only for lanes whose voxels
// changed.  Arithmetic order is the reference's (no FMA contraction, IEEE division).
#include <cstdlib>
#include <type_traits>
#include "kernels.cuh"
#include "exactdiv.cuh"

namespace dsb {

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

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

// Shared look-up tables: reciprocals of the integer voxel weights and c/255 for the colour bytes,
// both produced with IEEE division at kernel start (3 divisions per thread).
struct FuseTables { float rcp[kRcpTable]; float byte255[256]; };

__device__ __forceinline__ void fill_tables(FuseTables &T)
{
    for (int i = threadIdx.x; i < kRcpTable; i += blockDim.x) T.rcp[i] = 1.0f / (float)(i > 0 ? i : 1);
    for (int i = threadIdx.x; i < 256; i += blockDim.x) T.byte255[i] = (float)i / 255.0f;
}

// Depth part of ComputeUpdatedVoxelInfo (computeUpdatedVoxelDepthInfo / computeDeUpdatedVoxelDepthInfo,
// DeviceAgnostic/ITMSceneReconstructionEngine.h:17-77, 117-181).  Returns true when the colour part
// (:80-113, 184-217) must run for this voxel; that part is executed later by k_fuse on a CTA-wide
// compacted list, so that a warp does not pay for it when only a few of its voxels are near the surface.
template <bool DE, bool FAST>
__device__ __forceinline__ bool fuse_depth_pc(VoxelRegs &vox, Vec4 pc, const FuseParams &P, const float *__restrict__ depth,
                                              const FuseTables &T);

template <bool DE, bool FAST>
__device__ __forceinline__ bool fuse_depth(VoxelRegs &vox, float ptx, float pty, float ptz, const FuseParams &P,
                                           const float *__restrict__ depth, const FuseTables &T)
{
    if (P.stopMaxW && vox.w == P.maxW) return false;
    if (P.approximate && vox.w != 0) return false;
    Vec4 pt = {ptx, pty, ptz, 1.0f};
    return fuse_depth_pc<DE, FAST>(vox, mul(P.M_d, pt), P, depth, T);
}

// the same with the camera-space point supplied by the caller (k_fuse_warp shares the partial sums of M_d * pt between
// the voxels of a column; the summation order is the reference's, so pc is bit-identical)
template <bool DE, bool FAST>
__device__ __forceinline__ bool fuse_depth_pc(VoxelRegs &vox, Vec4 pc, const FuseParams &P, const float *__restrict__ depth,
                                              const FuseTables &T)
{
    typedef Exact<FAST> X;
    if (pc.z <= 0) return false;
    float qu, qv;
    X::pair(P.proj_d[0] * pc.x, P.proj_d[1] * pc.y, pc.z, qu, qv);
    float u = qu + P.proj_d[2];
    float v = qv + P.proj_d[3];
    if ((u < 1) || (u > P.W - 2) || (v < 1) || (v > P.H - 2)) return false;
    float dm = __ldg(depth + (int)(u + 0.5f) + (int)(v + 0.5f) * P.W);
    if (dm <= 0.0f) return false;
    float eta = dm - pc.z;
    if (eta < -P.mu) return false;
    const float q = X::by_host_rcp(eta, P.mu, P.rcpMu);
    {
        float oldF = X::by_int((float)vox.sdf, 32767.0f, P.rcp32767); int oldW = vox.w;
        float newF = (1.0f < q) ? 1.0f : q;
        int newW;
        if (P.depthWeighting) {
            // (int)((float)maxDistance / dm + 0.5) with a DOUBLE add in the reference: the sum is exact in double, so
            // the result is trunc(x) + (frac(x) >= 0.5), computed exactly in float (0 < x < 2^23 when it matters)
            float x = X::one((float)P.maxDistance, dm);
            if (x < 8388608.0f) { float tr = truncf(x); newW = (int)tr + ((x - tr >= 0.5f) ? 1 : 0); }
            else newW = 0x7fffffff;
            if (newW < 1) newW = 1;
            if (newW > P.maxNewW) newW = P.maxNewW;
        } else newW = 1;
        if (!DE) { newF = oldW * oldF + newW * newF; newW = oldW + newW; }
        else     { newF = oldW * oldF - newW * newF; newW = oldW - newW; }
        if (DE && newW == 0) { reset(vox); }      // defined behaviour for the reference's 0/0
        else {
            if ((unsigned)newW - 1u < (unsigned)(kRcpTable - 1)) newF = X::by_int(newF, (float)newW, T.rcp[newW]);
            else newF /= newW;
            newW = (newW < P.maxW) ? newW : P.maxW;
            vox.sdf = (short)(int)(newF * 32767.0f);
            vox.w = (unsigned char)newW;
            if (DE && vox.w == 0) reset(vox);
        }
    }
    return !((eta > P.mu) || (fabsf(q) > 0.25f));
}

template <bool DE, bool FAST>
__device__ __forceinline__ void fuse_colour(VoxelRegs &vox, float ptx, float pty, float ptz, const FuseParams &P,
                                            const uchar4 *__restrict__ rgb, const FuseTables &T)
{
    typedef Exact<FAST> X;
    float oldWc = (float)vox.wc;
    float oc0 = T.byte255[vox.r], oc1 = T.byte255[vox.g], oc2 = T.byte255[vox.b];
    Vec4 pt = {ptx, pty, ptz, 1.0f};
    Vec4 pr = mul(P.M_rgb, pt);
    float ur, vr;   // no z > 0 gate here in the reference; pair() falls back to plain division outside its window
    X::pair(P.proj_rgb[0] * pr.x, P.proj_rgb[1] * pr.y, pr.z, ur, vr);
    ur = ur + P.proj_rgb[2];
    vr = vr + P.proj_rgb[3];
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
    m0 = X::by_int(m0, 255.0f, P.rcp255); m1 = X::by_int(m1, 255.0f, P.rcp255); m2 = X::by_int(m2, 255.0f, P.rcp255);
    float newWc = 1;
    float n0, n1, n2;
    if (!DE) { n0 = oc0 * oldWc + m0 * newWc; n1 = oc1 * oldWc + m1 * newWc; n2 = oc2 * oldWc + m2 * newWc; newWc = oldWc + newWc; }
    else     { n0 = oc0 * oldWc - m0 * newWc; n1 = oc1 * oldWc - m1 * newWc; n2 = oc2 * oldWc - m2 * newWc; newWc = oldWc - newWc; }
    // De-fusing the last colour observation divides by zero and de-fusing from weight 0 leaves a weight of -1; the reference's
    // only implementation is CUDA, so the device's own conversions of inf / NaN / negative values ARE the reference behaviour
    // (cvt.rzi.s32.f32 saturates, NaN -> 0; `(uchar)float` is cvt.rzi.u32.f32: negative -> 0): nothing is special-cased here.
    if (newWc >= 1.0f) {      // integer-valued, <= 256
        const float y = T.rcp[(int)newWc];
        n0 = X::by_int(n0, newWc, y); n1 = X::by_int(n1, newWc, y); n2 = X::by_int(n2, newWc, y);
    } else { n0 /= newWc; n1 /= newWc; n2 /= newWc; }
    float mw = (float)(unsigned char)P.maxW;
    newWc = (newWc < mw) ? newWc : mw;
    vox.r = to_uchar(n0 * 255.0f); vox.g = to_uchar(n1 * 255.0f); vox.b = to_uchar(n2 * 255.0f);
    vox.wc = (unsigned char)newWc;        // the reference's `(uchar)newW` on a float (DeviceAgnostic/...:112,214)
}

// ---------------------------------------------------------------- warp-per-block kernel (round 1; de-integration and fall-back)
// One WARP fuses one 4 KiB voxel block: 8 z-slices of 64 voxels, lane l owns the voxel pair (x = 2(l&3), 2(l&3)+1; y = l>>2)
// of every slice, i.e. 16 voxels per lane per block.  Against the CTA-per-block kernel above this
//  * removes both CTA barriers per block (20 % of the stall samples in profiles/r01b) -- a warp never waits for another;
//  * amortises the list -> hash entry -> block prefetch pipeline and the queue bookkeeping over 16 voxels per lane, not 2;
//  * shares the x/y part of M_d * pt along a voxel column: pc = ((m0 x + m4 y) + m8 z) + m12 keeps the reference's
//    summation order, so only `+ m8 z` and `+ m12` remain per voxel (bit-identical).
// The colour part still runs on a compacted queue (per warp, drained densely after the 8 slices) on the warp's
// shared-memory copy of the block; 16-byte vectors are written back only where something changed.
constexpr int kFuseWarps = 4;

struct FuseWarpSmem {
    FuseTables T;
    uint4 vox[kFuseWarps][256];
    unsigned short queue[kFuseWarps][kBlock3];
    unsigned dirty[kFuseWarps][8];
};

template <bool DE, bool FAST, int MINB>
__global__ void __launch_bounds__(kFuseWarps * 32, MINB) k_fuse_warp(FuseParams P, dsb200_voxel *voxels,
                                                                   const dsb200_hash_entry *__restrict__ hash,
                                                                   const int *__restrict__ ids, const int *__restrict__ countPtr,
                                                                   const float *__restrict__ depth, const uchar4 *__restrict__ rgb,
                                                                   unsigned *shadow2)
{
    __shared__ FuseWarpSmem S;
    const int n = *countPtr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    fill_tables(S.T);
    if (lane < 8) S.dirty[warp][lane] = 0;
    __syncthreads();
    const FuseTables &T = S.T;
    uint4 *sVox = S.vox[warp];
    unsigned short *sQueue = S.queue[warp];
    unsigned *sDirty = S.dirty[warp];
    const int lx = (lane & 3) * 2, ly = lane >> 2;
    const int stride = gridDim.x * kFuseWarps;
    const HashEntry none = {0, 0, 0, 0, 0, -2, 0};
    int i = blockIdx.x * kFuseWarps + warp;
    // software pipeline over the warp's blocks: list entry 2 ahead, hash entry 1 ahead (+ L2 prefetch of its block)
    int slot2 = (i + 2 * stride < n) ? __ldg(ids + i + 2 * stride) : -1;
    HashEntry e1 = (i + stride < n) ? load_entry(hash, __ldg(ids + i + stride)) : none;
    HashEntry e0 = (i < n) ? load_entry(hash, __ldg(ids + i)) : none;
    for (; i < n; i += stride) {
        const int slot3 = (i + 3 * stride < n) ? __ldg(ids + i + 3 * stride) : -1;
        const HashEntry e2 = (slot2 >= 0) ? load_entry(hash, slot2) : none;
        if (e1.ptr >= 0) prefetch_l2(reinterpret_cast<const char *>(voxels + (size_t)e1.ptr * kBlock3) + lane * 128);
        if (e0.ptr >= 0) {
            const HashEntry e = e0;
            uint4 *base = reinterpret_cast<uint4 *>(voxels + (size_t)e.ptr * kBlock3);
            uint4 raw = base[lane];
            const int gx = e.px * kBlock + lx;
            const float ptx0 = (float)gx * P.voxelSize, ptx1 = (float)(gx + 1) * P.voxelSize;
            const float pty = (float)(e.py * kBlock + ly) * P.voxelSize;
            // (m0 x + m4 y) per row, for the two columns of this lane
            const float ax0 = P.M_d.m[0] * ptx0 + P.M_d.m[4] * pty, ay0 = P.M_d.m[1] * ptx0 + P.M_d.m[5] * pty, az0 = P.M_d.m[2] * ptx0 + P.M_d.m[6] * pty;
            const float ax1 = P.M_d.m[0] * ptx1 + P.M_d.m[4] * pty, ay1 = P.M_d.m[1] * ptx1 + P.M_d.m[5] * pty, az1 = P.M_d.m[2] * ptx1 + P.M_d.m[6] * pty;
            unsigned dirty = 0;
            int qn = 0;
#pragma unroll 1
            for (int z = 0; z < kBlock; z++) {
                uint4 nxt = raw;
                if (z + 1 < kBlock) nxt = base[(z + 1) * 32 + lane];
                const float ptz = (float)(e.pz * kBlock + z) * P.voxelSize;
                const float bx = P.M_d.m[8] * ptz, by = P.M_d.m[9] * ptz, bz = P.M_d.m[10] * ptz;
                VoxelRegs v0 = unpack(raw.x, raw.y), v1 = unpack(raw.z, raw.w);
                bool c0 = false, c1 = false;
                if (!((P.stopMaxW && v0.w == P.maxW) || (P.approximate && v0.w != 0))) {
                    Vec4 pc = {(ax0 + bx) + P.M_d.m[12], (ay0 + by) + P.M_d.m[13], (az0 + bz) + P.M_d.m[14], 1.0f};
                    c0 = fuse_depth_pc<DE, FAST>(v0, pc, P, depth, T);
                }
                if (!((P.stopMaxW && v1.w == P.maxW) || (P.approximate && v1.w != 0))) {
                    Vec4 pc = {(ax1 + bx) + P.M_d.m[12], (ay1 + by) + P.M_d.m[13], (az1 + bz) + P.M_d.m[14], 1.0f};
                    c1 = fuse_depth_pc<DE, FAST>(v1, pc, P, depth, T);
                }
                uint4 out = raw;
                pack(v0, out.x, out.y); pack(v1, out.z, out.w);
                if (out.x != raw.x || out.y != raw.y || out.z != raw.z || out.w != raw.w) {
                    dirty |= 1u << z;
                    // sdf shadow (read by the raycast): the colour part never touches the sdf, so it is final here
                    if (shadow2) shadow2[(size_t)e.ptr * 256 + z * 32 + lane] = (out.x & 0xffffu) | (out.z << 16);
                }
                sVox[z * 32 + lane] = out;
                const unsigned m0 = __ballot_sync(0xffffffffu, c0), m1 = __ballot_sync(0xffffffffu, c1);
                if (c0) sQueue[qn + __popc(m0 & lt)] = (unsigned short)(z * 64 + 2 * lane);
                if (c1) sQueue[qn + __popc(m0) + __popc(m1 & lt)] = (unsigned short)(z * 64 + 2 * lane + 1);
                qn += __popc(m0) + __popc(m1);
                raw = nxt;
            }
            __syncwarp();
            if (qn) {
                uint2 *sv = reinterpret_cast<uint2 *>(sVox);
                for (int k = lane; k < qn; k += 32) {
                    const int vi = sQueue[k];
                    uint2 w = sv[vi];
                    const uint2 before = w;
                    VoxelRegs vx = unpack(w.x, w.y);
                    fuse_colour<DE, FAST>(vx, (float)(e.px * kBlock + (vi & 7)) * P.voxelSize, (float)(e.py * kBlock + ((vi >> 3) & 7)) * P.voxelSize,
                                          (float)(e.pz * kBlock + (vi >> 6)) * P.voxelSize, P, rgb, T);
                    pack(vx, w.x, w.y);
                    if (w.x != before.x || w.y != before.y) {
                        sv[vi] = w;
                        atomicOr(&sDirty[vi >> 6], 1u << ((vi >> 1) & 31));
                    }
                }
                __syncwarp();
            }
#pragma unroll
            for (int z = 0; z < kBlock; z++) {
                const unsigned d = qn ? sDirty[z] : 0u;
                if (((dirty >> z) & 1u) || ((d >> lane) & 1u)) base[z * 32 + lane] = sVox[z * 32 + lane];
            }
            if (qn) {
                __syncwarp();
                if (lane < 8) sDirty[lane] = 0;
            }
            __syncwarp();
        }
        e0 = e1; e1 = e2; slot2 = slot3;
    }
}

// ---------------------------------------------------------------- flat variant (fuse_kernel = 2)
// Same decomposition as k_fuse_warp (one warp per block, 8 slices, a voxel pair per lane), but the depth part of both voxels is
// STRAIGHT-LINE code: every early return of computeUpdatedVoxelDepthInfo (:17-77) becomes a predicate, the depth pixel is always
// loaded (index 0 when the projection is rejected), and the result is selected at the end.  A warp executes every path of the
// branchy version anyway as soon as one lane takes it; without the 5-deep nest of divergent branches per voxel the two voxels'
// dependency chains (13 dependent FMAs for the projection alone) interleave, and the reconvergence bookkeeping disappears.
// The per-pixel part of the depth weighting, (int)(maxDistance / d + 0.5) clamped to [1, maxNewW] (:52-62), is taken from a byte
// image computed once per frame (k_depth_weights) instead of once per voxel.  Every operation on data is the one k_fuse_warp
// performs, in the same order: results are bit-identical (tests: DSB200_FUSE_KERNEL=2 runs the whole parity suite).
// Fusion (not de-fusion) with the exact-division fast paths only; launch_fuse falls back to k_fuse_warp otherwise.
__global__ void __launch_bounds__(256) k_depth_weights(const float *__restrict__ depth, unsigned char *wimg, int n, int maxDistance, int maxNewW)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float dm = depth[i];
        int newW = 1;
        if (dm > 0.0f) {
            const float x = (float)maxDistance / dm;
            if (x < 8388608.0f) { const float tr = truncf(x); newW = (int)tr + ((x - tr >= 0.5f) ? 1 : 0); }
            else newW = 0x7fffffff;
            if (newW < 1) newW = 1;
            if (newW > maxNewW) newW = maxNewW;
        }
        wimg[i] = (unsigned char)newW;
    }
}

// depth part on the packed low word of a voxel (sdf | w_depth << 16 | clr.r << 24); returns "the colour part must run"
// `outside` is raised when the voxel's camera-space z lies outside the operand window of the exact-division fast path
// (|z| < 2^-20 or > 2^20 metres: a voxel within a micrometre of the camera plane); the caller then redoes the slice with the
// general code.  Keeping that test out of this function keeps it free of branches.
template <bool WIMG>
__device__ __forceinline__ bool fuse_depth_flat(unsigned &lo, float pcx, float pcy, float pcz, const FuseParams &P,
                                                const float *__restrict__ depth, const unsigned char *__restrict__ wimg,
                                                const FuseTables &T, bool &outside)
{
    const int w = (int)((lo >> 16) & 0xffu);
    bool ok = !(pcz <= 0.0f);
    if (P.stopMaxW) ok = ok && (w != P.maxW);
    if (P.approximate) ok = ok && (w == 0);
    outside = outside || (ok && !div_range_ok(pcz));
    const float a0 = P.proj_d[0] * pcx, a1 = P.proj_d[1] * pcy;
    const float y = rcp_refined(pcz);
    const float qu = div_step2(a0, pcz, y), qv = div_step2(a1, pcz, y);
    const float u = qu + P.proj_d[2], v = qv + P.proj_d[3];
    ok = ok && !((u < 1) || (u > P.W - 2) || (v < 1) || (v > P.H - 2));
    const int idx = ok ? ((int)(u + 0.5f) + (int)(v + 0.5f) * P.W) : 0;
    const float dm = __ldg(depth + idx);
    const int newW = WIMG ? (int)__ldg(wimg + idx) : 1;
    ok = ok && !(dm <= 0.0f);
    const float eta = dm - pcz;
    ok = ok && !(eta < -P.mu);
    const float q = div_step2(eta, P.mu, P.rcpMu);
    const float oldF = div_step1((float)(short)(lo & 0xffffu), 32767.0f, P.rcp32767);
    const float newF = (1.0f < q) ? 1.0f : q;
    const float numer = (float)w * oldF + (float)newW * newF;
    const int ws = w + newW;                                             // 1 .. 255 + maxNewW <= 510
    const float F = div_step1(numer, (float)ws, T.rcp[ws]);
    const int w2 = (ws < P.maxW) ? ws : P.maxW;
    const unsigned nlo = (lo & 0xff000000u) | (((unsigned)w2 & 0xffu) << 16) | ((unsigned)(int)(F * 32767.0f) & 0xffffu);
    lo = ok ? nlo : lo;
    return ok && !((eta > P.mu) || (fabsf(q) > 0.25f));
}

// PACKED FP32 (sm_100: add / mul / fma .rn.f32x2, SASS FADD2 / FMUL2 / FFMA2): a lane's two voxels of a slice go through the depth
// part TOGETHER, one instruction per operation for both.  Each half of a packed operation is the IEEE round-to-nearest operation
// of the scalar code, so the result is bit-identical to two calls of fuse_depth_flat; what changes is the number of issue slots
// (the kernel is issue-bound): ~40 packed instead of ~80 scalar floating-point instructions per voxel pair.
__device__ __forceinline__ float2 bc2(float x) { return make_float2(x, x); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
// div_step1 / div_step2 (exactdiv.cuh) on both halves; nb = -b
__device__ __forceinline__ float2 div_step1_2(float2 a, float2 nb, float2 y)
{
    const float2 q = mul2(a, y);
    return fma2(fma2(nb, q, a), y, q);
}
__device__ __forceinline__ float2 div_step2_2(float2 a, float2 nb, float2 y)
{
    float2 q = mul2(a, y);
    q = fma2(fma2(nb, q, a), y, q);
    return fma2(fma2(nb, q, a), y, q);
}

template <bool WIMG>
__device__ __forceinline__ void fuse_depth_flat2(unsigned &loA, unsigned &loB, float2 pcx, float2 pcy, float2 pcz, const FuseParams &P,
                                                 const float *__restrict__ depth, const unsigned char *__restrict__ wimg,
                                                 const FuseTables &T, bool &outside, bool &colourA, bool &colourB)
{
    const int wA = (int)((loA >> 16) & 0xffu), wB = (int)((loB >> 16) & 0xffu);
    bool okA = !(pcz.x <= 0.0f), okB = !(pcz.y <= 0.0f);
    if (P.stopMaxW) { okA = okA && (wA != P.maxW); okB = okB && (wB != P.maxW); }
    if (P.approximate) { okA = okA && (wA == 0); okB = okB && (wB == 0); }
    outside = outside || (okA && !div_range_ok(pcz.x)) || (okB && !div_range_ok(pcz.y));
    const float2 nz = make_float2(-pcz.x, -pcz.y);
    const float2 a0 = mul2(bc2(P.proj_d[0]), pcx), a1 = mul2(bc2(P.proj_d[1]), pcy);
    float2 y = make_float2(rcp_approx(pcz.x), rcp_approx(pcz.y));          // rcp_refined on both halves
    y = fma2(y, fma2(nz, y, bc2(1.0f)), y);
    const float2 u = add2(div_step2_2(a0, nz, y), bc2(P.proj_d[2])), v = add2(div_step2_2(a1, nz, y), bc2(P.proj_d[3]));
    okA = okA && !((u.x < 1) || (u.x > P.W - 2) || (v.x < 1) || (v.x > P.H - 2));
    okB = okB && !((u.y < 1) || (u.y > P.W - 2) || (v.y < 1) || (v.y > P.H - 2));
    const float2 uh = add2(u, bc2(0.5f)), vh = add2(v, bc2(0.5f));
    const int idxA = okA ? ((int)uh.x + (int)vh.x * P.W) : 0, idxB = okB ? ((int)uh.y + (int)vh.y * P.W) : 0;
    const float2 dm = make_float2(__ldg(depth + idxA), __ldg(depth + idxB));
    const int newWA = WIMG ? (int)__ldg(wimg + idxA) : 1, newWB = WIMG ? (int)__ldg(wimg + idxB) : 1;
    okA = okA && !(dm.x <= 0.0f); okB = okB && !(dm.y <= 0.0f);
    const float2 eta = add2(dm, nz);                                       // dm - pcz
    okA = okA && !(eta.x < -P.mu); okB = okB && !(eta.y < -P.mu);
    const float2 q = div_step2_2(eta, bc2(-P.mu), bc2(P.rcpMu));
    const float2 oldF = div_step1_2(make_float2((float)(short)(loA & 0xffffu), (float)(short)(loB & 0xffffu)), bc2(-32767.0f), bc2(P.rcp32767));
    const float2 newF = make_float2((1.0f < q.x) ? 1.0f : q.x, (1.0f < q.y) ? 1.0f : q.y);
    // w * oldF + newW * newF: the two products packed, the sums SCALAR -- ptxas contracts a packed product feeding a packed sum into
    // FFMA2 whatever --fmad says and even for explicit mul.rn.f32x2 / add.rn.f32x2 (measured: 25 % of random operands then differ
    // from the reference's two roundings); a scalar FADD is left alone.  No other packed product here feeds a sum.
    const float2 pOld = mul2(make_float2((float)wA, (float)wB), oldF), pNew = mul2(make_float2((float)newWA, (float)newWB), newF);
    const float2 numer = make_float2(__fadd_rn(pOld.x, pNew.x), __fadd_rn(pOld.y, pNew.y));
    const int wsA = wA + newWA, wsB = wB + newWB;                          // 1 .. 255 + maxNewW <= 510
    const float2 F = div_step1_2(numer, make_float2(-(float)wsA, -(float)wsB), make_float2(T.rcp[wsA], T.rcp[wsB]));
    const float2 Fs = mul2(F, bc2(32767.0f));
    const int w2A = (wsA < P.maxW) ? wsA : P.maxW, w2B = (wsB < P.maxW) ? wsB : P.maxW;
    const unsigned nloA = (loA & 0xff000000u) | (((unsigned)w2A & 0xffu) << 16) | ((unsigned)(int)Fs.x & 0xffffu);
    const unsigned nloB = (loB & 0xff000000u) | (((unsigned)w2B & 0xffu) << 16) | ((unsigned)(int)Fs.y & 0xffffu);
    loA = okA ? nloA : loA; loB = okB ? nloB : loB;
    colourA = okA && !((eta.x > P.mu) || (fabsf(q.x) > 0.25f));
    colourB = okB && !((eta.y > P.mu) || (fabsf(q.y) > 0.25f));
}

// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) of one 4 KiB voxel block global -> shared, completion on an mbarrier.
// TMA = true: each warp owns a two-stage ring of block buffers; lane 0 issues the copy of the warp's NEXT block as soon as the
// previous occupant of that stage has been written back, so a block is in flight for a whole block's worth of arithmetic
// before it is needed, without holding registers (the LDG variant prefetches to L2 and then one slice ahead into registers).
// The block is updated in place in its stage buffer and dirty 16-byte vectors are written back from there.
__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bulk_load_block(void *dstSmem, const void *srcGlobal, unsigned bytes, unsigned long long *bar)
{
    // the stage buffer was last touched through the generic proxy (LDS / STS of the previous block): order those accesses before
    // the async-proxy write
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dstSmem)), "l"(srcGlobal), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}

struct FuseTmaSmem {
    FuseTables T;
    alignas(128) uint4 stage[kFuseWarps][2][256];      // two 4 KiB block buffers per warp
    unsigned long long bar[kFuseWarps][2];
    unsigned short queue[kFuseWarps][kBlock3];
    unsigned dirty[kFuseWarps][8];
};

__device__ __forceinline__ uint4 *fuse_stage(FuseWarpSmem &S, int warp, unsigned) { return S.vox[warp]; }
__device__ __forceinline__ uint4 *fuse_stage(FuseTmaSmem &S, int warp, unsigned s) { return S.stage[warp][s]; }
__device__ __forceinline__ unsigned long long &fuse_bar(FuseTmaSmem &S, int warp, unsigned s) { return S.bar[warp][s]; }
__device__ __forceinline__ unsigned long long &fuse_bar(FuseWarpSmem &S, int, unsigned) { return *reinterpret_cast<unsigned long long *>(S.dirty); }   // never used

template <bool WIMG, int MINB, bool TMA, bool PACKED>
__global__ void __launch_bounds__(kFuseWarps * 32, MINB) k_fuse_warp_flat(FuseParams P, dsb200_voxel *voxels,
                                                                        const dsb200_hash_entry *__restrict__ hash,
                                                                        const int *__restrict__ ids, const int *__restrict__ countPtr,
                                                                        const float *__restrict__ depth, const uchar4 *__restrict__ rgb,
                                                                        const unsigned char *__restrict__ wimg, unsigned *shadow2)
{
    typedef typename std::conditional<TMA, FuseTmaSmem, FuseWarpSmem>::type Smem;
    __shared__ Smem S;
    const int n = *countPtr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    fill_tables(S.T);
    if (lane < 8) S.dirty[warp][lane] = 0;
    if (TMA && lane < 2) mbar_init(&fuse_bar(S, warp, lane), 1);
    if (TMA) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const FuseTables &T = S.T;
    unsigned short *sQueue = S.queue[warp];
    unsigned *sDirty = S.dirty[warp];
    const int lx = (lane & 3) * 2, ly = lane >> 2;
    const int stride = gridDim.x * kFuseWarps;
    const HashEntry none = {0, 0, 0, 0, 0, -2, 0};
    int i = blockIdx.x * kFuseWarps + warp;
    int slot2 = (i + 2 * stride < n) ? __ldg(ids + i + 2 * stride) : -1;
    HashEntry e1 = (i + stride < n) ? load_entry(hash, __ldg(ids + i + stride)) : none;
    HashEntry e0 = (i < n) ? load_entry(hash, __ldg(ids + i)) : none;
    unsigned cur = 0, phase = 0;          // TMA: stage of the current block; bit s of `phase` = parity of stage s's next completion
    if (TMA && e0.ptr >= 0 && lane == 0)
        bulk_load_block(fuse_stage(S, warp, 0), voxels + (size_t)e0.ptr * kBlock3, 4096u, &fuse_bar(S, warp, 0));
    for (; i < n; i += stride) {
        const int slot3 = (i + 3 * stride < n) ? __ldg(ids + i + 3 * stride) : -1;
        const HashEntry e2 = (slot2 >= 0) ? load_entry(hash, slot2) : none;
        if (TMA) {
            // the other stage's block was written back in the previous iteration: start the next block's copy into it
            if (e1.ptr >= 0 && lane == 0)
                bulk_load_block(fuse_stage(S, warp, cur ^ 1u), voxels + (size_t)e1.ptr * kBlock3, 4096u, &fuse_bar(S, warp, cur ^ 1u));
        } else if (e1.ptr >= 0) prefetch_l2(reinterpret_cast<const char *>(voxels + (size_t)e1.ptr * kBlock3) + lane * 128);
        uint4 *sVox = fuse_stage(S, warp, cur);
        if (e0.ptr >= 0) {
            const HashEntry e = e0;
            uint4 *base = reinterpret_cast<uint4 *>(voxels + (size_t)e.ptr * kBlock3);
            if (TMA) { mbar_wait(&fuse_bar(S, warp, cur), (phase >> cur) & 1u); phase ^= 1u << cur; }
            const uint4 *src = TMA ? sVox : base;      // where the block's current contents are read from
            uint4 raw = src[lane];
            const int gx = e.px * kBlock + lx;
            const float ptx0 = (float)gx * P.voxelSize, ptx1 = (float)(gx + 1) * P.voxelSize;
            const float pty = (float)(e.py * kBlock + ly) * P.voxelSize;
            const float ax0 = P.M_d.m[0] * ptx0 + P.M_d.m[4] * pty, ay0 = P.M_d.m[1] * ptx0 + P.M_d.m[5] * pty, az0 = P.M_d.m[2] * ptx0 + P.M_d.m[6] * pty;
            const float ax1 = P.M_d.m[0] * ptx1 + P.M_d.m[4] * pty, ay1 = P.M_d.m[1] * ptx1 + P.M_d.m[5] * pty, az1 = P.M_d.m[2] * ptx1 + P.M_d.m[6] * pty;
            unsigned dirty = 0;
            int qn = 0;
#pragma unroll 2
            for (int z = 0; z < kBlock; z++) {
                uint4 nxt = raw;
                if (z + 1 < kBlock) nxt = src[(z + 1) * 32 + lane];
                const float ptz = (float)(e.pz * kBlock + z) * P.voxelSize;
                const float bx = P.M_d.m[8] * ptz, by = P.M_d.m[9] * ptz, bz = P.M_d.m[10] * ptz;
                unsigned lo0 = raw.x, lo1 = raw.z;
                bool outside = false;
                bool c0, c1;
                if (PACKED)
                    fuse_depth_flat2<WIMG>(lo0, lo1, add2(add2(make_float2(ax0, ax1), bc2(bx)), bc2(P.M_d.m[12])),
                                           add2(add2(make_float2(ay0, ay1), bc2(by)), bc2(P.M_d.m[13])),
                                           add2(add2(make_float2(az0, az1), bc2(bz)), bc2(P.M_d.m[14])), P, depth, wimg, T, outside, c0, c1);
                else {
                    c0 = fuse_depth_flat<WIMG>(lo0, (ax0 + bx) + P.M_d.m[12], (ay0 + by) + P.M_d.m[13], (az0 + bz) + P.M_d.m[14], P, depth, wimg, T, outside);
                    c1 = fuse_depth_flat<WIMG>(lo1, (ax1 + bx) + P.M_d.m[12], (ay1 + by) + P.M_d.m[13], (az1 + bz) + P.M_d.m[14], P, depth, wimg, T, outside);
                }
                if (__any_sync(0xffffffffu, outside)) {          // (never in practice) the slice again, general code
                    VoxelRegs v0 = unpack(raw.x, raw.y), v1 = unpack(raw.z, raw.w);
                    c0 = c1 = false;
                    if (!((P.stopMaxW && v0.w == P.maxW) || (P.approximate && v0.w != 0))) {
                        Vec4 pc = {(ax0 + bx) + P.M_d.m[12], (ay0 + by) + P.M_d.m[13], (az0 + bz) + P.M_d.m[14], 1.0f};
                        c0 = fuse_depth_pc<false, true>(v0, pc, P, depth, T);
                    }
                    if (!((P.stopMaxW && v1.w == P.maxW) || (P.approximate && v1.w != 0))) {
                        Vec4 pc = {(ax1 + bx) + P.M_d.m[12], (ay1 + by) + P.M_d.m[13], (az1 + bz) + P.M_d.m[14], 1.0f};
                        c1 = fuse_depth_pc<false, true>(v1, pc, P, depth, T);
                    }
                    uint4 o = raw;
                    pack(v0, o.x, o.y); pack(v1, o.z, o.w);
                    lo0 = o.x; lo1 = o.z;
                }
                if (lo0 != raw.x || lo1 != raw.z) {
                    dirty |= 1u << z;
                    if (shadow2) shadow2[(size_t)e.ptr * 256 + z * 32 + lane] = (lo0 & 0xffffu) | (lo1 << 16);
                }
                raw.x = lo0; raw.z = lo1;
                sVox[z * 32 + lane] = raw;
                const unsigned m0 = __ballot_sync(0xffffffffu, c0), m1 = __ballot_sync(0xffffffffu, c1);
                if (c0) sQueue[qn + __popc(m0 & lt)] = (unsigned short)(z * 64 + 2 * lane);
                if (c1) sQueue[qn + __popc(m0) + __popc(m1 & lt)] = (unsigned short)(z * 64 + 2 * lane + 1);
                qn += __popc(m0) + __popc(m1);
                raw = nxt;
            }
            __syncwarp();
            if (qn) {
                uint2 *sv = reinterpret_cast<uint2 *>(sVox);
                for (int k = lane; k < qn; k += 32) {
                    const int vi = sQueue[k];
                    uint2 w = sv[vi];
                    const uint2 before = w;
                    VoxelRegs vx = unpack(w.x, w.y);
                    fuse_colour<false, true>(vx, (float)(e.px * kBlock + (vi & 7)) * P.voxelSize, (float)(e.py * kBlock + ((vi >> 3) & 7)) * P.voxelSize,
                                             (float)(e.pz * kBlock + (vi >> 6)) * P.voxelSize, P, rgb, T);
                    pack(vx, w.x, w.y);
                    if (w.x != before.x || w.y != before.y) {
                        sv[vi] = w;
                        atomicOr(&sDirty[vi >> 6], 1u << ((vi >> 1) & 31));
                    }
                }
                __syncwarp();
            }
#pragma unroll
            for (int z = 0; z < kBlock; z++) {
                const unsigned d = qn ? sDirty[z] : 0u;
                if (((dirty >> z) & 1u) || ((d >> lane) & 1u)) base[z * 32 + lane] = sVox[z * 32 + lane];
            }
            if (qn) {
                __syncwarp();
                if (lane < 8) sDirty[lane] = 0;
            }
            __syncwarp();
        }
        if (TMA) cur ^= 1u;                 // stages alternate with the warp's blocks, processed or not
        e0 = e1; e1 = e2; slot2 = slot3;
    }
}

void launch_fuse(cudaStream_t st, const FuseParams &P, bool deintegrate, dsb200_voxel *voxels,
                 const dsb200_hash_entry *hash, const int *ids, const int *count, const float *depth, const uint8_t *rgb,
                 short *sdfShadow, unsigned char *weightScratch)
{
    unsigned *shadow2 = reinterpret_cast<unsigned *>(sdfShadow);
    const uchar4 *rgb4 = reinterpret_cast<const uchar4 *>(rgb);
    const int variant = tunables().fuse_kernel;   // 1: round-1 warp-per-block kernel; 2: + straight-line depth part; 3: + TMA bulk block loads; 4: + packed FP32 (default)
    if (variant >= 2 && !deintegrate && P.fastDiv && P.maxNewW >= 1 && P.maxNewW <= 255 && (!P.depthWeighting || weightScratch)) {
        const int g = kNumSMs * tunables().fuse_ctas;
        const int minb = tunables().fuse_minb;     // register budget: 65536 / (128 * minb)
#define DSB_FLAT4(W_, M_, T_, K_) DSB_LAUNCH((k_fuse_warp_flat<W_, M_, T_, K_>), g, kFuseWarps * 32, 0, st, P, voxels, hash, ids, count, depth, rgb4, W_ ? weightScratch : nullptr, shadow2)
#define DSB_FLAT(W_, M_, T_) DSB_FLAT4(W_, M_, T_, false)
        if (P.depthWeighting) DSB_LAUNCH(k_depth_weights, kNumSMs * 4, 256, 0, st, depth, weightScratch, P.W * P.H, P.maxDistance, P.maxNewW);
        if (variant >= 4) {                // TMA bulk loads + packed FP32 depth part (FADD2 / FMUL2 / FFMA2)
            if (P.depthWeighting) { if (minb >= 5) DSB_FLAT4(true, 5, true, true); else DSB_FLAT4(true, 4, true, true); }
            else { if (minb >= 5) DSB_FLAT4(false, 5, true, true); else DSB_FLAT4(false, 4, true, true); }
        } else if (variant == 3) {         // TMA bulk loads: 39 KiB of shared memory per CTA -> at most 5 CTAs per SM
            if (P.depthWeighting) { if (minb >= 5) DSB_FLAT(true, 5, true); else DSB_FLAT(true, 4, true); }
            else { if (minb >= 5) DSB_FLAT(false, 5, true); else DSB_FLAT(false, 4, true); }
        } else if (P.depthWeighting) {
            if (minb >= 8) DSB_FLAT(true, 8, false); else if (minb >= 6) DSB_FLAT(true, 6, false); else DSB_FLAT(true, 4, false);
        } else {
            if (minb >= 8) DSB_FLAT(false, 8, false); else if (minb >= 6) DSB_FLAT(false, 6, false); else DSB_FLAT(false, 4, false);
        }
#undef DSB_FLAT
#undef DSB_FLAT4
        return;
    }
    {
        const int g = kNumSMs * 8;
        if (deintegrate) {
            if (P.fastDiv) DSB_LAUNCH((k_fuse_warp<true, true, 8>), g, kFuseWarps * 32, 0, st, P, voxels, hash, ids, count, depth, rgb4, shadow2);
            else DSB_LAUNCH((k_fuse_warp<true, false, 8>), g, kFuseWarps * 32, 0, st, P, voxels, hash, ids, count, depth, rgb4, shadow2);
        } else {
            if (P.fastDiv) DSB_LAUNCH((k_fuse_warp<false, true, 8>), g, kFuseWarps * 32, 0, st, P, voxels, hash, ids, count, depth, rgb4, shadow2);
            else DSB_LAUNCH((k_fuse_warp<false, false, 8>), g, kFuseWarps * 32, 0, st, P, voxels, hash, ids, count, depth, rgb4, shadow2);
        }
    }
}

// ---------------------------------------------------------------- exact-division self test
__device__ __forceinline__ unsigned long long splitmix(unsigned long long &x)
{
    unsigned long long z = (x += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
// float with a uniformly random significand and sign, exponent uniform in [loExp, hiExp]
__device__ __forceinline__ float random_float(unsigned long long &st, int loExp, int hiExp)
{
    unsigned long long r = splitmix(st);
    unsigned mant = (unsigned)r & 0x7fffffu, sign = (unsigned)(r >> 23) & 1u;
    int ex = loExp + (int)((r >> 24) % (unsigned long long)(hiExp - loExp + 1));
    return __uint_as_float((sign << 31) | ((unsigned)(ex + 127) << 23) | mant);
}
__device__ __forceinline__ bool same_bits(float a, float b) { return __float_as_uint(a) == __float_as_uint(b); }

__global__ void __launch_bounds__(256) k_selftest_division(unsigned long long n, unsigned long long seed, unsigned long long *mismatches)
{
    typedef Exact<true> X;
    unsigned long long bad = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned long long st = seed * 0x2545f4914f6cdd1dull + i * 0x9e3779b97f4a7c15ull;
        // (1) shared run-time divisor: numerators in [2^-84, 2^68] or 0, divisor anywhere in [2^-30, 2^30] (window + fall-back)
        float a0 = random_float(st, -84, 68), a1 = random_float(st, -84, 68), b = random_float(st, -30, 30);
        if ((i & 63) == 0) a0 = 0.0f;
        if ((i & 1023) == 1) b = 0.0f;
        float q0, q1;
        X::pair(a0, a1, b, q0, q1);
        bad += !same_bits(q0, a0 / b); bad += !same_bits(q1, a1 / b);
        bad += !same_bits(X::one(a1, b), a1 / b);
        // realistic projection operands: z in [0.05, 400], numerators up to 1e6
        float z = fabsf(random_float(st, -5, 8)), nx = random_float(st, -20, 20);
        X::pair(nx, a1, z, q0, q1);
        bad += !same_bits(q0, nx / z);
        // (2) host reciprocal, two steps: any divisor whose significand is not all ones
        float mu = fabsf(random_float(st, -20, 20));
        if ((__float_as_uint(mu) & 0x7fffffu) == 0x7fffffu) mu = 0.08f;
        float eta = random_float(st, -40, 10);
        bad += !same_bits(X::by_host_rcp(eta, mu, 1.0f / mu), eta / mu);
        // (3) integer divisors with one correction step
        unsigned long long r = splitmix(st);
        float w = (float)(1 + (int)(r % 511ull)), num = random_float(st, -60, 20);
        bad += !same_bits(X::by_int(num, w, 1.0f / w), num / w);
        bad += !same_bits(X::by_int(num, 255.0f, 1.0f / 255.0f), num / 255.0f);
        float s16 = (float)((int)((r >> 16) & 0xffffull) - 32768);
        bad += !same_bits(X::by_int(s16, 32767.0f, 1.0f / 32767.0f), s16 / 32767.0f);
        // (4) depth weighting rounding: trunc + frac test == (int)((double)x + 0.5)
        float x = fabsf(random_float(st, -10, 22));
        float tr = truncf(x); int fastW = (int)tr + ((x - tr >= 0.5f) ? 1 : 0);
        bad += fastW != (int)((double)x + 0.5);
    }
    for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(mismatches, bad);
}

void launch_selftest_division(cudaStream_t st, unsigned long long n, unsigned long long seed, unsigned long long *mismatches)
{
    DSB_LAUNCH(k_selftest_division, kNumSMs * 8, 256, 0, st, n, seed, mismatches);
}

// ResetScene (CPU .cpp:25-45; CUDA .cu:111-148): pure store bandwidth, 16-byte stores.
__global__ void __launch_bounds__(256) k_reset_scene(uint4 *voxels2, size_t numVoxelPairs, int *hashWords, int total,
                                                     int *allocList, int numBlocks, int *excessList, int excessSize,
                                                     uint8_t *visType, unsigned *reqKey, Counters *counters, uint4 *shadow8)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    const uint4 blank = make_uint4(0x00007fffu, 0u, 0x00007fffu, 0u);   // {sdf 32767, w 0, clr 0, wc 0} x2
    for (size_t i = tid; i < numVoxelPairs; i += stride) voxels2[i] = blank;
    if (shadow8) for (size_t i = tid; i < numVoxelPairs / 4; i += stride) shadow8[i] = make_uint4(0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu);
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
        counters->noVisible = 0; counters->noVisiblePrev = 0; counters->renderBlocks = 0; counters->freedBlocks = 0; counters->snapshotOverflow = 0; counters->totalFreed = 0;
    }
}

void launch_reset_scene(cudaStream_t st, dsb200_voxel *voxels, size_t numVoxels, dsb200_hash_entry *hash, int total,
                        int *allocList, int numBlocks, int *excessList, int excessSize, uint8_t *visTypeOrNull,
                        unsigned *reqKey, Counters *counters, short *sdfShadow)
{
    DSB_LAUNCH(k_reset_scene, kNumSMs * 16, 256, 0, st, reinterpret_cast<uint4 *>(voxels), numVoxels / 2,
               reinterpret_cast<int *>(hash), total, allocList, numBlocks, excessList, excessSize, visTypeOrNull, reqKey,
               counters, reinterpret_cast<uint4 *>(sdfShadow));
}

// sdf shadow <- voxel array (a scene this engine has not been maintaining the shadow for)
__global__ void __launch_bounds__(256) k_rebuild_shadow(const uint4 *__restrict__ voxels2, size_t numVoxelPairs, unsigned *shadow2)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < numVoxelPairs; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = voxels2[i];
        shadow2[i] = (v.x & 0xffffu) | (v.z << 16);
    }
}
void launch_rebuild_shadow(cudaStream_t st, const dsb200_voxel *voxels, size_t numVoxels, short *sdfShadow)
{
    DSB_LAUNCH(k_rebuild_shadow, kNumSMs * 16, 256, 0, st, reinterpret_cast<const uint4 *>(voxels), numVoxels / 2,
               reinterpret_cast<unsigned *>(sdfShadow));
}

} // namespace dsb
