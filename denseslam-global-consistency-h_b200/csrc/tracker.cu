// tracker.cu -- the data-parallel half of the ICP depth tracker.
//
// The reference splits ITMDepthTracker into a host Levenberg-Marquardt loop (ITMLib/Engine/ITMDepthTracker.cpp:153-211,
// device independent) and one device-specific virtual, ComputeGandH, plus the low-level engine's FilterSubsampleWithHoles
// that builds the depth pyramid (ITMDepthTracker.cpp:66-76).  These two are what this file provides:
//
//   k_filter_subsample_holes   DeviceAgnostic/ITMLowLevelEngine.h:26-47 (float image, holes = depth <= 0)
//   k_tracker_gh<SHORT, ROT>   DeviceAgnostic/ITMDepthTracker.h:9-104 per pixel (bit-exact: same operation order, no
//                              contraction) + the sums of DeviceSpecific/CPU/ITMDepthTracker_CPU.cpp:13-79.
//
// Summation: the reference CPU engine adds the per-pixel terms serially in float, its CUDA engine adds them in a
// block tree followed by float atomics (order not defined).  Here every term is accumulated in double along a FIXED
// tree: grid-stride pixels per thread, xor-shuffle per warp, warps in order per CTA, CTA partials in eight interleaved
// chains combined in order by the last CTA to finish.  The result does not depend on scheduling, and is the correctly-rounded sum to about 1e-13 relative,
// which the parity tests check against the oracle's double-precision serial sum (and, loosely, its float serial sum).
#include <mutex>

#include "kernels.cuh"

namespace dsb {

namespace {

__global__ void __launch_bounds__(256) k_filter_subsample_holes(const float *__restrict__ in, int oldW, float *__restrict__ out,
                                                                int newW, int newH)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= newW || y >= newH) return;
    const float2 r0 = *reinterpret_cast<const float2 *>(in + (size_t)(2 * y) * oldW + 2 * x);
    const float2 r1 = *reinterpret_cast<const float2 *>(in + (size_t)(2 * y + 1) * oldW + 2 * x);
    float acc = 0.0f, good = 0.0f;
    if (r0.x > 0.0f) { acc += r0.x; good++; }
    if (r0.y > 0.0f) { acc += r0.y; good++; }
    if (r1.x > 0.0f) { acc += r1.x; good++; }
    if (r1.y > 0.0f) { acc += r1.y; good++; }
    if (good > 0) acc /= good;
    out[(size_t)y * newW + x] = acc;
}

// odd input width: the float2 loads above would be misaligned on odd rows
__global__ void __launch_bounds__(256) k_filter_subsample_holes_scalar(const float *__restrict__ in, int oldW,
                                                                       float *__restrict__ out, int newW, int newH)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= newW || y >= newH) return;
    const float *p = in + (size_t)(2 * y) * oldW + 2 * x;
    float acc = 0.0f, good = 0.0f, v;
    v = p[0]; if (v > 0.0f) { acc += v; good++; }
    v = p[1]; if (v > 0.0f) { acc += v; good++; }
    v = p[oldW]; if (v > 0.0f) { acc += v; good++; }
    v = p[oldW + 1]; if (v > 0.0f) { acc += v; good++; }
    if (good > 0) acc /= good;
    out[(size_t)y * newW + x] = acc;
}

// interpolateBilinear_withHoles, DeviceAgnostic/ITMPixelUtils.h:41-71 (Vector4f source; the (short) casts are the
// reference's).  Only x, y, z of the result are used by the tracker besides the hole flag.
__device__ __forceinline__ bool bilinear_holes(const float4 *__restrict__ src, float px, float py, int W, Vec3 &r)
{
    const short ix = (short)floorf(px), iy = (short)floorf(py);
    const float dx = px - (float)ix, dy = py - (float)iy;
    const float4 a = __ldg(src + ix + iy * W), b = __ldg(src + (ix + 1) + iy * W);
    const float4 c = __ldg(src + ix + (iy + 1) * W), d = __ldg(src + (ix + 1) + (iy + 1) * W);
    if (a.w < 0 || b.w < 0 || c.w < 0 || d.w < 0) return false;
    // the reference writes a*(1-dx)*(1-dy) + b*dx*(1-dy) + c*(1-dx)*dy + d*dx*dy, products left to right
    r.x = a.x * (1.0f - dx) * (1.0f - dy) + b.x * dx * (1.0f - dy) + c.x * (1.0f - dx) * dy + d.x * dx * dy;
    r.y = a.y * (1.0f - dx) * (1.0f - dy) + b.y * dx * (1.0f - dy) + c.y * (1.0f - dx) * dy + d.y * dx * dy;
    r.z = a.z * (1.0f - dx) * (1.0f - dy) + b.z * dx * (1.0f - dy) + c.z * (1.0f - dx) * dy + d.z * dx * dy;
    return true;
}

} // namespace

struct TrackerParams {
    Mat4 approxInvPose, scenePose;
    float viewIntr[4], sceneIntr[4];
    float distThresh;
    int viewW, viewH, sceneW, sceneH;
};

namespace {

constexpr int kGhThreads = 256;
constexpr int kGhSlots = 32;        // 21 hessian + 6 nabla + f + count, padded

// computePerPointGH_Depth_Ab, DeviceAgnostic/ITMDepthTracker.h:9-74
template <bool SHORT, bool ROT>
__device__ __forceinline__ bool per_point_ab(const TrackerParams &P, const float4 *__restrict__ points,
                                             const float4 *__restrict__ normals, int x, int y, float depth, float *A, float &b)
{
    if (depth <= 1e-8f) return false;
    Vec4 p;
    p.x = depth * (((float)x - P.viewIntr[2]) / P.viewIntr[0]);
    p.y = depth * (((float)y - P.viewIntr[3]) / P.viewIntr[1]);
    p.z = depth; p.w = 1.0f;
    p = mul(P.approxInvPose, p);
    p.w = 1.0f;
    const Vec4 q = mul(P.scenePose, p);
    if (q.z <= 0.0f) return false;
    const float u = P.sceneIntr[0] * q.x / q.z + P.sceneIntr[2];
    const float v = P.sceneIntr[1] * q.y / q.z + P.sceneIntr[3];
    if (!((u >= 0.0f) && (u <= (float)(P.sceneW - 2)) && (v >= 0.0f) && (v <= (float)(P.sceneH - 2)))) return false;
    Vec3 c;
    if (!bilinear_holes(points, u, v, P.sceneW, c)) return false;
    // the interpolated w of the points map is >= 0 whenever its four corners are: nothing more to test (:35)
    const float dx = c.x - p.x, dy = c.y - p.y, dz = c.z - p.z;
    const float dist = dx * dx + dy * dy + dz * dz;
    if (dist > P.distThresh) return false;
    Vec3 n;
    if (!bilinear_holes(normals, u, v, P.sceneW, n)) { n.x = 0; n.y = 0; n.z = 0; }   // holes give a zero normal (:57-59 of PixelUtils)
    b = n.x * dx + n.y * dy + n.z * dz;
    if (SHORT && !ROT) { A[0] = n.x; A[1] = n.y; A[2] = n.z; }
    else {
        A[0] = +p.z * n.y - p.y * n.z;
        A[1] = -p.z * n.x + p.x * n.z;
        A[2] = +p.y * n.x - p.x * n.y;
        if (!SHORT) { A[3] = n.x; A[4] = n.y; A[5] = n.z; }
    }
    return true;
}

template <bool SHORT, bool ROT>
__global__ void __launch_bounds__(kGhThreads) k_tracker_gh(TrackerParams P, const float *__restrict__ depth,
                                                           const float4 *__restrict__ points, const float4 *__restrict__ normals,
                                                           double *__restrict__ partials, unsigned *__restrict__ ticket,
                                                           double *__restrict__ result, double seq)
{
    constexpr int NP = SHORT ? 3 : 6, NH = SHORT ? 6 : 21, NV = NH + NP + 1;
    double acc[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) acc[i] = 0.0;
    int count = 0;

    const int n = P.viewW * P.viewH;
    for (int i = blockIdx.x * kGhThreads + threadIdx.x; i < n; i += gridDim.x * kGhThreads) {
        const int y = i / P.viewW, x = i - y * P.viewW;
        float A[NP], b;
        if (!per_point_ab<SHORT, ROT>(P, points, normals, x, y, __ldg(depth + i), A, b)) continue;
        count++;
        // computePerPointGH_Depth, DeviceAgnostic/ITMDepthTracker.h:76-104: float products, summed in double
        int k = 0;
#pragma unroll
        for (int r = 0; r < NP; r++) {
#pragma unroll
            for (int c = 0; c <= r; c++) acc[k++] += (double)(A[r] * A[c]);
        }
#pragma unroll
        for (int r = 0; r < NP; r++) acc[NH + r] += (double)(b * A[r]);
        acc[NH + NP] += (double)(b * b);
    }

    // warp tree (xor shuffles: every lane ends with the same value), then warps in order, then CTAs in order
#pragma unroll
    for (int i = 0; i < NV; i++) {
#pragma unroll
        for (int o = 16; o; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
    }
    count = __reduce_add_sync(0xffffffffu, count);

    __shared__ double sh[kGhThreads / 32][kGhSlots];
    __shared__ bool last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; i++) sh[warp][i] = acc[i];
        sh[warp][NV] = (double)count;
    }
    __syncthreads();
    if (threadIdx.x <= NV) {
        double s = 0.0;
        for (int w = 0; w < kGhThreads / 32; w++) s += sh[w][threadIdx.x];
        partials[(size_t)blockIdx.x * kGhSlots + threadIdx.x] = s;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!last) return;
    __threadfence();
    // CTA partials in a fixed order, eight interleaved chains per slot (independent loads in flight), chains combined in order
    {
        const int slot = threadIdx.x & 31, chain = threadIdx.x >> 5;
        double s = 0.0;
        if (slot <= NV) {
#pragma unroll 4
            for (unsigned cta = chain; cta < gridDim.x; cta += kGhThreads / 32) s += __ldcg(partials + (size_t)cta * kGhSlots + slot);
        }
        sh[chain][slot] = s;
    }
    __syncthreads();
    if (threadIdx.x <= NV) {
        double s = 0.0;
        for (int c = 0; c < kGhThreads / 32; c++) s += sh[c][threadIdx.x];
        result[threadIdx.x] = s;           // mapped host memory
        __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *ticket = 0;                       // re-armed for the next launch on this stream
        // sequence number last: the host polls this word instead of paying a stream synchronisation per evaluation
        *reinterpret_cast<volatile double *>(result + kGhSlots - 1) = seq;      // ordered after the sums by the fences + barrier above
    }
}

} // namespace

void launch_filter_subsample_holes(cudaStream_t st, const float *in, int oldW, int oldH, float *out)
{
    const int newW = oldW / 2, newH = oldH / 2;
    if (newW <= 0 || newH <= 0) return;
    dim3 block(32, 8), grid(div_up(newW, 32), div_up(newH, 8));
    if ((oldW & 1) == 0 && ((uintptr_t)in & 7) == 0) DSB_LAUNCH(k_filter_subsample_holes, grid, block, 0, st, in, oldW, out, newW, newH);
    else DSB_LAUNCH(k_filter_subsample_holes_scalar, grid, block, 0, st, in, oldW, out, newW, newH);
}

int tracker_gh_max_ctas() { return kNumSMs * 2; }   // two 256-thread CTAs per SM fit the 6-parameter variant's registers
int tracker_gh_slots() { return kGhSlots; }

// result: [0, NH) hessian lower triangle (row-major r, c <= r), [NH, NH+NP) nabla, [NH+NP] sum of b*b, [NH+NP+1] valid count
void launch_tracker_gh(cudaStream_t st, int iterationType, const float *approxInvPose, const float *scenePose,
                       const float *viewIntr, const float *sceneIntr, float distThresh, const float *depth, int viewW, int viewH,
                       const float *points, const float *normals, int sceneW, int sceneH, double *partials, unsigned *ticket,
                       double *result, double seq)
{
    TrackerParams P;
    for (int i = 0; i < 16; i++) { P.approxInvPose.m[i] = approxInvPose[i]; P.scenePose.m[i] = scenePose[i]; }
    for (int i = 0; i < 4; i++) { P.viewIntr[i] = viewIntr[i]; P.sceneIntr[i] = sceneIntr[i]; }
    P.distThresh = distThresh; P.viewW = viewW; P.viewH = viewH; P.sceneW = sceneW; P.sceneH = sceneH;
    int grid = div_up(viewW * viewH, kGhThreads);
    if (grid > tracker_gh_max_ctas()) grid = tracker_gh_max_ctas();
    const float4 *pts = reinterpret_cast<const float4 *>(points), *nrm = reinterpret_cast<const float4 *>(normals);
    if (iterationType == DSB200_TRACKER_ITERATION_ROTATION)
        DSB_LAUNCH((k_tracker_gh<true, true>), grid, kGhThreads, 0, st, P, depth, pts, nrm, partials, ticket, result, seq);
    else if (iterationType == DSB200_TRACKER_ITERATION_TRANSLATION)
        DSB_LAUNCH((k_tracker_gh<true, false>), grid, kGhThreads, 0, st, P, depth, pts, nrm, partials, ticket, result, seq);
    else
        DSB_LAUNCH((k_tracker_gh<false, false>), grid, kGhThreads, 0, st, P, depth, pts, nrm, partials, ticket, result, seq);
}


// ---------------------------------------------------------------- the allocation-free evaluation
// dsb200_tracker_compute_g_and_h_static (include/dsb200.h): the same kernel, its reduction scratch (CTA partials + ticket) in a
// file-scope device array instead of a dsb200_tracker's cudaMalloc'ed buffers, the sums written to a caller-owned DEVICE buffer
// instead of the tracker's mapped host memory.  No counterpart in the reference; added for
// callers that cannot allocate (CUDA graph capture, serving loops).  Every GPU that loads libdsb200 holds the array until exit.
namespace {
constexpr int kStaticMaxCtas = kNumSMs * 2;                                   // = tracker_gh_max_ctas(): the grid never exceeds it
constexpr size_t kStaticPartialBytes = (size_t)kStaticMaxCtas * kGhSlots * sizeof(double);   // 296 * 32 * 8 = 75,776 B
constexpr size_t kStaticScratchBytes = kStaticPartialBytes + 256;            // + the ticket, on its own 256-byte line
constexpr int kStaticMaxDevices = 64;
}
static __device__ __align__(256) unsigned char g_tracker_static_scratch[kStaticScratchBytes];

int fail_with(int code, const char *msg);      // capi.cu: sets dsb200_last_error

namespace {
// In host code `g_tracker_static_scratch` names a host-side stand-in, not device memory: the device address is asked for once
// per device, with that device current, and cached.
std::mutex g_static_mutex;
unsigned char *g_static_ptr[kStaticMaxDevices] = {};
int static_scratch_for_current_device(unsigned char **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_with(DSB200_ERR_CUDA, cudaGetErrorString(e));
    if (dev < 0 || dev >= kStaticMaxDevices) return fail_with(DSB200_ERR_INVALID, "tracker_compute_g_and_h_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_tracker_static_scratch);
        if (e != cudaSuccess || !p) return fail_with(DSB200_ERR_CUDA, e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned null");
        g_static_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_static_ptr[dev];
    return DSB200_OK;
}
} // namespace

} // namespace dsb

extern "C" int dsb200_tracker_compute_g_and_h_static(const float *depth, int32_t vW, int32_t vH, const float *vIntr, const float *points,
                                                     const float *normals, int32_t sW, int32_t sH, const float *sIntr,
                                                     const float *scenePose, int32_t iterationType, float distThresh,
                                                     const float *approxInvPose, double *sums, void *stream)
{
    using namespace dsb;
    if (!depth || !vIntr || !points || !normals || !sIntr || !scenePose || !approxInvPose || !sums)
        return fail_with(DSB200_ERR_INVALID, "tracker_compute_g_and_h_static: null argument");
    if (vW <= 0 || vH <= 0 || sW < 2 || sH < 2 || sW > 32767 || sH > 32767 || ((uintptr_t)points & 15) || ((uintptr_t)normals & 15) ||
        ((uintptr_t)sums & 7))
        return fail_with(DSB200_ERR_INVALID, "tracker_compute_g_and_h_static: bad image size or unaligned maps");
    if ((long long)vW * vH > (long long)DSB200_TRACKER_STATIC_MAX_PIXELS)
        return fail_with(DSB200_ERR_INVALID, "tracker_compute_g_and_h_static: view image larger than DSB200_TRACKER_STATIC_MAX_PIXELS (4,194,304 pixels)");
    if (iterationType != DSB200_TRACKER_ITERATION_ROTATION && iterationType != DSB200_TRACKER_ITERATION_TRANSLATION &&
        iterationType != DSB200_TRACKER_ITERATION_BOTH)
        return fail_with(DSB200_ERR_INVALID, "tracker_compute_g_and_h_static: unknown iteration type");
    static_assert(kStaticMaxCtas * kGhSlots * sizeof(double) + 256 == DSB200_TRACKER_STATIC_SCRATCH_BYTES, "header states the array's size");
    unsigned char *scratch = nullptr;
    { int rc = static_scratch_for_current_device(&scratch); if (rc) return rc; }
    cudaStream_t st = (cudaStream_t)stream;
    double *partials = reinterpret_cast<double *>(scratch);
    unsigned *ticket = reinterpret_cast<unsigned *>(scratch + kStaticPartialBytes);
    // the array keeps the last call's contents: the partials of the CTAs this launch uses are all overwritten before they are
    // read; the ticket must start at 0 (the kernel re-arms it, but an aborted earlier launch would not have)
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(sums, 0, kGhSlots * sizeof(double), st);    // the slots this iteration type does not use read 0
    if (e != cudaSuccess) return fail_with(DSB200_ERR_CUDA, cudaGetErrorString(e));
    launch_tracker_gh(st, iterationType, approxInvPose, scenePose, vIntr, sIntr, distThresh, depth, vW, vH, points, normals, sW, sH,
                      partials, ticket, sums, 0.0);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_with(DSB200_ERR_CUDA, cudaGetErrorString(e));
    return DSB200_OK;
}
