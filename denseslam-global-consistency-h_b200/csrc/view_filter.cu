// view_filter.cu -- the view builder's optional bilateral depth filter.
//
//   k_filter_depth   filterDepth, DeviceAgnostic/ITMViewBuilder.h:30-53, applied like ITMViewBuilder_CPU::DepthFiltering
//                    (CPU/ITMViewBuilder_CPU.cpp:119-130): output cleared, interior [2, W-2) x [2, H-2) filtered;
//   five passes ping-ponging between the view's depth image and a scratch image, result copied back (UpdateView :49-58).
//
// Arithmetic follows the reference expression by expression (float, no contraction, IEEE division and square root).  The
// one library call, exp(float), resolves to glibc's expf on the reference's host.  That function is NOT correctly rounded
// (a degree-3 polynomial in double after a 32-entry table reduction, ~2^-33 relative error), so a more accurate exp disagrees
// with it on about one weight in 10^4, and five passes spread every such difference over a 21x21 neighbourhood.  expf_host()
// below therefore restates glibc's published algorithm (sysdeps/ieee754/flt-32/e_expf.c, glibc >= 2.27: table 2^(i/32),
// InvLn2N, the three scaled coefficients, the 0x1.8p52 shift trick) with the FMA contraction of its x86-64 FMA build; the
// remaining disagreement was one argument in 4e8 on the build host (a value 5 double-ulps from a float tie).  tests/test_gpu_parity.py
// checks the function itself against the host C library bit for bit on 200 000 arguments, and for the filter asks for >= 99.5 %
// of the pixels bit-identical to the oracle (glibc expf on the test host) and 1e-6 relative elsewhere.
#include "kernels.cuh"
#include <cmath>
#include <cstring>

namespace dsb {

namespace {

constexpr int kFx = 32, kFy = 8;     // CTA tile; the 5x5 window is staged in shared memory with a 2-pixel apron

__constant__ unsigned long long c_exp2fTab[32];     // asuint64(2^(i/32)) - (i << 47), filled by ensure_exp_table()

__device__ __forceinline__ float expf_host(float x)
{
    const unsigned abstop = (__float_as_uint(x) >> 20) & 0x7ffu;
    if (abstop >= 0x42bu) {                              // |x| >= 88 (or NaN / Inf)
        if (__float_as_uint(x) == 0xff800000u) return 0.0f;
        if (abstop >= 0x7f8u) return x + x;
        if (x > 0x1.62e42ep6f) return __int_as_float(0x7f800000);                 // overflow
        if (x < -0x1.9fe368p6f) return 0.0f;                                      // underflow
        if (x < -0x1.9d1d9ep6f) return __int_as_float(1);                         // "may underflow": 0x1.4p-75f squared = 2^-149
    }
    const double xd = (double)x;
    double z = 0x1.71547652b82fep+5 * xd;                // InvLn2N = 0x1.71547652b82fep+0 * 32
    double kd = z + 0x1.8p+52;                           // SHIFT: round to nearest integer in the low mantissa bits
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd -= 0x1.8p+52;
    const double r = z - kd;
    const unsigned long long t = c_exp2fTab[ki & 31u] + (ki << 47);
    const double sc = __longlong_as_double((long long)t);
    // C = {0x1.c6af84b912394p-5 / 32^3, 0x1.ebfce50fac4f3p-3 / 32^2, 0x1.62e42ff0c52d6p-1 / 32}
    z = fma(0x1.c6af84b912394p-5 / 32 / 32 / 32, r, 0x1.ebfce50fac4f3p-3 / 32 / 32);
    const double r2 = r * r;
    double y = fma(0x1.62e42ff0c52d6p-1 / 32, r, 1.0);
    y = fma(z, r2, y);
    y = y * sc;
    return (float)y;
}

__global__ void __launch_bounds__(kFx * kFy) k_filter_depth(const float *__restrict__ in, float *__restrict__ out, int W, int H)
{
    __shared__ float tile[kFy + 4][kFx + 4];
    const int x0 = blockIdx.x * kFx, y0 = blockIdx.y * kFy;
    for (int i = threadIdx.y * kFx + threadIdx.x; i < (kFy + 4) * (kFx + 4); i += kFx * kFy) {
        const int ty = i / (kFx + 4), tx = i - ty * (kFx + 4);
        const int gx = x0 + tx - 2, gy = y0 + ty - 2;
        tile[ty][tx] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? __ldg(in + gx + (size_t)gy * W) : 0.0f;
    }
    __syncthreads();
    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x >= W || y >= H) return;
    if (x < 2 || x >= W - 2 || y < 2 || y >= H - 2) { out[x + (size_t)y * W] = 0.0f; return; }   // image_out->Clear()
    const float z = tile[threadIdx.y + 2][threadIdx.x + 2];
    if (z < 0.0f) { out[x + (size_t)y * W] = -1.0f; return; }
    const float kSigmaL = 1.2232f;   // MEAN_SIGMA_L
    const float sigma_z = 1.0f / (0.0012f + 0.0019f * (z - 0.4f) * (z - 0.4f) + 0.0001f / sqrtf(z) * 0.25f);
    float final_depth = 0.0f, w_sum = 0.0f;
#pragma unroll
    for (int i = -2; i <= 2; i++) {
#pragma unroll
        for (int j = -2; j <= 2; j++) {
            const float tmpz = tile[threadIdx.y + 2 + i][threadIdx.x + 2 + j];
            if (tmpz < 0.0f) continue;
            float dz = tmpz - z; dz *= dz;
            const float w = expf_host(-0.5f * ((float)(abs(i) + abs(j)) * kSigmaL * kSigmaL + dz * sigma_z * sigma_z));
            w_sum += w;
            final_depth += w * tmpz;
        }
    }
    final_depth /= w_sum;
    out[x + (size_t)y * W] = final_depth;
}

} // namespace

// 2^(i/32) correctly rounded to double (via the 64-bit-mantissa long double exp2), stored the way glibc's table is
static void ensure_exp_table()
{
    static bool done = false;
    if (done) return;
    unsigned long long tab[32];
    for (int i = 0; i < 32; i++) {
        const double v = (double)exp2l((long double)i / 32.0L);
        unsigned long long bits; memcpy(&bits, &v, 8);
        tab[i] = bits - ((unsigned long long)i << 47);
    }
    cudaMemcpyToSymbol(c_exp2fTab, tab, sizeof(tab));
    done = true;
}

namespace {
__global__ void k_expf_selftest(const float *__restrict__ x, float *__restrict__ y, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) y[i] = expf_host(x[i]);
}
}

void launch_expf_selftest(cudaStream_t st, const float *x, float *y, int n)
{
    ensure_exp_table();
    DSB_LAUNCH(k_expf_selftest, kNumSMs * 4, 256, 0, st, x, y, n);
}

namespace {
// convertDisparityToDepth, DeviceAgnostic/ITMViewBuilder.h:7-20 (Kinect-style raw disparities)
__global__ void __launch_bounds__(256) k_disparity_to_depth(const short *__restrict__ in, float *__restrict__ out, float cx, float cy,
                                                            float fx, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float t = cx - (float)in[i];
        const float depth = (t == 0) ? 0.0f : 8.0f * cy * fx / t;
        out[i] = (depth > 0) ? depth : -1.0f;
    }
}
}

void launch_disparity_to_depth(cudaStream_t st, const short *in, float *out, float paramX, float paramY, float fx, int n)
{
    int grid = div_up(n, 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    DSB_LAUNCH(k_disparity_to_depth, grid, 256, 0, st, in, out, paramX, paramY, fx, n);
}

void launch_filter_depth(cudaStream_t st, const float *in, float *out, int W, int H)
{
    ensure_exp_table();
    dim3 block(kFx, kFy), grid(div_up(W, kFx), div_up(H, kFy));
    DSB_LAUNCH(k_filter_depth, grid, block, 0, st, in, out, W, H);
}

namespace {

// acos(float) in the reference resolves to glibc's acosf, which up to glibc 2.39 is the fdlibm routine (sysdeps/ieee754/flt-32/
// e_acosf.c): a rational approximation R(z) = P(z) / Q(z) evaluated in float, one correctly rounded square root and a split
// pi/2.  Every step is a single IEEE float operation, so with --fmad=false the restatement below is bit-identical to the host
// library: checked on the build host over every float in [-1, 1] (2 130 706 434 arguments, 0 mismatches), and on the device by
// tests/test_gpu_parity.py through dsb200_selftest_acosf.  (glibc >= 2.40 ships a correctly rounded acosf; the test would say so.)
__device__ __forceinline__ float acosf_host(float x)
{
    const float pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f;
    const float pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f, pS3 = -4.0055535734e-02f,
                pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f;
    const float qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f, qS3 = -6.8828397989e-01f, qS4 = 7.7038154006e-02f;
    const int hx = __float_as_int(x), ix = hx & 0x7fffffff;
    if (ix == 0x3f800000) return (hx > 0) ? 0.0f : pi + 2.0f * pio2_lo;           // |x| == 1
    if (ix > 0x3f800000) return (x - x) / (x - x);                                // |x| > 1 or NaN
    if (ix < 0x3f000000) {                                                        // |x| < 0.5
        if (ix <= 0x32800000) return pio2_hi + pio2_lo;                           // |x| <= 2^-26
        const float z = x * x;
        const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
        const float q = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
        const float r = p / q;
        return pio2_hi - (x - (pio2_lo - x * r));
    }
    if (hx < 0) {                                                                 // x < -0.5
        const float z = (1.0f + x) * 0.5f;
        const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
        const float q = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
        const float s = sqrtf(z), r = p / q, w = r * s - pio2_lo;
        return pi - 2.0f * (s + w);
    }
    const float z = (1.0f - x) * 0.5f, s = sqrtf(z);                              // x > 0.5
    const float df = __int_as_float(__float_as_int(s) & 0xfffff000);
    const float c = (z - df * df) / (s + df);
    const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
    const float q = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
    const float r = p / q, w = r * s + c;
    return 2.0f * (df + w);
}

__global__ void k_acosf_selftest(const float *__restrict__ x, float *__restrict__ y, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) y[i] = acosf_host(x[i]);
}

// computeNormalAndWeight, DeviceAgnostic/ITMViewBuilder.h:55-112, over the interior [2, W-2) x [2, H-2) like
// ITMViewBuilder_CPU::ComputeNormalAndWeights (CPU/ITMViewBuilder_CPU.cpp:132-143): pixels outside it are not written, and a
// rejected pixel only gets normal.w = -1 and sigmaZ = -1 (its normal.xyz keeps what the image held).  The un-projection multiplies
// by the focal length where a pinhole model divides; that is the reference's arithmetic and is kept.
__global__ void __launch_bounds__(kFx * kFy) k_normal_and_weight(const float *__restrict__ depth, float4 *__restrict__ normal_out,
                                                                 float *__restrict__ sigmaZ_out, int W, int H, float4 intr)
{
    const int x = blockIdx.x * kFx + threadIdx.x, y = blockIdx.y * kFy + threadIdx.y;
    if (x < 2 || x >= W - 2 || y < 2 || y >= H - 2) return;
    const size_t idx = x + (size_t)y * W;
    const float z = __ldg(depth + idx);
    bool ok = !(z < 0.0f);
    float zxp = 0.0f, zxm = 0.0f, zyp = 0.0f, zym = 0.0f;
    if (ok) {
        zxp = __ldg(depth + idx + 1); zyp = __ldg(depth + idx + W);
        zxm = __ldg(depth + idx - 1); zym = __ldg(depth + idx - W);
        ok = !(zxp <= 0 || zyp <= 0 || zxm <= 0 || zym <= 0);
    }
    float nx = 0.0f, ny = 0.0f, nz = 0.0f;
    if (ok) {
        const float fx_ = (float)x, fy_ = (float)y;
        // unprojected neighbours (x multiplies intr.x, y multiplies intr.y; principal point intr.z, intr.w)
        const float xp_x = zxp * ((fx_ + 1.0f) - intr.z) * intr.x, xp_y = zxp * (fy_ - intr.w) * intr.y;
        const float xm_x = zxm * ((fx_ - 1.0f) - intr.z) * intr.x, xm_y = zxm * (fy_ - intr.w) * intr.y;
        const float yp_x = zyp * (fx_ - intr.z) * intr.x, yp_y = zyp * ((fy_ + 1.0f) - intr.w) * intr.y;
        const float ym_x = zym * (fx_ - intr.z) * intr.x, ym_y = zym * ((fy_ - 1.0f) - intr.w) * intr.y;
        const float dxx = xp_x - xm_x, dxy = xp_y - xm_y, dxz = zxp - zxm;
        const float dyx = yp_x - ym_x, dyy = yp_y - ym_y, dyz = zyp - zym;
        nx = dxy * dyz - dxz * dyy;
        ny = dxz * dyx - dxx * dyz;
        nz = dxx * dyy - dxy * dyx;
        ok = !(nx == 0.0f && ny == 0.0f && nz == 0.0f);
    }
    if (!ok) {
        reinterpret_cast<float *>(normal_out + idx)[3] = -1.0f;
        sigmaZ_out[idx] = -1.0f;
        return;
    }
    const float norm = 1.0f / sqrtf(nx * nx + ny * ny + nz * nz);
    nx *= norm; ny *= norm; nz *= norm;
    normal_out[idx] = make_float4(nx, ny, nz, 1.0f);
    const float theta = acosf_host(nz);
    const float kPi = 3.1415926535897932384626433832795f;          // ORUtils/MathUtils.h:28, a float
    const float theta_diff = theta / (kPi * 0.5f - theta);
    sigmaZ_out[idx] = 0.0012f + 0.0019f * (z - 0.4f) * (z - 0.4f) + 0.0001f / sqrtf(z) * theta_diff * theta_diff;
}

} // namespace

void launch_acosf_selftest(cudaStream_t st, const float *x, float *y, int n)
{
    DSB_LAUNCH(k_acosf_selftest, kNumSMs * 4, 256, 0, st, x, y, n);
}

void launch_normal_and_weight(cudaStream_t st, const float *depth, float *normal4, float *sigmaZ, int W, int H, const float *intr)
{
    dim3 block(kFx, kFy), grid(div_up(W, kFx), div_up(H, kFy));
    DSB_LAUNCH(k_normal_and_weight, grid, block, 0, st, depth, reinterpret_cast<float4 *>(normal4), sigmaZ, W, H,
               make_float4(intr[0], intr[1], intr[2], intr[3]));
}

} // namespace dsb
