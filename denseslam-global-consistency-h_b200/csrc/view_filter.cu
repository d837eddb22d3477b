// view_filter.cu -- the view builder's optional bilateral depth filter.
//
//   k_filter_depth   filterDepth, DeviceAgnostic/ITMViewBuilder.h:30-53, applied like ITMViewBuilder_CPU::DepthFiltering
//                    (CPU/ITMViewBuilder_CPU.cpp:119-130): output cleared, interior [2, W-2) x [2, H-2) filtered;
//   five passes ping-ponging between the view's depth image and a scratch image, result copied back (UpdateView :49-58).
//
// Arithmetic follows the reference expression by expression (float, no contraction, IEEE division and square root).  The
// one library call, exp(float), is glibc's expf on the reference's host: a double-precision evaluation rounded to float.
// Here it is CUDA's double exp rounded to float, which agrees with any < 1 ulp-double evaluation except where the exact
// value lies within ~2^-29 of a float rounding boundary (about one weight in 10^8); tests/test_gpu_parity.py therefore
// checks this kernel against the oracle with a 1e-6 relative tolerance and requires >= 99.99 % of the pixels bit-identical.
#include "kernels.cuh"

namespace dsb {

namespace {

constexpr int kFx = 32, kFy = 8;     // CTA tile; the 5x5 window is staged in shared memory with a 2-pixel apron

__device__ __forceinline__ float exp_like_host(float x) { return (float)exp((double)x); }

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
            const float w = exp_like_host(-0.5f * ((float)(abs(i) + abs(j)) * kSigmaL * kSigmaL + dz * sigma_z * sigma_z));
            w_sum += w;
            final_depth += w * tmpz;
        }
    }
    final_depth /= w_sum;
    out[x + (size_t)y * W] = final_depth;
}

} // namespace

void launch_filter_depth(cudaStream_t st, const float *in, float *out, int W, int H)
{
    dim3 block(kFx, kFy), grid(div_up(W, kFx), div_up(H, kFy));
    DSB_LAUNCH(k_filter_depth, grid, block, 0, st, in, out, W, H);
}

} // namespace dsb
