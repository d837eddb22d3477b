// visibility.cuh -- frustum test of one voxel block (shared by allocation and FindVisibleBlocks).
#pragma once
#include "common.cuh"

namespace dsb {

// checkPointVisibility / checkBlockVisibility (DA/ITMSceneReconstructionEngine.h:491-569)
__device__ __forceinline__ void point_visibility(bool &vis, bool &visEnl, Vec4 p, const Mat4 &M, const float *proj,
                                                 int W, int H, bool useSwapping)
{
    Vec4 q = mul(M, p);
    if (q.z < 1e-10f) return;
    float u = proj[0] * q.x / q.z + proj[2];
    float v = proj[1] * q.y / q.z + proj[3];
    if (u >= 0 && u < W && v >= 0 && v < H) { vis = true; visEnl = true; }
    else if (useSwapping) {
        int lx = -W / 8, ly = W + W / 8, lz = -H / 8, lw = H + H / 8;
        if (u >= lx && u < ly && v >= lz && v < lw) visEnl = true;
    }
}

__device__ __forceinline__ void block_visibility(bool &vis, bool &visEnl, short bx, short by, short bz, const Mat4 &M,
                                 const float *proj, float voxelSize, int W, int H, bool useSwapping)
{
    const float f = (float)kBlock * voxelSize;
    Vec4 p;
    vis = false; visEnl = false;
    p.x = (float)bx * f; p.y = (float)by * f; p.z = (float)bz * f; p.w = 1.0f;
    point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.z += f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.y += f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.x += f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.z -= f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.y -= f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.x -= f; p.y += f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;
    p.x += f; p.y -= f; p.z += f; point_visibility(vis, visEnl, p, M, proj, W, H, useSwapping);
}

} // namespace dsb
