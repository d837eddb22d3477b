// oracle/refcuda_shim.cu -- TEST / BENCH INFRASTRUCTURE ONLY (never linked into the product).
//
// A thin C ABI around the UNMODIFIED reference CUDA engines, compiled in place from /root/reference by
// oracle/Makefile (target `refcuda`) for sm_100 into
//   oracle/_ref/libitmref_cuda.so        IEEE build (--fmad=false, no fast math): the set-level PIN for the entry points
//                                        the reference implements only in CUDA (decay, slide window, de-integration,
//                                        history queues)
//   oracle/_ref/libitmref_cuda_fast.so   the reference's own flags (--use_fast_math): the GPU COMPARATOR of SURVEY 8(d)
// No reference source is copied: this file #includes the reference headers and forwards calls.
//
// Reference classes driven here:
//   ITMSceneReconstructionEngine_CUDA<ITMVoxel,ITMVoxelBlockHash>
//       ITMLib/Engine/DeviceSpecific/CUDA/ITMSceneReconstructionEngine_CUDA.cu:111-1080
//   ITMVisualisationEngine_CUDA<ITMVoxel,ITMVoxelBlockHash>
//       ITMLib/Engine/DeviceSpecific/CUDA/ITMVisualisationEngine_CUDA.cu:137-581
// The engines use the legacy default stream and blocking copies, so every call is synchronous on return.
#include <queue>
#include <map>
#include "ITMLib/Engine/ITMSceneReconstructionEngine.h"
// The history queues (CUDA/ITMSceneReconstructionEngine_CUDA.h:41-48) are private and have no accessor; the tests compare
// their sizes, so this translation unit (only) reads the class declaration with `private` spelt `public`.  The engine
// itself is compiled from the unmodified reference source; access specifiers do not change the object layout.
#define private public
#include "ITMLib/Engine/DeviceSpecific/CUDA/ITMSceneReconstructionEngine_CUDA.h"
#undef private
#include "ITMLib/Engine/DeviceSpecific/CUDA/ITMVisualisationEngine_CUDA.h"
#include "ITMLib/Objects/ITMRenderState_VH.h"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <fcntl.h>
#include <unistd.h>

using namespace ITMLib::Engine;
using namespace ITMLib::Objects;

namespace {
struct CuCtx {
    ITMSceneParams params;
    ITMScene<ITMVoxel, ITMVoxelBlockHash> *scene;
    ITMSceneReconstructionEngine_CUDA<ITMVoxel, ITMVoxelBlockHash> *recon;
    ITMVisualisationEngine_CUDA<ITMVoxel, ITMVoxelBlockHash> *vis;
    ITMRGBDCalib calib;
    ITMView *view;
    ITMTrackingState *tracking;
    ITMRenderState_VH *rs, *rsFree;
    ITMUChar4Image *outImg;
    ITMFloatImage *outFloat;
    Vector2i imgSize;
    void *stage; size_t stageBytes;      // device staging for block gathers
    int *idsDev; size_t idsCap;
    cudaEvent_t ev[6];
    std::string error;
};
Matrix4f toM(const float *m) { Matrix4f M; for (int i = 0; i < 16; i++) M.m[i] = m[i]; return M; }

// The reference prints a status block per call (CUDA .cu:341-349, 939-951); keep it out of the caller's stdout.
struct Quiet {
    int saved;
    Quiet() { fflush(stdout); saved = dup(1); int nul = open("/dev/null", O_WRONLY); dup2(nul, 1); close(nul); }
    ~Quiet() { fflush(stdout); dup2(saved, 1); close(saved); }
};

__global__ void gather_blocks(const uint4 *vba, const int *ptrs, int n, uint4 *out)
{
    // one 4 KiB voxel block = 256 uint4
    for (int b = blockIdx.x; b < n; b += gridDim.x) {
        const int p = ptrs[b];
        if (p < 0) continue;
        out[(size_t)b * 256 + threadIdx.x] = vba[(size_t)p * 256 + threadIdx.x];
    }
}
} // namespace

#define CU_OK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { c->error = cudaGetErrorString(e_); return -2; } } while (0)
#define GUARD(stmt) do { try { Quiet q_; stmt; } catch (const std::exception &ex) { c->error = ex.what(); return -1; } } while (0)

extern "C" {

int itmcu_device_count() { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }
int itmcu_no_total_entries() { return ITMVoxelBlockHash::noTotalEntries; }
int itmcu_local_block_num() { return SDF_LOCAL_BLOCK_NUM; }
int itmcu_bucket_num() { return SDF_BUCKET_NUM; }
int itmcu_excess_list_size() { return SDF_EXCESS_LIST_SIZE; }
#ifdef __FAST_MATH__
int itmcu_fast_math() { return 1; }
#else
int itmcu_fast_math() { return 0; }
#endif

void *itmcu_create(int W, int H, float voxelSize, float mu, int maxW, float vfMin, float vfMax, int stopAtMaxW, int useSwapping)
{
    Quiet q;
    CuCtx *c = new CuCtx{ITMSceneParams(mu, maxW, voxelSize, vfMin, vfMax, stopAtMaxW != 0)};
    c->imgSize = Vector2i(W, H);
    c->scene = new ITMScene<ITMVoxel, ITMVoxelBlockHash>(&c->params, useSwapping != 0, MEMORYDEVICE_CUDA);
    c->recon = new ITMSceneReconstructionEngine_CUDA<ITMVoxel, ITMVoxelBlockHash>();
    c->vis = new ITMVisualisationEngine_CUDA<ITMVoxel, ITMVoxelBlockHash>();
    c->view = new ITMView(&c->calib, c->imgSize, c->imgSize, true);
    c->tracking = new ITMTrackingState(c->imgSize, MEMORYDEVICE_CUDA);
    c->rs = c->vis->CreateRenderState(c->scene, c->imgSize);
    c->rsFree = c->vis->CreateRenderState(c->scene, c->imgSize);
    c->outImg = new ITMUChar4Image(c->imgSize, true, true);
    c->outFloat = new ITMFloatImage(c->imgSize, true, true);
    c->stage = nullptr; c->stageBytes = 0; c->idsDev = nullptr; c->idsCap = 0;
    for (int i = 0; i < 6; i++) cudaEventCreate(&c->ev[i]);
    c->recon->ResetScene(c->scene);
    cudaDeviceSynchronize();
    return c;
}

void itmcu_destroy(void *h)
{
    CuCtx *c = (CuCtx *)h;
    cudaDeviceSynchronize();
    for (int i = 0; i < 6; i++) cudaEventDestroy(c->ev[i]);
    cudaFree(c->stage); cudaFree(c->idsDev);
    delete c->outFloat; delete c->outImg; delete c->rsFree; delete c->rs; delete c->tracking;
    delete c->view; delete c->vis; delete c->recon; delete c->scene; delete c;
}

const char *itmcu_last_error(void *h) { return ((CuCtx *)h)->error.c_str(); }

int itmcu_reset(void *h)
{
    CuCtx *c = (CuCtx *)h;
    GUARD(c->recon->ResetScene(c->scene));
    c->rs->noVisibleEntries = 0;
    CU_OK(cudaMemset(c->rs->GetEntriesVisibleType(), 0, ITMVoxelBlockHash::noTotalEntries));
    CU_OK(cudaDeviceSynchronize());
    return 0;
}

void itmcu_set_calib(void *h, const float *intrD, const float *intrRGB, const float *rgb2depth)
{
    CuCtx *c = (CuCtx *)h;
    c->view->calib->intrinsics_d.SetFrom(intrD[0], intrD[1], intrD[2], intrD[3], c->imgSize.x, c->imgSize.y);
    c->view->calib->intrinsics_rgb.SetFrom(intrRGB[0], intrRGB[1], intrRGB[2], intrRGB[3], c->imgSize.x, c->imgSize.y);
    c->view->calib->trafo_rgb_to_depth.SetFrom(toM(rgb2depth));
}

// depth / rgba: HOST (fromDevice = 0) or DEVICE (fromDevice = 1) arrays
int itmcu_set_view(void *h, const float *depth, const uint8_t *rgba, double timestamp, int fromDevice)
{
    CuCtx *c = (CuCtx *)h;
    const size_t n = (size_t)c->imgSize.x * c->imgSize.y;
    const cudaMemcpyKind k = fromDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    CU_OK(cudaMemcpy(c->view->depth->GetData(MEMORYDEVICE_CUDA), depth, n * sizeof(float), k));
    CU_OK(cudaMemcpy(c->view->rgb->GetData(MEMORYDEVICE_CUDA), rgba, n * 4, k));
    c->view->mTimeStamp = timestamp;
    return 0;
}

void itmcu_set_pose(void *h, const float *M) { ((CuCtx *)h)->tracking->pose_d->SetM(toM(M)); }
void itmcu_set_requires_full_rendering(void *h, int on) { ((CuCtx *)h)->tracking->requiresFullRendering = on != 0; }
void itmcu_set_weights(void *h, int depthWeighting, int maxNewW, int maxDistance)
{
    WeightParams wp; wp.depthWeighting = depthWeighting != 0; wp.maxNewW = maxNewW; wp.maxDistance = maxDistance;
    ((CuCtx *)h)->recon->SetFusionWeightParams(wp);
}

int itmcu_allocate(void *h, int onlyUpdateVisibleList, int isDefusion)
{
    CuCtx *c = (CuCtx *)h;
    GUARD(c->recon->AllocateSceneFromDepth(c->scene, c->view, c->tracking, c->rs, onlyUpdateVisibleList != 0, isDefusion != 0));
    return 0;
}
int itmcu_integrate(void *h) { CuCtx *c = (CuCtx *)h; GUARD(c->recon->IntegrateIntoScene(c->scene, c->view, c->tracking, c->rs)); CU_OK(cudaDeviceSynchronize()); return 0; }
int itmcu_deintegrate(void *h) { CuCtx *c = (CuCtx *)h; GUARD(c->recon->DeIntegrateIntoScene(c->scene, c->view, c->tracking)); CU_OK(cudaDeviceSynchronize()); return 0; }
int itmcu_decay(void *h, int maxWeight, int minAge) { CuCtx *c = (CuCtx *)h; GUARD(c->recon->Decay(c->scene, c->rs, maxWeight, minAge, false)); return 0; }
int itmcu_decay_defusion(void *h, int maxWeight, int minAge) { CuCtx *c = (CuCtx *)h; GUARD(c->recon->DecayDefusionPart(c->scene, c->rs, maxWeight, minAge, false)); return 0; }
int itmcu_slide_window(void *h, int maxAge) { CuCtx *c = (CuCtx *)h; GUARD(c->recon->SlideWindow(c->scene, c->rs, maxAge)); return 0; }
int itmcu_slide_window_defusion(void *h, int maxAge, int maxSize) { CuCtx *c = (CuCtx *)h; GUARD(c->recon->SlideWindowDefusionPart(c->scene, c->rs, maxAge, maxSize)); return 0; }
// sizes of frameVisibleBlocks, frameVisibleBlocksForSlideWindow, frameVisibleBlocksDefusion,
// frameVisibleBlocksForSlideWindowDefusion, mDefusionBlockDataBase
void itmcu_history_sizes(void *h, int *out)
{
    CuCtx *c = (CuCtx *)h;
    out[0] = (int)c->recon->frameVisibleBlocks.size(); out[1] = (int)c->recon->frameVisibleBlocksForSlideWindow.size();
    out[2] = (int)c->recon->frameVisibleBlocksDefusion.size(); out[3] = (int)c->recon->frameVisibleBlocksForSlideWindowDefusion.size();
    out[4] = (int)c->recon->mDefusionBlockDataBase.size();
}
long long itmcu_decayed_block_count(void *h) { return (long long)((CuCtx *)h)->recon->GetDecayedBlockCount(); }

int itmcu_create_expected_depths(void *h)
{
    CuCtx *c = (CuCtx *)h;
    GUARD(c->vis->CreateExpectedDepths(c->scene, c->tracking->pose_d, &c->view->calib->intrinsics_d, c->rs));
    return 0;
}
int itmcu_create_icp_maps(void *h) { CuCtx *c = (CuCtx *)h; GUARD(c->vis->CreateICPMaps(c->scene, c->view, c->tracking, c->rs)); CU_OK(cudaDeviceSynchronize()); return 0; }
int itmcu_find_surface(void *h)
{
    CuCtx *c = (CuCtx *)h;
    GUARD(c->vis->FindSurface(c->scene, c->tracking->pose_d, &c->view->calib->intrinsics_d, c->rs));
    CU_OK(cudaDeviceSynchronize());
    return 0;
}
int itmcu_render_free_view(void *h, const float *M, const float *intr, int renderType)
{
    CuCtx *c = (CuCtx *)h;
    ITMPose pose; pose.SetM(toM(M));
    ITMIntrinsics in; in.SetFrom(intr[0], intr[1], intr[2], intr[3], c->imgSize.x, c->imgSize.y);
    GUARD(c->vis->FindVisibleBlocks(c->scene, &pose, &in, c->rsFree);
          c->vis->CreateExpectedDepths(c->scene, &pose, &in, c->rsFree);
          c->vis->RenderImage(c->scene, &pose, &in, c->rsFree, c->outImg, c->outFloat, (IITMVisualisationEngine::RenderImageType)renderType));
    CU_OK(cudaDeviceSynchronize());
    return 0;
}

// One DenseSLAM keyframe through the reference's engines in ITMDenseMapper::ProcessFrame + ITMTrackingController::Prepare
// order (ITMDenseMapper.cpp:56-81, ITMTrackingController.cpp:21-58) on the view / pose already set; ms[0..3] = device time
// of allocate, integrate, expected depths, raycast + ICP maps (CUDA events on the legacy stream the engines use), ms[4] = host
// wall time of the four calls (the engines block on 12-byte counter copies in between).
int itmcu_process_frame_timed(void *h, float *ms)
{
    CuCtx *c = (CuCtx *)h;
    try {
        Quiet q;
        const auto t0 = std::chrono::steady_clock::now();
        cudaEventRecord(c->ev[0], 0);
        c->recon->AllocateSceneFromDepth(c->scene, c->view, c->tracking, c->rs, false, false);
        cudaEventRecord(c->ev[1], 0);
        c->recon->IntegrateIntoScene(c->scene, c->view, c->tracking, c->rs);
        cudaEventRecord(c->ev[2], 0);
        c->vis->CreateExpectedDepths(c->scene, c->tracking->pose_d, &c->view->calib->intrinsics_d, c->rs);
        cudaEventRecord(c->ev[3], 0);
        c->vis->CreateICPMaps(c->scene, c->view, c->tracking, c->rs);
        cudaEventRecord(c->ev[4], 0);
        cudaEventSynchronize(c->ev[4]);
        const auto t1 = std::chrono::steady_clock::now();
        for (int i = 0; i < 4; i++) cudaEventElapsedTime(&ms[i], c->ev[i], c->ev[i + 1]);
        ms[4] = std::chrono::duration<float, std::milli>(t1 - t0).count();
    } catch (const std::exception &ex) { c->error = ex.what(); return -1; }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { c->error = cudaGetErrorString(e); return -2; }
    return 0;
}

// ---- state downloads (host buffers supplied by the caller) ----
int itmcu_last_free_block(void *h) { return ((CuCtx *)h)->scene->localVBA.lastFreeBlockId; }
int itmcu_last_free_excess(void *h) { return ((CuCtx *)h)->scene->index.GetLastFreeExcessListId(); }
int itmcu_no_visible(void *h, int freeView) { CuCtx *c = (CuCtx *)h; return (freeView ? c->rsFree : c->rs)->noVisibleEntries; }

int itmcu_get_hash(void *h, void *out)
{
    CuCtx *c = (CuCtx *)h;
    CU_OK(cudaMemcpy(out, c->scene->index.GetEntries(), (size_t)ITMVoxelBlockHash::noTotalEntries * sizeof(ITMHashEntry), cudaMemcpyDeviceToHost));
    return 0;
}
int itmcu_get_alloc_list(void *h, int *out)
{
    CuCtx *c = (CuCtx *)h;
    CU_OK(cudaMemcpy(out, c->scene->localVBA.GetAllocationList(), (size_t)SDF_LOCAL_BLOCK_NUM * sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}
int itmcu_get_visible_ids(void *h, int freeView, int *out)
{
    CuCtx *c = (CuCtx *)h;
    ITMRenderState_VH *rs = freeView ? c->rsFree : c->rs;
    if (rs->noVisibleEntries > 0)
        CU_OK(cudaMemcpy(out, rs->GetVisibleEntryIDs(), (size_t)rs->noVisibleEntries * sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}
int itmcu_get_visible_type(void *h, uint8_t *out)
{
    CuCtx *c = (CuCtx *)h;
    CU_OK(cudaMemcpy(out, c->rs->GetEntriesVisibleType(), ITMVoxelBlockHash::noTotalEntries, cudaMemcpyDeviceToHost));
    return 0;
}
// voxel blocks ptrs[0..n) of the VBA, 4096 bytes each, in list order (a negative ptr leaves its slot untouched)
int itmcu_get_blocks(void *h, const int *ptrs, int n, void *out)
{
    CuCtx *c = (CuCtx *)h;
    if (n <= 0) return 0;
    const size_t bytes = (size_t)n * 4096;
    if (bytes > c->stageBytes) { cudaFree(c->stage); c->stage = nullptr; c->stageBytes = 0; CU_OK(cudaMalloc(&c->stage, bytes)); c->stageBytes = bytes; }
    if ((size_t)n > c->idsCap) { cudaFree(c->idsDev); c->idsDev = nullptr; c->idsCap = 0; CU_OK(cudaMalloc(&c->idsDev, (size_t)n * 4)); c->idsCap = n; }
    CU_OK(cudaMemcpy(c->idsDev, ptrs, (size_t)n * 4, cudaMemcpyHostToDevice));
    CU_OK(cudaMemset(c->stage, 0, bytes));
    gather_blocks<<<n < 4096 ? n : 4096, 256>>>((const uint4 *)c->scene->localVBA.GetVoxelBlocks(), c->idsDev, n, (uint4 *)c->stage);
    CU_OK(cudaGetLastError());
    CU_OK(cudaMemcpy(out, c->stage, bytes, cudaMemcpyDeviceToHost));
    return 0;
}
static int get_image(CuCtx *c, const void *dev, size_t bytes, void *out) { CU_OK(cudaMemcpy(out, dev, bytes, cudaMemcpyDeviceToHost)); return 0; }
int itmcu_get_range_image(void *h, int freeView, float *out)
{
    CuCtx *c = (CuCtx *)h; ITMRenderState_VH *rs = freeView ? c->rsFree : c->rs;
    return get_image(c, rs->renderingRangeImage->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 8, out);
}
int itmcu_get_raycast_result(void *h, int freeView, float *out)
{
    CuCtx *c = (CuCtx *)h; ITMRenderState_VH *rs = freeView ? c->rsFree : c->rs;
    return get_image(c, rs->raycastResult->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 16, out);
}
int itmcu_get_icp_points(void *h, float *out)
{
    CuCtx *c = (CuCtx *)h;
    return get_image(c, c->tracking->pointCloud->locations->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 16, out);
}
int itmcu_get_icp_normals(void *h, float *out)
{
    CuCtx *c = (CuCtx *)h;
    return get_image(c, c->tracking->pointCloud->colours->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 16, out);
}
int itmcu_get_icp_rendering(void *h, uint8_t *out)
{
    CuCtx *c = (CuCtx *)h;
    return get_image(c, c->rs->raycastImage->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 4, out);
}
int itmcu_get_out_image(void *h, uint8_t *out)
{
    CuCtx *c = (CuCtx *)h;
    return get_image(c, c->outImg->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 4, out);
}
int itmcu_get_out_depth(void *h, float *out)
{
    CuCtx *c = (CuCtx *)h;
    return get_image(c, c->outFloat->GetData(MEMORYDEVICE_CUDA), (size_t)c->imgSize.x * c->imgSize.y * 4, out);
}

} // extern "C"
