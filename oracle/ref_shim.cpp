// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A thin C ABI around the UNMODIFIED reference CPU engines, compiled in place from
// /root/reference by oracle/Makefile into oracle/_ref/libitmref.so.  It exists so the
// tests can (1) pin oracle/restate.cpp against the real reference and (2) generate
// the golden fixtures under tests/golden/.  No reference source is copied: this file
// only #includes the reference headers and forwards calls.
//
// Reference classes driven here:
//   ITMSceneReconstructionEngine_CPU<ITMVoxel,ITMVoxelBlockHash>
//       ITMLib/Engine/DeviceSpecific/CPU/ITMSceneReconstructionEngine_CPU.cpp:25-342
//   ITMVisualisationEngine_CPU<ITMVoxel,ITMVoxelBlockHash>
//       ITMLib/Engine/DeviceSpecific/CPU/ITMVisualisationEngine_CPU.cpp:27-180,296-459
//   ITMSwappingEngine_CPU<ITMVoxel,ITMVoxelBlockHash> on the scene's ITMGlobalCache
//       ITMLib/Engine/DeviceSpecific/CPU/ITMSwappingEngine_CPU.cpp:22-297, Objects/ITMGlobalCache.h
//   ITMMeshingEngine_CPU<ITMVoxel,ITMVoxelBlockHash>
//       ITMLib/Engine/DeviceSpecific/CPU/ITMMeshingEngine_CPU.cpp:20-83
//   ITMDepthTracker_CPU (+ the shared host loop ITMLib/Engine/ITMDepthTracker.cpp) and ITMLowLevelEngine_CPU
//       ITMLib/Engine/DeviceSpecific/CPU/ITMDepthTracker_CPU.cpp:13-79, CPU/ITMLowLevelEngine_CPU.cpp:50-62
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMSceneReconstructionEngine_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMVisualisationEngine_CPU.h"
#include "ITMLib/Objects/ITMRenderState_VH.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMViewBuilder_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMDepthTracker_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMLowLevelEngine_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMSwappingEngine_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMMeshingEngine_CPU.h"
#include "ITMLib/Engine/DeviceAgnostic/ITMMeshingEngine.h"   // edgeTable / triangleTable (static const: this TU gets its own copy)

#include <cstring>
#include <cstdint>

using namespace ITMLib::Engine;
using namespace ITMLib::Objects;

namespace {
struct RefCtx {
    ITMSceneParams params;
    ITMScene<ITMVoxel, ITMVoxelBlockHash> *scene;
    ITMSceneReconstructionEngine_CPU<ITMVoxel, ITMVoxelBlockHash> *recon;
    ITMVisualisationEngine_CPU<ITMVoxel, ITMVoxelBlockHash> *vis;
    ITMRGBDCalib calib;
    ITMView *view;
    ITMTrackingState *tracking;
    ITMRenderState_VH *rs;       // tracking-path render state (visible list from Allocate)
    ITMRenderState_VH *rsFree;   // free-view render state (FindVisibleBlocks)
    ITMUChar4Image *outImg;
    ITMFloatImage *outFloat;
    Vector2i imgSize;
};
Matrix4f toM(const float *m) { Matrix4f M; for (int i = 0; i < 16; i++) M.m[i] = m[i]; return M; }
}

extern "C" {

int itmref_sizeof_voxel() { return (int)sizeof(ITMVoxel); }
int itmref_sizeof_hash_entry() { return (int)sizeof(ITMHashEntry); }
int itmref_no_total_entries() { return ITMVoxelBlockHash::noTotalEntries; }
int itmref_local_block_num() { return SDF_LOCAL_BLOCK_NUM; }
int itmref_bucket_num() { return SDF_BUCKET_NUM; }
int itmref_excess_list_size() { return SDF_EXCESS_LIST_SIZE; }

void *itmref_create(int W, int H, float voxelSize, float mu, int maxW, float vfMin, float vfMax,
                    int stopAtMaxW, int useSwapping)
{
    RefCtx *c = new RefCtx{ITMSceneParams(mu, maxW, voxelSize, vfMin, vfMax, stopAtMaxW != 0)};
    c->imgSize = Vector2i(W, H);
    c->scene = new ITMScene<ITMVoxel, ITMVoxelBlockHash>(&c->params, useSwapping != 0, MEMORYDEVICE_CPU);
    c->recon = new ITMSceneReconstructionEngine_CPU<ITMVoxel, ITMVoxelBlockHash>();
    c->vis = new ITMVisualisationEngine_CPU<ITMVoxel, ITMVoxelBlockHash>();
    c->view = new ITMView(&c->calib, c->imgSize, c->imgSize, false);
    c->tracking = new ITMTrackingState(c->imgSize, MEMORYDEVICE_CPU);
    c->rs = c->vis->CreateRenderState(c->scene, c->imgSize);
    c->rsFree = c->vis->CreateRenderState(c->scene, c->imgSize);
    c->outImg = new ITMUChar4Image(c->imgSize, true, false);
    c->outFloat = new ITMFloatImage(c->imgSize, true, false);
    c->recon->ResetScene(c->scene);
    return c;
}

void itmref_destroy(void *h)
{
    RefCtx *c = (RefCtx *)h;
    delete c->outFloat; delete c->outImg; delete c->rsFree; delete c->rs; delete c->tracking;
    delete c->view; delete c->vis; delete c->recon; delete c->scene; delete c;
}

void itmref_reset(void *h) { RefCtx *c = (RefCtx *)h; c->recon->ResetScene(c->scene); c->rs->noVisibleEntries = 0; }

// intr = (fx, fy, cx, cy); rgb2depth = column-major 4x4 (trafo_rgb_to_depth.calib)
void itmref_set_calib(void *h, const float *intrD, const float *intrRGB, const float *rgb2depth)
{
    RefCtx *c = (RefCtx *)h;
    c->view->calib->intrinsics_d.SetFrom(intrD[0], intrD[1], intrD[2], intrD[3], c->imgSize.x, c->imgSize.y);
    c->view->calib->intrinsics_rgb.SetFrom(intrRGB[0], intrRGB[1], intrRGB[2], intrRGB[3], c->imgSize.x, c->imgSize.y);
    c->view->calib->trafo_rgb_to_depth.SetFrom(toM(rgb2depth));
}

void itmref_set_view(void *h, const float *depth, const uint8_t *rgba, double timestamp)
{
    RefCtx *c = (RefCtx *)h;
    size_t n = (size_t)c->imgSize.x * c->imgSize.y;
    memcpy(c->view->depth->GetData(MEMORYDEVICE_CPU), depth, n * sizeof(float));
    memcpy(c->view->rgb->GetData(MEMORYDEVICE_CPU), rgba, n * 4);
    c->view->mTimeStamp = timestamp;
}

// ITMViewBuilder_CPU::UpdateView with the affine depth calibration DenseSLAM configures (DS/InfiniTamDriver.cpp:78-81,286):
// raw int16 depth + RGBA in, view->depth (float) out
void itmref_update_view_raw_ex(void *h, const short *rawDepth, const uint8_t *rgba, float scale, float offset, double timestamp,
                               int useBilateralFilter);
void itmref_update_view_raw(void *h, const short *rawDepth, const uint8_t *rgba, float scale, float offset, double timestamp)
{
    itmref_update_view_raw_ex(h, rawDepth, rgba, scale, offset, timestamp, 0);
}
void itmref_update_view_raw_ex(void *h, const short *rawDepth, const uint8_t *rgba, float scale, float offset, double timestamp,
                               int useBilateralFilter)
{
    RefCtx *c = (RefCtx *)h;
    c->calib.disparityCalib.SetFrom(scale, offset, ITMDisparityCalib::TRAFO_AFFINE);
    ITMViewBuilder_CPU vb(&c->calib);
    ITMUChar4Image rgb(c->imgSize, true, false);
    ITMShortImage raw(c->imgSize, true, false);
    size_t n = (size_t)c->imgSize.x * c->imgSize.y;
    memcpy(rgb.GetData(MEMORYDEVICE_CPU), rgba, n * 4);
    memcpy(raw.GetData(MEMORYDEVICE_CPU), rawDepth, n * sizeof(short));
    ITMView *built = NULL;        // a fresh view makes the builder allocate its own staging images (UpdateView :12-27)
    vb.UpdateView(&built, &rgb, &raw, timestamp, useBilateralFilter != 0);
    memcpy(c->view->depth->GetData(MEMORYDEVICE_CPU), built->depth->GetData(MEMORYDEVICE_CPU), n * sizeof(float));
    memcpy(c->view->rgb->GetData(MEMORYDEVICE_CPU), built->rgb->GetData(MEMORYDEVICE_CPU), n * 4);
    c->view->mTimeStamp = timestamp;
    delete built;
}
// ITMViewBuilder_CPU::ConvertDisparityToDepth on raw arrays (Kinect-style disparity calibration)
void itmref_convert_disparity_to_depth(const short *disparity, float *depth, int W, int H, float paramX, float paramY, float fx)
{
    ITMRGBDCalib calib;
    ITMViewBuilder_CPU vb(&calib);
    ITMShortImage in(Vector2i(W, H), true, false);
    ITMFloatImage out(Vector2i(W, H), true, false);
    memcpy(in.GetData(MEMORYDEVICE_CPU), disparity, sizeof(short) * (size_t)W * H);
    ITMIntrinsics intr; intr.SetFrom(fx, fx, W * 0.5f, H * 0.5f, W, H);
    vb.ConvertDisparityToDepth(&out, &in, &intr, Vector2f(paramX, paramY));
    memcpy(depth, out.GetData(MEMORYDEVICE_CPU), sizeof(float) * (size_t)W * H);
}
// ITMViewBuilder_CPU::ComputeNormalAndWeights on raw arrays; normal4 / sigmaZ come in pre-filled (the reference writes only
// part of them) and go out updated
void itmref_compute_normal_and_weights(const float *depth, float *normal4, float *sigmaZ, int W, int H, const float *intr)
{
    ITMRGBDCalib calib;
    ITMViewBuilder_CPU vb(&calib);
    const size_t n = (size_t)W * H;
    ITMFloatImage in(Vector2i(W, H), true, false), sz(Vector2i(W, H), true, false);
    ITMFloat4Image nrm(Vector2i(W, H), true, false);
    memcpy(in.GetData(MEMORYDEVICE_CPU), depth, sizeof(float) * n);
    memcpy(sz.GetData(MEMORYDEVICE_CPU), sigmaZ, sizeof(float) * n);
    memcpy(nrm.GetData(MEMORYDEVICE_CPU), normal4, sizeof(float) * 4 * n);
    vb.ComputeNormalAndWeights(&nrm, &sz, &in, Vector4f(intr[0], intr[1], intr[2], intr[3]));
    memcpy(sigmaZ, sz.GetData(MEMORYDEVICE_CPU), sizeof(float) * n);
    memcpy(normal4, nrm.GetData(MEMORYDEVICE_CPU), sizeof(float) * 4 * n);
}
// ITMSwappingEngine_CPU on the scene's global cache (the context must have been created with useSwapping)
void itmref_swap_integrate_global_into_local(void *h)
{
    RefCtx *c = (RefCtx *)h;
    ITMSwappingEngine_CPU<ITMVoxel, ITMVoxelBlockHash> sw;
    sw.IntegrateGlobalIntoLocal(c->scene, c->rs);
}
void itmref_swap_save_to_global_memory(void *h, int withRenderState)
{
    RefCtx *c = (RefCtx *)h;
    ITMSwappingEngine_CPU<ITMVoxel, ITMVoxelBlockHash> sw;
    if (withRenderState) sw.SaveToGlobalMemory(c->scene, c->rs); else sw.SaveToGlobalMemory(c->scene);
}
uint8_t *itmref_swap_states(void *h) { return (uint8_t *)((RefCtx *)h)->scene->globalCache->GetSwapStates(false); }
int itmref_swap_has_stored(void *h, int entry) { return ((RefCtx *)h)->scene->globalCache->HasStoredData(entry) ? 1 : 0; }
const void *itmref_swap_stored_block(void *h, int entry) { return ((RefCtx *)h)->scene->globalCache->GetStoredVoxelBlock(entry); }
void itmref_swap_has_stored_all(void *h, uint8_t *out)
{
    ITMGlobalCache<ITMVoxel> *g = ((RefCtx *)h)->scene->globalCache;
    for (int i = 0; i < g->noTotalEntries; i++) out[i] = g->HasStoredData(i) ? 1 : 0;
}
int itmref_sizeof_swap_state() { return (int)sizeof(ITMHashSwapState); }
int itmref_transfer_block_num() { return SDF_TRANSFER_BLOCK_NUM; }
const float *itmref_view_depth(void *h) { return ((RefCtx *)h)->view->depth->GetData(MEMORYDEVICE_CPU); }

void itmref_set_pose(void *h, const float *M) { ((RefCtx *)h)->tracking->pose_d->SetM(toM(M)); }
void itmref_get_pose(void *h, float *M, float *invM)
{
    RefCtx *c = (RefCtx *)h;
    Matrix4f m = c->tracking->pose_d->GetM(), im = c->tracking->pose_d->GetInvM();
    memcpy(M, m.m, 64); memcpy(invM, im.m, 64);
}
int itmref_mat4_inv(const float *M, float *out) { Matrix4f o; bool ok = toM(M).inv(o); memcpy(out, o.m, 64); return ok; }
void itmref_mat4_mul(const float *A, const float *B, float *out) { Matrix4f o = toM(A) * toM(B); memcpy(out, o.m, 64); }

void itmref_allocate(void *h, int onlyUpdateVisibleList)
{
    RefCtx *c = (RefCtx *)h;
    c->recon->AllocateSceneFromDepth(c->scene, c->view, c->tracking, c->rs, onlyUpdateVisibleList != 0, false);
}
void itmref_integrate(void *h) { RefCtx *c = (RefCtx *)h; c->recon->IntegrateIntoScene(c->scene, c->view, c->tracking, c->rs); }
void itmref_create_expected_depths(void *h)
{
    RefCtx *c = (RefCtx *)h;
    c->vis->CreateExpectedDepths(c->scene, c->tracking->pose_d, &c->view->calib->intrinsics_d, c->rs);
}
void itmref_create_icp_maps(void *h) { RefCtx *c = (RefCtx *)h; c->vis->CreateICPMaps(c->scene, c->view, c->tracking, c->rs); }
void itmref_find_surface(void *h)
{
    RefCtx *c = (RefCtx *)h;
    c->vis->FindSurface(c->scene, c->tracking->pose_d, &c->view->calib->intrinsics_d, c->rs);
}

// approximate raycast: ITMTrackingController::Prepare's ForwardRender branch (ITMTrackingController.cpp:40-47)
void itmref_forward_render(void *h) { RefCtx *c = (RefCtx *)h; c->vis->ForwardRender(c->scene, c->view, c->tracking, c->rs); }
void *itmref_forward_projection(void *h) { RefCtx *c = (RefCtx *)h; return c->rs->forwardProjection->GetData(MEMORYDEVICE_CPU); }
int itmref_no_missing_points(void *h) { RefCtx *c = (RefCtx *)h; return c->rs->noFwdProjMissingPoints; }

// free-view path: ITMMainEngine::GetImage (ITMLib/Engine/ITMMainEngine.cpp:272-277)
void itmref_render_free_view(void *h, const float *M, const float *intr, int renderType)
{
    RefCtx *c = (RefCtx *)h;
    ITMPose pose; pose.SetM(toM(M));
    ITMIntrinsics in; in.SetFrom(intr[0], intr[1], intr[2], intr[3], c->imgSize.x, c->imgSize.y);
    c->vis->FindVisibleBlocks(c->scene, &pose, &in, c->rsFree);
    c->vis->CreateExpectedDepths(c->scene, &pose, &in, c->rsFree);
    c->vis->RenderImage(c->scene, &pose, &in, c->rsFree, c->outImg, c->outFloat,
                        (IITMVisualisationEngine::RenderImageType)renderType);
}
int itmref_count_visible_blocks(void *h, int minId, int maxId)
{
    RefCtx *c = (RefCtx *)h; return c->vis->CountVisibleBlocks(c->scene, c->rs, minId, maxId);
}

// ---- state accessors (pointers into the reference's own buffers) ----
void *itmref_hash_table(void *h) { return ((RefCtx *)h)->scene->index.GetEntries(); }
void *itmref_vba(void *h) { return ((RefCtx *)h)->scene->localVBA.GetVoxelBlocks(); }
int *itmref_alloc_list(void *h) { return ((RefCtx *)h)->scene->localVBA.GetAllocationList(); }
int *itmref_excess_list(void *h) { return ((RefCtx *)h)->scene->index.GetExcessAllocationList(); }
int itmref_last_free_block(void *h) { return ((RefCtx *)h)->scene->localVBA.lastFreeBlockId; }
int itmref_last_free_excess(void *h) { return ((RefCtx *)h)->scene->index.GetLastFreeExcessListId(); }
int itmref_no_visible(void *h, int freeView) { RefCtx *c = (RefCtx *)h; return (freeView ? c->rsFree : c->rs)->noVisibleEntries; }
int *itmref_visible_ids(void *h, int freeView) { RefCtx *c = (RefCtx *)h; return (freeView ? c->rsFree : c->rs)->GetVisibleEntryIDs(); }
uint8_t *itmref_visible_type(void *h) { return ((RefCtx *)h)->rs->GetEntriesVisibleType(); }
float *itmref_range_image(void *h, int freeView)
{
    RefCtx *c = (RefCtx *)h; return (float *)(freeView ? c->rsFree : c->rs)->renderingRangeImage->GetData(MEMORYDEVICE_CPU);
}
float *itmref_raycast_result(void *h, int freeView)
{
    RefCtx *c = (RefCtx *)h; return (float *)(freeView ? c->rsFree : c->rs)->raycastResult->GetData(MEMORYDEVICE_CPU);
}
float *itmref_icp_points(void *h) { return (float *)((RefCtx *)h)->tracking->pointCloud->locations->GetData(MEMORYDEVICE_CPU); }
float *itmref_icp_normals(void *h) { return (float *)((RefCtx *)h)->tracking->pointCloud->colours->GetData(MEMORYDEVICE_CPU); }
uint8_t *itmref_icp_rendering(void *h) { return (uint8_t *)((RefCtx *)h)->rs->raycastImage->GetData(MEMORYDEVICE_CPU); }
uint8_t *itmref_out_image(void *h) { return (uint8_t *)((RefCtx *)h)->outImg->GetData(MEMORYDEVICE_CPU); }

} // extern "C"

// ---- ICP depth tracker ----
namespace {
// ComputeGandH is protected: a subclass that fills the protected evaluation state from raw arrays and calls it
struct TrackerProbe : public ITMDepthTracker_CPU {
    TrackerProbe(Vector2i size, TrackerIterationType *regime, const ITMLowLevelEngine *ll)
        : ITMDepthTracker_CPU(size, regime, 1, 0, 0.01f, 1e-3f, ll) {}
    int probe(const float *depth, int vW, int vH, const float *vIntr, const float *points, const float *normals, int sW, int sH,
              const float *sIntr, const float *scenePoseM, const float *approxInvPose, float dt, int type, float *f, float *nabla,
              float *hessian)
    {
        ITMTemplatedHierarchyLevel<ITMFloatImage> vl(Vector2i(vW, vH), 0, (TrackerIterationType)type, MEMORYDEVICE_CPU);
        ITMSceneHierarchyLevel sl(Vector2i(sW, sH), 0, (TrackerIterationType)type, MEMORYDEVICE_CPU);
        memcpy(vl.depth->GetData(MEMORYDEVICE_CPU), depth, sizeof(float) * (size_t)vW * vH);
        memcpy(sl.pointsMap->GetData(MEMORYDEVICE_CPU), points, sizeof(float) * 4 * (size_t)sW * sH);
        memcpy(sl.normalsMap->GetData(MEMORYDEVICE_CPU), normals, sizeof(float) * 4 * (size_t)sW * sH);
        vl.intrinsics = Vector4f(vIntr[0], vIntr[1], vIntr[2], vIntr[3]);
        sl.intrinsics = Vector4f(sIntr[0], sIntr[1], sIntr[2], sIntr[3]);
        this->levelId = 0; this->distThresh[0] = dt; this->iterationType = (TrackerIterationType)type;
        this->scenePose = toM(scenePoseM); this->sceneHierarchyLevel = &sl; this->viewHierarchyLevel = &vl;
        return this->ComputeGandH(*f, nabla, hessian, toM(approxInvPose));
    }
};
}

extern "C" {

int itmref_tracker_compute_g_and_h(const float *depth, int vW, int vH, const float *vIntr, const float *points,
                                   const float *normals, int sW, int sH, const float *sIntr, const float *scenePose,
                                   const float *approxInvPose, float distThresh, int iterationType, float *f, float *nabla,
                                   float *hessian)
{
    ITMLowLevelEngine_CPU ll;
    TrackerIterationType regime[1] = {(TrackerIterationType)iterationType};
    TrackerProbe t(Vector2i(vW, vH), regime, &ll);
    return t.probe(depth, vW, vH, vIntr, points, normals, sW, sH, sIntr, scenePose, approxInvPose, distThresh, iterationType, f,
                   nabla, hessian);
}

void itmref_filter_subsample_with_holes(const float *in, int w, int h, float *out)
{
    ITMLowLevelEngine_CPU ll;
    ITMFloatImage src(Vector2i(w, h), true, false), dst(Vector2i(w / 2, h / 2), true, false);
    memcpy(src.GetData(MEMORYDEVICE_CPU), in, sizeof(float) * (size_t)w * h);
    ll.FilterSubsampleWithHoles(&dst, &src);
    memcpy(out, dst.GetData(MEMORYDEVICE_CPU), sizeof(float) * (size_t)(w / 2) * (h / 2));
}

// ITMPose::SetM + Coerce; params = (tx, ty, tz, rx, ry, rz)
void itmref_pose_coerce(const float *Min, float *Mout, float *params)
{
    ITMPose p; p.SetM(toM(Min)); p.Coerce();
    Matrix4f m = p.GetM(); memcpy(Mout, m.m, 64);
    for (int i = 0; i < 6; i++) params[i] = p.GetParams()[i];
}
void itmref_pose_from_params(const float *params, float *Mout)
{
    ITMPose p(params[0], params[1], params[2], params[3], params[4], params[5]);
    Matrix4f m = p.GetM(); memcpy(Mout, m.m, 64);
}

// pose_pointCloud (normally written by CreateICPMaps, ITMVisualisationEngine_CPU.cpp) for maps loaded from a fixture
void itmref_set_scene_pose(void *h, const float *M) { ((RefCtx *)h)->tracking->pose_pointCloud->SetM(toM(M)); }

// The whole tracker: ITMDepthTracker::TrackCamera (ITMDepthTracker.cpp:153-211) on the context's view and tracking state
// (point cloud + pose_pointCloud from the last CreateICPMaps; pose_d is the initial guess and receives the result).
void itmref_track_camera(void *h, int levels, const int *regime, int noICPRunTillLevel, float distThresh, float terminationThreshold)
{
    RefCtx *c = (RefCtx *)h;
    ITMLowLevelEngine_CPU ll;
    TrackerIterationType r[16];
    for (int i = 0; i < levels && i < 16; i++) r[i] = (TrackerIterationType)regime[i];
    ITMDepthTracker_CPU t(c->imgSize, r, levels, noICPRunTillLevel, distThresh, terminationThreshold, &ll);
    t.TrackCamera(c->tracking, c->view);
}

} // extern "C"


// ---- meshing: ITMMeshingEngine_CPU::MeshScene on the context's scene.  out: room for maxOut ITMMesh::Triangle records
// (72 bytes each); returns noTotalTriangles (the reference's own buffer holds SDF_LOCAL_BLOCK_NUM * 64 of them).
extern "C" {
int itmref_sizeof_triangle() { return (int)sizeof(ITMMesh::Triangle); }
unsigned itmref_max_triangles() { return ITMMesh::noMaxTriangles; }
unsigned itmref_mesh_scene(void *h, void *out, unsigned maxOut)
{
    RefCtx *c = (RefCtx *)h;
    ITMMesh *mesh = new ITMMesh(MEMORYDEVICE_CPU);
    ITMMeshingEngine_CPU<ITMVoxel, ITMVoxelBlockHash> mesher;
    mesher.MeshScene(mesh, c->scene);
    unsigned n = mesh->noTotalTriangles;
    unsigned copy = n + 1 < maxOut ? n + 1 : maxOut;      // + the slot past the count (holds the overflow triangle when capped)
    memcpy(out, mesh->triangles->GetData(MEMORYDEVICE_CPU), sizeof(ITMMesh::Triangle) * (size_t)copy);
    delete mesh;
    return n;
}
// the reference's marching-cubes tables as compiled into its mesher (for tests/test_meshing.py)
void itmref_mc_tables(int *edge, int *tri)
{
    for (int i = 0; i < 256; i++) { edge[i] = edgeTable[i]; for (int k = 0; k < 16; k++) tri[i * 16 + k] = triangleTable[i][k]; }
}
}
