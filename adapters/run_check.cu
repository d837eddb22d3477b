// adapters/run_check.cu -- the drop-in EXECUTED: the reference's own objects (ITMScene, ITMView, ITMTrackingState,
// ITMRenderState_VH, all MEMORYDEVICE_CUDA) driven through the *_B200 adapter subclasses in the call order of
// ITMDenseMapper::ProcessFrame (ITMLib/Engine/ITMDenseMapper.cpp:56-81: AllocateSceneFromDepth, IntegrateIntoScene) and
// ITMTrackingController::Prepare (ITMLib/Engine/ITMTrackingController.cpp:21-58: CreateExpectedDepths, CreateICPMaps), and
// compared IN THE SAME PROCESS with the reference's ITMSceneReconstructionEngine_CPU / ITMVisualisationEngine_CPU on a
// MEMORYDEVICE_CPU scene fed the same frames.  Also: the free-view path of ITMMainEngine::GetImage (ITMMainEngine.cpp:272-277),
// the exhaustion -> std::runtime_error contract (CUDA .cu:350-358), the host mirrors DenseSLAM reads after every call
// (DS/InfiniTamDriver.h:210,344-346) and the pruning entry points.
//
// TEST INFRASTRUCTURE: built by `make -C adapters run_check` (only where /root/reference exists) into
// oracle/_ref/adapter_run_check; tests/test_gpu_adapters.py writes the frames file and runs it on the GPU box.
//
// frames file: int32 W, H, n; float intr[4], voxelSize, mu; int32 maxW; float vfMin, vfMax; then per frame
//              float depth[W*H]; uint8 rgba[W*H*4]; float M[16]; double timestamp
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ITMSceneReconstructionEngine_B200.h"
#include "ITMVisualisationEngine_B200.h"
#include "ITMViewBuilder_B200.h"
#include "ITMMeshingEngine_B200.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMSceneReconstructionEngine_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMVisualisationEngine_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMViewBuilder_CPU.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMMeshingEngine_CPU.h"

using namespace ITMLib::Engine;
using namespace ITMLib::Objects;

static int failures = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); if (++failures > 20) exit(1); } } while (0)

struct Side {
    ITMScene<ITMVoxel, ITMVoxelBlockHash> *scene;
    ITMSceneReconstructionEngine<ITMVoxel, ITMVoxelBlockHash> *recon;
    ITMVisualisationEngine<ITMVoxel, ITMVoxelBlockHash> *vis;
    ITMView *view;
    ITMTrackingState *trk;
    ITMRenderState_VH *rs, *rsFree;
    ITMUChar4Image *outImg;
    ITMFloatImage *outFloat;
    bool gpu;
};

template <class T> static std::vector<T> pull(const T *p, size_t n, bool gpu)
{
    std::vector<T> v(n);
    if (!n) return v;
    if (gpu) { if (cudaMemcpy(v.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost) != cudaSuccess) { printf("cudaMemcpy failed\n"); exit(2); } }
    else memcpy(v.data(), p, n * sizeof(T));
    return v;
}
template <class T> static void push(T *dst, const T *src, size_t n, bool gpu)
{
    if (gpu) cudaMemcpy(dst, src, n * sizeof(T), cudaMemcpyHostToDevice); else memcpy(dst, src, n * sizeof(T));
}
static MemoryDeviceType mem(bool gpu) { return gpu ? MEMORYDEVICE_CUDA : MEMORYDEVICE_CPU; }

static Side make_side(bool gpu, const ITMSceneParams *params, const ITMRGBDCalib *calib, Vector2i size)
{
    Side s; s.gpu = gpu;
    s.scene = new ITMScene<ITMVoxel, ITMVoxelBlockHash>(params, false, mem(gpu));
    if (gpu) {
        s.recon = new ITMSceneReconstructionEngine_B200<ITMVoxel, ITMVoxelBlockHash>();
        s.vis = new ITMVisualisationEngine_B200<ITMVoxel, ITMVoxelBlockHash>(size);
    } else {
        s.recon = new ITMSceneReconstructionEngine_CPU<ITMVoxel, ITMVoxelBlockHash>();
        s.vis = new ITMVisualisationEngine_CPU<ITMVoxel, ITMVoxelBlockHash>();
    }
    s.view = new ITMView(calib, size, size, gpu);
    s.trk = new ITMTrackingState(size, mem(gpu));
    s.rs = (ITMRenderState_VH *)s.vis->CreateRenderState(s.scene, size);
    s.rsFree = (ITMRenderState_VH *)s.vis->CreateRenderState(s.scene, size);
    s.outImg = new ITMUChar4Image(size, true, gpu);
    s.outFloat = new ITMFloatImage(size, true, gpu);
    s.recon->ResetScene(s.scene);
    return s;
}

static void compare_state(const Side &a, const Side &b, int frame, size_t npx)
{
    const int total = ITMVoxelBlockHash::noTotalEntries;
    EXPECT(a.scene->localVBA.lastFreeBlockId == b.scene->localVBA.lastFreeBlockId, "frame %d lastFreeBlockId %d vs %d", frame,
           a.scene->localVBA.lastFreeBlockId, b.scene->localVBA.lastFreeBlockId);
    EXPECT(a.scene->index.GetLastFreeExcessListId() == b.scene->index.GetLastFreeExcessListId(), "frame %d lastFreeExcessListId", frame);
    EXPECT(a.rs->noVisibleEntries == b.rs->noVisibleEntries, "frame %d noVisibleEntries %d vs %d", frame, a.rs->noVisibleEntries, b.rs->noVisibleEntries);
    std::vector<ITMHashEntry> ha = pull(a.scene->index.GetEntries(), total, a.gpu), hb = pull(b.scene->index.GetEntries(), total, b.gpu);
    int bad = 0, live = 0;
    for (int i = 0; i < total; i++) {
        if (ha[i].pos.x != hb[i].pos.x || ha[i].pos.y != hb[i].pos.y || ha[i].pos.z != hb[i].pos.z || ha[i].offset != hb[i].offset || ha[i].ptr != hb[i].ptr) bad++;
        if (ha[i].ptr >= 0) live++;
    }
    EXPECT(bad == 0, "frame %d: %d hash entries differ", frame, bad);
    const int nv = a.rs->noVisibleEntries < b.rs->noVisibleEntries ? a.rs->noVisibleEntries : b.rs->noVisibleEntries;
    std::vector<int> va = pull(a.rs->GetVisibleEntryIDs(), nv, a.gpu), vb = pull(b.rs->GetVisibleEntryIDs(), nv, b.gpu);
    EXPECT(va == vb, "frame %d: visible lists differ", frame);
    std::vector<uchar> ta = pull(a.rs->GetEntriesVisibleType(), total, a.gpu), tb = pull(b.rs->GetEntriesVisibleType(), total, b.gpu);
    EXPECT(ta == tb, "frame %d: entriesVisibleType differs", frame);
    // voxels of every allocated block (sdf, w_depth, clr, w_color; the 8th byte is padding)
    long vbad = 0;
    std::vector<ITMVoxel> xa(SDF_BLOCK_SIZE3), xb(SDF_BLOCK_SIZE3);
    for (int i = 0; i < total && bad == 0; i++) {
        if (ha[i].ptr < 0) continue;
        xa = pull(a.scene->localVBA.GetVoxelBlocks() + (size_t)ha[i].ptr * SDF_BLOCK_SIZE3, SDF_BLOCK_SIZE3, a.gpu);
        xb = pull(b.scene->localVBA.GetVoxelBlocks() + (size_t)hb[i].ptr * SDF_BLOCK_SIZE3, SDF_BLOCK_SIZE3, b.gpu);
        for (int k = 0; k < SDF_BLOCK_SIZE3; k++)
            if (xa[k].sdf != xb[k].sdf || xa[k].w_depth != xb[k].w_depth || xa[k].w_color != xb[k].w_color || xa[k].clr.x != xb[k].clr.x ||
                xa[k].clr.y != xb[k].clr.y || xa[k].clr.z != xb[k].clr.z) vbad++;
    }
    EXPECT(vbad == 0, "frame %d: %ld voxels differ", frame, vbad);
    // images
    std::vector<Vector4f> ra = pull(a.rs->raycastResult->GetData(mem(a.gpu)), npx, a.gpu), rb = pull(b.rs->raycastResult->GetData(mem(b.gpu)), npx, b.gpu);
    EXPECT(memcmp(ra.data(), rb.data(), npx * sizeof(Vector4f)) == 0, "frame %d: raycastResult differs", frame);
    std::vector<Vector4f> pa = pull(a.trk->pointCloud->locations->GetData(mem(a.gpu)), npx, a.gpu), pb = pull(b.trk->pointCloud->locations->GetData(mem(b.gpu)), npx, b.gpu);
    EXPECT(memcmp(pa.data(), pb.data(), npx * sizeof(Vector4f)) == 0, "frame %d: ICP points differ", frame);
    std::vector<Vector4f> na = pull(a.trk->pointCloud->colours->GetData(mem(a.gpu)), npx, a.gpu), nb = pull(b.trk->pointCloud->colours->GetData(mem(b.gpu)), npx, b.gpu);
    EXPECT(memcmp(na.data(), nb.data(), npx * sizeof(Vector4f)) == 0, "frame %d: ICP normals differ", frame);
    std::vector<Vector4u> ga = pull(a.rs->raycastImage->GetData(mem(a.gpu)), npx, a.gpu), gb = pull(b.rs->raycastImage->GetData(mem(b.gpu)), npx, b.gpu);
    EXPECT(memcmp(ga.data(), gb.data(), npx * sizeof(Vector4u)) == 0, "frame %d: rendering differs", frame);
    Matrix4f Ma = a.trk->pose_pointCloud->GetM(), Mb = b.trk->pose_pointCloud->GetM();
    EXPECT(memcmp(Ma.m, Mb.m, 64) == 0, "frame %d: pose_pointCloud differs", frame);
    printf("frame %d: %d blocks allocated, %d visible, lastFreeBlockId %d -- CPU engine == B200 adapter\n", frame, live, a.rs->noVisibleEntries,
           a.scene->localVBA.lastFreeBlockId);
}

int main(int argc, char **argv)
{
    if (argc < 2) { printf("usage: %s frames.bin\n", argv[0]); return 2; }
    FILE *f = fopen(argv[1], "rb");
    if (!f) { printf("cannot open %s\n", argv[1]); return 2; }
    int W, H, n, maxW; float intr[4], voxelSize, mu, vfMin, vfMax;
    if (fread(&W, 4, 1, f) != 1 || fread(&H, 4, 1, f) != 1 || fread(&n, 4, 1, f) != 1 || fread(intr, 4, 4, f) != 4 || fread(&voxelSize, 4, 1, f) != 1 ||
        fread(&mu, 4, 1, f) != 1 || fread(&maxW, 4, 1, f) != 1 || fread(&vfMin, 4, 1, f) != 1 || fread(&vfMax, 4, 1, f) != 1) return 2;
    const size_t npx = (size_t)W * H;
    const Vector2i size(W, H);
    ITMSceneParams params(mu, maxW, voxelSize, vfMin, vfMax, false);
    ITMRGBDCalib calib;
    calib.intrinsics_d.SetFrom(intr[0], intr[1], intr[2], intr[3], W, H);
    calib.intrinsics_rgb.SetFrom(intr[0], intr[1], intr[2], intr[3], W, H);
    Side cpu = make_side(false, &params, &calib, size), gpu = make_side(true, &params, &calib, size);
    Side *sides[2] = {&cpu, &gpu};
    std::vector<float> depth(npx); std::vector<uchar> rgba(npx * 4); float M[16]; double ts;
    Matrix4f lastM;
    for (int i = 0; i < n; i++) {
        if (fread(depth.data(), 4, npx, f) != npx || fread(rgba.data(), 1, npx * 4, f) != npx * 4 || fread(M, 4, 16, f) != 16 || fread(&ts, 8, 1, f) != 1) return 2;
        for (int k = 0; k < 16; k++) lastM.m[k] = M[k];
        for (Side *s : sides) {
            push(s->view->depth->GetData(mem(s->gpu)), depth.data(), npx, s->gpu);
            push(s->view->rgb->GetData(mem(s->gpu)), (const Vector4u *)rgba.data(), npx, s->gpu);
            s->view->mTimeStamp = ts;
            s->trk->pose_d->SetM(lastM);
            // ITMDenseMapper::ProcessFrame
            s->recon->AllocateSceneFromDepth(s->scene, s->view, s->trk, s->rs);
            s->recon->IntegrateIntoScene(s->scene, s->view, s->trk, s->rs);
            // ITMTrackingController::Prepare
            s->vis->CreateExpectedDepths(s->scene, s->trk->pose_d, &(s->view->calib->intrinsics_d), s->rs);
            s->vis->CreateICPMaps(s->scene, s->view, s->trk, s->rs);
        }
        compare_state(cpu, gpu, i, npx);
    }
    fclose(f);

    // free-view path (ITMMainEngine::GetImage): FindVisibleBlocks, CreateExpectedDepths, RenderImage
    {
        ITMPose pose; pose.SetM(lastM);
        const IITMVisualisationEngine::RenderImageType types[3] = {IITMVisualisationEngine::RENDER_SHADED_GREYSCALE,
            IITMVisualisationEngine::RENDER_COLOUR_FROM_VOLUME, IITMVisualisationEngine::RENDER_COLOUR_FROM_NORMAL};
        for (int t = 0; t < 3; t++) {
            for (Side *s : sides) {
                s->vis->FindVisibleBlocks(s->scene, &pose, &(s->view->calib->intrinsics_d), s->rsFree);
                s->vis->CreateExpectedDepths(s->scene, &pose, &(s->view->calib->intrinsics_d), s->rsFree);
                s->vis->RenderImage(s->scene, &pose, &(s->view->calib->intrinsics_d), s->rsFree, s->outImg, s->outFloat, types[t]);
            }
            EXPECT(cpu.rsFree->noVisibleEntries == gpu.rsFree->noVisibleEntries, "free view: %d vs %d visible", cpu.rsFree->noVisibleEntries, gpu.rsFree->noVisibleEntries);
            std::vector<Vector4u> a = pull(cpu.outImg->GetData(MEMORYDEVICE_CPU), npx, false), b = pull(gpu.outImg->GetData(MEMORYDEVICE_CUDA), npx, true);
            EXPECT(memcmp(a.data(), b.data(), npx * 4) == 0, "free view: rendered image of type %d differs", (int)types[t]);
        }
        printf("free view: %d visible blocks, 3 render types -- CPU engine == B200 adapter\n", gpu.rsFree->noVisibleEntries);
        EXPECT(gpu.vis->CountVisibleBlocks(gpu.scene, gpu.rs, 0, SDF_LOCAL_BLOCK_NUM) == cpu.vis->CountVisibleBlocks(cpu.scene, cpu.rs, 0, SDF_LOCAL_BLOCK_NUM), "CountVisibleBlocks");
    }

    // view builder: raw int16 millimetres -> float metres (DS/InfiniTamDriver.cpp:280-288 path)
    {
        calib.disparityCalib.SetFrom(1.0f / 1000.0f, 0.0f, ITMDisparityCalib::TRAFO_AFFINE);
        ITMViewBuilder_CPU vbc(&calib); ITMViewBuilder_B200 vbg(&calib);
        ITMUChar4Image rgbImg(size, true, false); ITMShortImage raw(size, true, false);
        memcpy(rgbImg.GetData(MEMORYDEVICE_CPU), rgba.data(), npx * 4);
        short *r = raw.GetData(MEMORYDEVICE_CPU);
        for (size_t k = 0; k < npx; k++) r[k] = depth[k] > 0 ? (short)(depth[k] * 1000.0f + 0.5f) : 0;
        ITMView *vc = NULL, *vg = NULL;
        vbc.UpdateView(&vc, &rgbImg, &raw, 1.5, false);
        vbg.UpdateView(&vg, &rgbImg, &raw, 1.5, false);
        std::vector<float> a = pull(vc->depth->GetData(MEMORYDEVICE_CPU), npx, false), b = pull(vg->depth->GetData(MEMORYDEVICE_CUDA), npx, true);
        EXPECT(memcmp(a.data(), b.data(), npx * 4) == 0, "view builder: depth differs");
        EXPECT(vg->mTimeStamp == 1.5, "view builder: timestamp");
        printf("view builder: UpdateView -- CPU builder == B200 adapter\n");
        delete vc; delete vg;
    }

    // meshing (ITMMainEngine::SaveSceneToMesh path): ITMMesh on the device through the adapter vs ITMMeshingEngine_CPU
    {
        ITMMesh *mc = new ITMMesh(MEMORYDEVICE_CPU), *mg = new ITMMesh(MEMORYDEVICE_CUDA);
        ITMMeshingEngine_CPU<ITMVoxel, ITMVoxelBlockHash> meshCpu; ITMMeshingEngine_B200<ITMVoxel, ITMVoxelBlockHash> meshGpu;
        meshCpu.MeshScene(mc, cpu.scene);
        meshGpu.MeshScene(mg, gpu.scene);
        EXPECT(mc->noTotalTriangles == mg->noTotalTriangles && mc->noTotalTriangles > 100, "meshing: %u vs %u triangles", mc->noTotalTriangles, mg->noTotalTriangles);
        std::vector<ITMMesh::Triangle> a = pull(mc->triangles->GetData(MEMORYDEVICE_CPU), mc->noTotalTriangles, false),
                                       b = pull(mg->triangles->GetData(MEMORYDEVICE_CUDA), mc->noTotalTriangles, true);
        EXPECT(memcmp(a.data(), b.data(), a.size() * sizeof(ITMMesh::Triangle)) == 0, "meshing: triangles differ");
        printf("meshing: %u triangles -- ITMMeshingEngine_CPU == B200 adapter\n", mg->noTotalTriangles);
        delete mc; delete mg;
    }

    // pruning entry points execute through the adapter and keep the host mirrors valid (the CPU engine throws for these)
    {
        const int before = gpu.scene->localVBA.lastFreeBlockId;
        for (int k = 0; k < 3; k++) gpu.recon->SlideWindow(gpu.scene, gpu.rs, 2);
        const int after = gpu.scene->localVBA.lastFreeBlockId;
        EXPECT(after > before, "SlideWindow freed nothing (%d -> %d)", before, after);
        EXPECT((long)gpu.recon->GetDecayedBlockCount() == (long)(after - before), "GetDecayedBlockCount %ld vs %d", (long)gpu.recon->GetDecayedBlockCount(), after - before);
        gpu.recon->Decay(gpu.scene, gpu.rs, 1, 0, false);
        bool threw = false;
        try { cpu.recon->SlideWindow(cpu.scene, cpu.rs, 2); } catch (const std::runtime_error &) { threw = true; }
        EXPECT(threw, "the reference CPU engine was expected to throw for SlideWindow");
        printf("pruning: SlideWindow x3 freed %d blocks, GetDecayedBlockCount agrees with the lastFreeBlockId mirror\n", after - before);
    }

    // exhaustion: the reference throws std::runtime_error with this text (CUDA .cu:350-358)
    {
        gpu.recon->ResetScene(gpu.scene);
        gpu.rs->noVisibleEntries = 0;
        cudaMemset(gpu.rs->GetEntriesVisibleType(), 0, ITMVoxelBlockHash::noTotalEntries);
        gpu.scene->localVBA.lastFreeBlockId = 9;          // ten free blocks left
        std::string what;
        try { gpu.recon->AllocateSceneFromDepth(gpu.scene, gpu.view, gpu.trk, gpu.rs); } catch (const std::runtime_error &e) { what = e.what(); }
        EXPECT(what.find("run out of space in the Voxel Block Array") != std::string::npos, "exhaustion message: '%s'", what.c_str());
        EXPECT(gpu.scene->localVBA.lastFreeBlockId < 0, "exhaustion: lastFreeBlockId mirror %d", gpu.scene->localVBA.lastFreeBlockId);
        printf("exhaustion: std::runtime_error(\"%s\")\n", what.c_str());
    }

    if (failures) { printf("%d FAILURES\n", failures); return 1; }
    printf("adapters: OK\n");
    return 0;
}
