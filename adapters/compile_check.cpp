// compile_check.cpp -- instantiates both adapters against the UNMODIFIED reference headers (run where
// /root/reference exists: `make -C adapters check`).  Proves the adapters override every pure virtual of the
// two interfaces with matching signatures, and that the reference's struct layouts match include/dsb200.h.
#include "ITMSceneReconstructionEngine_B200.h"
#include "ITMVisualisationEngine_B200.h"
#include "ITMDepthTracker_B200.h"
#include "ITMViewBuilder_B200.h"
#include "ITMSwappingEngine_B200.h"
#include "ITMMeshingEngine_B200.h"
#include "ITMLib/Engine/DeviceSpecific/CPU/ITMLowLevelEngine_CPU.h"

using namespace ITMLib::Engine;

template class ITMLib::Engine::ITMSceneReconstructionEngine_B200<ITMVoxel, ITMVoxelIndex>;
template class ITMLib::Engine::ITMVisualisationEngine_B200<ITMVoxel, ITMVoxelIndex>;
template class ITMLib::Engine::ITMSwappingEngine_B200<ITMVoxel, ITMVoxelIndex>;
template class ITMLib::Engine::ITMMeshingEngine_B200<ITMVoxel, ITMVoxelIndex>;

static_assert(sizeof(ITMVoxel) == 8, "ITMVoxel_s_rgb");
static_assert(sizeof(ITMHashEntry) == 20, "ITMHashEntry");
static_assert(SDF_LOCAL_BLOCK_NUM == DSB200_DEFAULT_LOCAL_BLOCKS && SDF_BUCKET_NUM == DSB200_DEFAULT_BUCKETS &&
              SDF_EXCESS_LIST_SIZE == DSB200_DEFAULT_EXCESS, "default capacities");
static_assert((int)IITMVisualisationEngine::RENDER_SHADED_GREYSCALE == DSB200_RENDER_SHADED_GREYSCALE &&
              (int)IITMVisualisationEngine::RENDER_COLOUR_FROM_VOLUME == DSB200_RENDER_COLOUR_FROM_VOLUME &&
              (int)IITMVisualisationEngine::RENDER_COLOUR_FROM_NORMAL == DSB200_RENDER_COLOUR_FROM_NORMAL &&
              (int)IITMVisualisationEngine::RENDER_DEPTH_MAP == DSB200_RENDER_DEPTH_MAP, "RenderImageType values");

// what ITMDenseMapper / ITMMainEngine would do with deviceType == DEVICE_B200 (never executed here: no GPU)
ITMSceneReconstructionEngine<ITMVoxel, ITMVoxelIndex> *make_recon() { return new ITMSceneReconstructionEngine_B200<ITMVoxel, ITMVoxelIndex>(); }
IITMVisualisationEngine *make_vis(Vector2i imgSize) { return new ITMVisualisationEngine_B200<ITMVoxel, ITMVoxelIndex>(imgSize); }
// ITMTrackerFactory.h:147 with deviceType == DEVICE_B200 (the low-level engine would be ITMLowLevelEngine_CUDA there)
ITMTracker *make_tracker(Vector2i imgSize, TrackerIterationType *regime, const ITMLowLevelEngine *lowLevelEngine)
{
	return new ITMDepthTracker_B200(imgSize, regime, 5, 0, 0.01f, 1e-3f, lowLevelEngine);
}
static_assert((int)TRACKER_ITERATION_ROTATION == DSB200_TRACKER_ITERATION_ROTATION && (int)TRACKER_ITERATION_TRANSLATION == DSB200_TRACKER_ITERATION_TRANSLATION &&
              (int)TRACKER_ITERATION_BOTH == DSB200_TRACKER_ITERATION_BOTH && (int)TRACKER_ITERATION_NONE == DSB200_TRACKER_ITERATION_NONE, "TrackerIterationType values");
// ITMMainEngine.cpp:57-81 / DenseSLAM InfiniTamDriver.cpp:286 with deviceType == DEVICE_B200
ITMViewBuilder *make_view_builder(const ITMRGBDCalib *calib) { return new ITMViewBuilder_B200(calib); }
// ITMDenseMapper.cpp:20-33 with deviceType == DEVICE_B200 and settings->useSwapping
ITMSwappingEngine<ITMVoxel, ITMVoxelIndex> *make_swapping() { return new ITMSwappingEngine_B200<ITMVoxel, ITMVoxelIndex>(); }
static_assert(SDF_TRANSFER_BLOCK_NUM == DSB200_DEFAULT_TRANSFER_BLOCKS && sizeof(ITMHashSwapState) == 1, "swapping constants");
// ITMMainEngine.cpp:57-81 with deviceType == DEVICE_B200 (meshingEngine)
ITMMeshingEngine<ITMVoxel, ITMVoxelIndex> *make_mesher() { return new ITMMeshingEngine_B200<ITMVoxel, ITMVoxelIndex>(); }
