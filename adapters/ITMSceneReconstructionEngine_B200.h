// ITMSceneReconstructionEngine_B200.h -- drop-in subclass of the reference's reconstruction-engine interface
// (ITMLib/Engine/ITMSceneReconstructionEngine.h:36-96) that forwards every virtual to libdsb200's C ABI.
//
// Compile this inside the reference tree (include path = InfiniTAM/), link with -ldsb200.  It replaces
// ITMSceneReconstructionEngine_CUDA<TVoxel,ITMVoxelBlockHash> in the switch of
// ITMLib/Engine/ITMDenseMapper.cpp:16-37.  The objects (ITMScene, ITMRenderState_VH,
// ITMView, ITMTrackingState) stay the reference's own, allocated with MEMORYDEVICE_CUDA; this adapter only
// copies their raw device pointers into the POD structs of include/dsb200.h.
#pragma once

#include <stdexcept>
#include <string>

#include "ITMLib/Engine/ITMSceneReconstructionEngine.h"
#include "ITMLib/Objects/ITMRenderState_VH.h"
#include "dsb200.h"

namespace ITMLib
{
	namespace Engine
	{
		namespace b200
		{
			inline void check(int rc)
			{
				if (rc == DSB200_OK) return;
				// same failure style as the reference: logic errors throw std::runtime_error (CUDA .cu:350-358)
				throw std::runtime_error(std::string(dsb200_last_error()));
			}

			inline void copy16(float *dst, const Matrix4f &M) { for (int i = 0; i < 16; i++) dst[i] = M.m[i]; }

			template<class TVoxel>
			inline dsb200_scene wrap(ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				static_assert(sizeof(TVoxel) == sizeof(dsb200_voxel), "libdsb200 is built for ITMVoxel_s_rgb (8 bytes)");
				static_assert(sizeof(ITMHashEntry) == sizeof(dsb200_hash_entry), "ITMHashEntry layout");
				dsb200_scene s;
				s.voxels = reinterpret_cast<dsb200_voxel*>(scene->localVBA.GetVoxelBlocks());
				s.hash = reinterpret_cast<dsb200_hash_entry*>(scene->index.GetEntries());
				s.alloc_list = scene->localVBA.GetAllocationList();
				s.excess_list = scene->index.GetExcessAllocationList();
				s.num_blocks = SDF_LOCAL_BLOCK_NUM; s.bucket_num = SDF_BUCKET_NUM; s.excess_size = SDF_EXCESS_LIST_SIZE;
				s.last_free_block = scene->localVBA.lastFreeBlockId;
				s.last_free_excess = scene->index.GetLastFreeExcessListId();
				s.params.voxel_size = scene->sceneParams->voxelSize; s.params.mu = scene->sceneParams->mu;
				s.params.max_w = scene->sceneParams->maxW;
				s.params.vf_min = scene->sceneParams->viewFrustum_min; s.params.vf_max = scene->sceneParams->viewFrustum_max;
				s.params.stop_integrating_at_max_w = scene->sceneParams->stopIntegratingAtMaxW ? 1 : 0;
				s.use_swapping = scene->useSwapping ? 1 : 0;
				return s;
			}

			// host mirrors the callers read after every call (DS/InfiniTamDriver.h:210,344-346)
			template<class TVoxel>
			inline void unwrap(const dsb200_scene &s, ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				scene->localVBA.lastFreeBlockId = s.last_free_block;
				scene->index.SetLastFreeExcessListId(s.last_free_excess);
			}

			inline dsb200_render_state wrap(ITMRenderState_VH *rs)
			{
				dsb200_render_state r;
				r.visible_ids = rs->GetVisibleEntryIDs();
				r.visible_type = rs->GetEntriesVisibleType();
				r.visible_capacity = SDF_LOCAL_BLOCK_NUM;
				r.no_visible = rs->noVisibleEntries;
				r.range_image = reinterpret_cast<float*>(rs->renderingRangeImage->GetData(MEMORYDEVICE_CUDA));
				r.raycast_result = reinterpret_cast<float*>(rs->raycastResult->GetData(MEMORYDEVICE_CUDA));
				r.raycast_image = reinterpret_cast<uint8_t*>(rs->raycastImage->GetData(MEMORYDEVICE_CUDA));
				r.width = rs->raycastResult->noDims.x; r.height = rs->raycastResult->noDims.y;
				return r;
			}

			inline dsb200_view wrap(const ITMView *view)
			{
				dsb200_view v;
				v.depth = view->depth->GetData(MEMORYDEVICE_CUDA);
				v.rgb = reinterpret_cast<const uint8_t*>(view->rgb->GetData(MEMORYDEVICE_CUDA));
				v.depth_w = view->depth->noDims.x; v.depth_h = view->depth->noDims.y;
				v.rgb_w = view->rgb->noDims.x; v.rgb_h = view->rgb->noDims.y;
				const Vector4f pd = view->calib->intrinsics_d.projectionParamsSimple.all;
				const Vector4f pr = view->calib->intrinsics_rgb.projectionParamsSimple.all;
				v.intr_d[0] = pd.x; v.intr_d[1] = pd.y; v.intr_d[2] = pd.z; v.intr_d[3] = pd.w;
				v.intr_rgb[0] = pr.x; v.intr_rgb[1] = pr.y; v.intr_rgb[2] = pr.z; v.intr_rgb[3] = pr.w;
				copy16(v.rgb_to_depth_inv, view->calib->trafo_rgb_to_depth.calib_inv);
				v.timestamp = view->mTimeStamp;
				return v;
			}

			inline dsb200_tracking wrap(const ITMTrackingState *ts)
			{
				dsb200_tracking t;
				copy16(t.M_d, ts->pose_d->GetM());
				copy16(t.invM_d, ts->pose_d->GetInvM());
				t.requires_full_rendering = ts->requiresFullRendering ? 1 : 0;
				t.icp_points = reinterpret_cast<float*>(ts->pointCloud->locations->GetData(MEMORYDEVICE_CUDA));
				t.icp_normals = reinterpret_cast<float*>(ts->pointCloud->colours->GetData(MEMORYDEVICE_CUDA));
				return t;
			}
		}

		template<class TVoxel, class TIndex>
		class ITMSceneReconstructionEngine_B200;   // only the voxel-block-hash index is on the hot path

		template<class TVoxel>
		class ITMSceneReconstructionEngine_B200<TVoxel, ITMVoxelBlockHash> : public ITMSceneReconstructionEngine<TVoxel, ITMVoxelBlockHash>
		{
		private:
			dsb200_recon *engine;

		public:
			explicit ITMSceneReconstructionEngine_B200(void *cudaStream = NULL) : engine(NULL)
			{
				b200::check(dsb200_recon_create(&engine, SDF_LOCAL_BLOCK_NUM, SDF_BUCKET_NUM, SDF_EXCESS_LIST_SIZE, cudaStream));
			}
			~ITMSceneReconstructionEngine_B200(void) { dsb200_recon_destroy(engine); }

			dsb200_recon *Handle(void) { return engine; }

			void ResetScene(ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				dsb200_scene s = b200::wrap(scene);
				b200::check(dsb200_reset_scene(engine, &s));
				b200::unwrap(s, scene);
			}

			void AllocateSceneFromDepth(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMView *view, const ITMTrackingState *trackingState,
				const ITMRenderState *renderState, bool onlyUpdateVisibleList = false, bool isDefusion = false)
			{
				ITMRenderState_VH *rs_vh = (ITMRenderState_VH*)renderState;   // same cast as the reference (.cu:160)
				dsb200_scene s = b200::wrap(scene); dsb200_render_state r = b200::wrap(rs_vh);
				dsb200_view v = b200::wrap(view); dsb200_tracking t = b200::wrap(trackingState);
				int rc = dsb200_allocate_from_depth(engine, &s, &v, &t, &r, onlyUpdateVisibleList ? 1 : 0, isDefusion ? 1 : 0);
				b200::unwrap(s, scene); rs_vh->noVisibleEntries = r.no_visible;
				b200::check(rc);   // exhaustion -> std::runtime_error with the reference's message
			}

			void IntegrateIntoScene(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMView *view, const ITMTrackingState *trackingState,
				const ITMRenderState *renderState)
			{
				dsb200_weight_params w;
				WeightParams p = this->GetFusionWeightParams();
				w.depth_weighting = p.depthWeighting ? 1 : 0; w.max_new_w = p.maxNewW; w.max_distance = p.maxDistance;
				b200::check(dsb200_set_fusion_weight_params(engine, &w));
				dsb200_scene s = b200::wrap(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				dsb200_view v = b200::wrap(view); dsb200_tracking t = b200::wrap(trackingState);
				b200::check(dsb200_integrate(engine, &s, &v, &t, &r));
			}

			void DeIntegrateIntoScene(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMView *view, const ITMTrackingState *trackingState)
			{
				dsb200_scene s = b200::wrap(scene); dsb200_view v = b200::wrap(view); dsb200_tracking t = b200::wrap(trackingState);
				b200::check(dsb200_deintegrate(engine, &s, &v, &t));
			}

			void Decay(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMRenderState *renderState, int maxWeight, int minAge, bool forceAllVoxels)
			{
				dsb200_scene s = b200::wrap(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				b200::check(dsb200_decay(engine, &s, &r, maxWeight, minAge, forceAllVoxels ? 1 : 0));
				b200::unwrap(s, scene);
			}

			void DecayDefusionPart(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMRenderState *renderState, int maxWeight, int minAge, bool forceAllVoxels)
			{
				dsb200_scene s = b200::wrap(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				b200::check(dsb200_decay_defusion(engine, &s, &r, maxWeight, minAge, forceAllVoxels ? 1 : 0));
				b200::unwrap(s, scene);
			}

			void SlideWindow(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMRenderState *renderState, int maxAge)
			{
				dsb200_scene s = b200::wrap(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				b200::check(dsb200_slide_window(engine, &s, &r, maxAge));
				b200::unwrap(s, scene);
			}

			void SlideWindowDefusionPart(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMRenderState *renderState, int maxAge, int maxSize)
			{
				dsb200_scene s = b200::wrap(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				b200::check(dsb200_slide_window_defusion(engine, &s, &r, maxAge, maxSize));
				b200::unwrap(s, scene);
			}

			size_t GetDecayedBlockCount() { return dsb200_decayed_block_count(engine); }
		};
	}
}
