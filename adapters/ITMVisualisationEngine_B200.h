// ITMVisualisationEngine_B200.h -- drop-in subclass of the reference's visualisation-engine interface
// (ITMLib/Engine/ITMVisualisationEngine.h:60-124) forwarding to libdsb200's C ABI.  Replaces
// ITMVisualisationEngine_CUDA<TVoxel,ITMVoxelBlockHash> in ITMLib/Engine/ITMMainEngine.cpp:57-81.
#pragma once

#include "ITMLib/Engine/ITMVisualisationEngine.h"
#include "ITMSceneReconstructionEngine_B200.h"

namespace ITMLib
{
	namespace Engine
	{
		template<class TVoxel, class TIndex>
		class ITMVisualisationEngine_B200;

		template<class TVoxel>
		class ITMVisualisationEngine_B200<TVoxel, ITMVoxelBlockHash> : public ITMVisualisationEngine<TVoxel, ITMVoxelBlockHash>
		{
		private:
			dsb200_vis *engine;   // scratch only; `mutable` semantics live behind the handle, so the virtuals stay const

			static void intr4(float *dst, const ITMIntrinsics *intrinsics)
			{
				const Vector4f p = intrinsics->projectionParamsSimple.all;
				dst[0] = p.x; dst[1] = p.y; dst[2] = p.z; dst[3] = p.w;
			}
			// the scene is const in this interface; the C ABI takes it by const pointer too
			static dsb200_scene wrapc(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				return b200::wrap(const_cast<ITMScene<TVoxel, ITMVoxelBlockHash>*>(scene));
			}

		public:
			explicit ITMVisualisationEngine_B200(const Vector2i &imgSize, void *cudaStream = NULL) : engine(NULL)
			{
				b200::check(dsb200_vis_create(&engine, SDF_LOCAL_BLOCK_NUM, imgSize.x, imgSize.y, cudaStream));
			}
			~ITMVisualisationEngine_B200(void) { dsb200_vis_destroy(engine); }

			dsb200_vis *Handle(void) const { return engine; }

			ITMRenderState_VH *CreateRenderState(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const Vector2i &imgSize) const
			{
				// exactly what the reference does (CUDA .cu:137-142): the render state stays a reference object
				return new ITMRenderState_VH(ITMVoxelBlockHash::noTotalEntries, imgSize, scene->sceneParams->viewFrustum_min,
					scene->sceneParams->viewFrustum_max, MEMORYDEVICE_CUDA);
			}

			void FindVisibleBlocks(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMPose *pose, const ITMIntrinsics *intrinsics,
				ITMRenderState *renderState) const
			{
				ITMRenderState_VH *rs_vh = (ITMRenderState_VH*)renderState;
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap(rs_vh);
				float M[16], in[4]; b200::copy16(M, pose->GetM()); intr4(in, intrinsics);
				b200::check(dsb200_find_visible_blocks(engine, &s, M, in, &r));
				rs_vh->noVisibleEntries = r.no_visible;
			}

			int CountVisibleBlocks(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMRenderState *renderState, int minBlockId = 0,
				int maxBlockId = SDF_LOCAL_BLOCK_NUM) const
			{
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				int n = 0;
				b200::check(dsb200_count_visible_blocks(engine, &s, &r, minBlockId, maxBlockId, &n));
				return n;
			}

			void CreateExpectedDepths(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMPose *pose, const ITMIntrinsics *intrinsics,
				ITMRenderState *renderState) const
			{
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				float M[16], in[4]; b200::copy16(M, pose->GetM()); intr4(in, intrinsics);
				b200::check(dsb200_create_expected_depths(engine, &s, M, in, &r));
			}

			void RenderImage(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMPose *pose, const ITMIntrinsics *intrinsics,
				const ITMRenderState *renderState, ITMUChar4Image *outputImage, ITMFloatImage *outputFloatImage,
				IITMVisualisationEngine::RenderImageType type = IITMVisualisationEngine::RENDER_SHADED_GREYSCALE) const
			{
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				float M[16], invM[16], in[4];
				b200::copy16(M, pose->GetM()); b200::copy16(invM, pose->GetInvM()); intr4(in, intrinsics);
				b200::check(dsb200_render_image(engine, &s, M, invM, in, &r,
					reinterpret_cast<uint8_t*>(outputImage->GetData(MEMORYDEVICE_CUDA)),
					outputFloatImage ? outputFloatImage->GetData(MEMORYDEVICE_CUDA) : NULL, (int)type));
			}

			void FindSurface(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMPose *pose, const ITMIntrinsics *intrinsics,
				const ITMRenderState *renderState) const
			{
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				float invM[16], in[4]; b200::copy16(invM, pose->GetInvM()); intr4(in, intrinsics);
				b200::check(dsb200_find_surface(engine, &s, invM, in, &r));
			}

			void CreatePointCloud(const ITMScene<TVoxel, ITMVoxelBlockHash>*, const ITMView*, ITMTrackingState*, ITMRenderState*, bool) const
			{
				b200::check(dsb200_create_point_cloud(engine));   // colour tracker: never selected by DenseSLAM -> throws
			}

			void CreateICPMaps(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMView *view, ITMTrackingState *trackingState,
				ITMRenderState *renderState) const
			{
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				dsb200_view v = b200::wrap(view); dsb200_tracking t = b200::wrap(trackingState);
				b200::check(dsb200_create_icp_maps(engine, &s, &v, &t, &r));
				trackingState->pose_pointCloud->SetFrom(trackingState->pose_d);   // CUDA .cu:470
			}

			void ForwardRender(const ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const ITMView *view, ITMTrackingState *trackingState,
				ITMRenderState *renderState) const
			{
				dsb200_scene s = wrapc(scene); dsb200_render_state r = b200::wrap((ITMRenderState_VH*)renderState);
				dsb200_view v = b200::wrap(view); dsb200_tracking t = b200::wrap(trackingState);
				b200::check(dsb200_forward_render(engine, &s, &v, &t, &r));
			}
		};
	}
}
