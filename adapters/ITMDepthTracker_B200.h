// ITMDepthTracker_B200.h -- drop-in subclass of the reference's ICP depth tracker (ITMLib/Engine/ITMDepthTracker.h:26-70)
// whose one device-specific virtual, ComputeGandH, forwards to libdsb200.  Replaces ITMDepthTracker_CUDA
// (ITMLib/Engine/DeviceSpecific/CUDA/ITMDepthTracker_CUDA.{h,cu}) in ITMLib/Engine/ITMTrackerFactory.h:147,270.
// The Levenberg-Marquardt loop, the pyramid (lowLevelEngine->FilterSubsampleWithHoles on MEMORYDEVICE_CUDA images) and
// the pose update stay the reference's own host code (ITMDepthTracker.cpp).
#pragma once

#include "ITMLib/Engine/ITMDepthTracker.h"
#include "ITMSceneReconstructionEngine_B200.h"

namespace ITMLib
{
	namespace Engine
	{
		class ITMDepthTracker_B200 : public ITMDepthTracker
		{
		private:
			dsb200_tracker *tracker;   // reduction scratch + stream; the images stay the reference's MemoryBlocks

		protected:
			int ComputeGandH(float &f, float *nabla, float *hessian, Matrix4f approxInvPose)
			{
				const Vector2i viewSize = viewHierarchyLevel->depth->noDims, sceneSize = sceneHierarchyLevel->pointsMap->noDims;
				const Vector4f vi = viewHierarchyLevel->intrinsics, si = sceneHierarchyLevel->intrinsics;
				const float viewIntr[4] = { vi.x, vi.y, vi.z, vi.w }, sceneIntr[4] = { si.x, si.y, si.z, si.w };
				float pose[16], inv[16]; int32_t noValidPoints = 0;
				b200::copy16(pose, scenePose); b200::copy16(inv, approxInvPose);
				b200::check(dsb200_tracker_compute_g_and_h_on(tracker,
					viewHierarchyLevel->depth->GetData(MEMORYDEVICE_CUDA), viewSize.x, viewSize.y, viewIntr,
					reinterpret_cast<const float*>(sceneHierarchyLevel->pointsMap->GetData(MEMORYDEVICE_CUDA)),
					reinterpret_cast<const float*>(sceneHierarchyLevel->normalsMap->GetData(MEMORYDEVICE_CUDA)),
					sceneSize.x, sceneSize.y, sceneIntr, pose, (int32_t)iterationType, distThresh[levelId], inv,
					&f, nabla, hessian, &noValidPoints));
				return noValidPoints;
			}

		public:
			ITMDepthTracker_B200(Vector2i imgSize, TrackerIterationType *trackingRegime, int noHierarchyLevels, int noICPRunTillLevel,
				float distThresh, float terminationThreshold, const ITMLowLevelEngine *lowLevelEngine, void *cudaStream = NULL)
				: ITMDepthTracker(imgSize, trackingRegime, noHierarchyLevels, noICPRunTillLevel, distThresh, terminationThreshold,
					lowLevelEngine, MEMORYDEVICE_CUDA), tracker(NULL)
			{
				// one level: the pyramid belongs to the base class here, the handle only carries the scratch
				b200::check(dsb200_tracker_create(&tracker, imgSize.x, imgSize.y, 1, cudaStream));
			}
			~ITMDepthTracker_B200(void) { dsb200_tracker_destroy(tracker); }
		};
	}
}
