// ITMViewBuilder_B200.h -- drop-in subclass of the reference's view builder (ITMLib/Engine/ITMViewBuilder.h:20-55) whose device
// steps forward to libdsb200.  Replaces ITMViewBuilder_CUDA (ITMLib/Engine/DeviceSpecific/CUDA/ITMViewBuilder_CUDA.{h,cu}) in
// ITMLib/Engine/ITMMainEngine.cpp:57-81 and DenseSLAM/InfiniTamDriver.cpp:286.  The host side of UpdateView (allocation on first
// use, the two host-to-device SetFrom copies, the calibration switch) is the reference's own sequence (CUDA .cu:32-81).
#pragma once

#include "ITMLib/Engine/ITMViewBuilder.h"
#include "ITMSceneReconstructionEngine_B200.h"

namespace ITMLib
{
	namespace Engine
	{
		class ITMViewBuilder_B200 : public ITMViewBuilder
		{
		private:
			void *stream;

		public:
			void ConvertDisparityToDepth(ITMFloatImage *depth_out, const ITMShortImage *disp_in, const ITMIntrinsics *depthIntrinsics,
				Vector2f disparityCalibParams)
			{
				b200::check(dsb200_convert_disparity_to_depth(disp_in->GetData(MEMORYDEVICE_CUDA), depth_out->GetData(MEMORYDEVICE_CUDA),
					disp_in->noDims.x, disp_in->noDims.y, disparityCalibParams.x, disparityCalibParams.y,
					depthIntrinsics->projectionParamsSimple.fx, stream));
			}

			void ConvertDepthAffineToFloat(ITMFloatImage *depth_out, const ITMShortImage *depth_in, Vector2f depthCalibParams)
			{
				dsb200_view v; memset(&v, 0, sizeof(v));
				v.depth = depth_out->GetData(MEMORYDEVICE_CUDA); v.depth_w = depth_in->noDims.x; v.depth_h = depth_in->noDims.y;
				b200::check(dsb200_view_from_raw(depth_in->GetData(MEMORYDEVICE_CUDA), NULL, depthCalibParams.x, depthCalibParams.y, &v, stream));
			}

			void DepthFiltering(ITMFloatImage *image_out, const ITMFloatImage *image_in)
			{
				b200::check(dsb200_depth_filter_pass(image_in->GetData(MEMORYDEVICE_CUDA), image_out->GetData(MEMORYDEVICE_CUDA),
					image_in->noDims.x, image_in->noDims.y, stream));
			}

			// the sensor-noise model's depth normals and uncertainties; read only by the weighted-ICP tracker (TRACKER_WICP,
			// ITMWeightedICPTracker.cpp:67) and the uncertainty preview (ITMMainEngine.cpp:233), neither of which DenseSLAM selects
			void ComputeNormalAndWeights(ITMFloat4Image *normal_out, ITMFloatImage *sigmaZ_out, const ITMFloatImage *depth_in, Vector4f intrinsic)
			{
				const float k[4] = { intrinsic.x, intrinsic.y, intrinsic.z, intrinsic.w };
				b200::check(dsb200_compute_normal_and_weights(depth_in->GetData(MEMORYDEVICE_CUDA),
					reinterpret_cast<float*>(normal_out->GetData(MEMORYDEVICE_CUDA)), sigmaZ_out->GetData(MEMORYDEVICE_CUDA),
					depth_in->noDims.x, depth_in->noDims.y, k, stream));
			}

			// raw int16 depth + RGBA from the host: upload, convert with the view's calibration, optionally filter
			void UpdateView(ITMView **view_ptr, ITMUChar4Image *rgbImage, ITMShortImage *rawDepthImage, double timestamp, bool useBilateralFilter,
				bool modelSensorNoise = false)
			{
				const bool fresh = (*view_ptr == NULL);
				if (fresh) *view_ptr = new ITMView(calib, rgbImage->noDims, rawDepthImage->noDims, true);
				if (fresh) resizeStaging(rawDepthImage->noDims);
				if (fresh && modelSensorNoise)
				{
					(*view_ptr)->depthNormal = new ITMFloat4Image(rawDepthImage->noDims, true, true);
					(*view_ptr)->depthUncertainty = new ITMFloatImage(rawDepthImage->noDims, true, true);
				}
				ITMView *view = *view_ptr;
				view->mTimeStamp = timestamp;
				view->rgb->SetFrom(rgbImage, ORUtils::MemoryBlock<Vector4u>::CPU_TO_CUDA);
				shortImage->SetFrom(rawDepthImage, ORUtils::MemoryBlock<short>::CPU_TO_CUDA);

				const ITMDisparityCalib &dc = view->calib->disparityCalib;
				if (dc.type == ITMDisparityCalib::TRAFO_KINECT) ConvertDisparityToDepth(view->depth, shortImage, &(view->calib->intrinsics_d), dc.params);
				else if (dc.type == ITMDisparityCalib::TRAFO_AFFINE) ConvertDepthAffineToFloat(view->depth, shortImage, dc.params);

				if (useBilateralFilter)
				{
					// the reference's five DepthFiltering passes and the copy back, as one call
					dsb200_view v; memset(&v, 0, sizeof(v));
					v.depth = view->depth->GetData(MEMORYDEVICE_CUDA); v.depth_w = view->depth->noDims.x; v.depth_h = view->depth->noDims.y;
					b200::check(dsb200_view_filter_depth(&v, floatImage->GetData(MEMORYDEVICE_CUDA), stream));
				}
				if (modelSensorNoise)
					ComputeNormalAndWeights(view->depthNormal, view->depthUncertainty, view->depth, view->calib->intrinsics_d.projectionParamsSimple.all);
				b200::check(dsb200_stream_synchronize(stream));
			}

			// float depth already in metres: only the two uploads
			void UpdateView(ITMView **view_ptr, ITMUChar4Image *rgbImage, ITMFloatImage *depthImage)
			{
				if (*view_ptr == NULL) *view_ptr = new ITMView(calib, rgbImage->noDims, depthImage->noDims, true);
				(*view_ptr)->rgb->UpdateDeviceFromHost();
				(*view_ptr)->depth->UpdateDeviceFromHost();
			}

			// with an IMU sample attached to the view
			void UpdateView(ITMView **view_ptr, ITMUChar4Image *rgbImage, ITMShortImage *depthImage, double timestamp, bool useBilateralFilter,
				ITMIMUMeasurement *imuMeasurement)
			{
				if (*view_ptr == NULL)
				{
					*view_ptr = new ITMViewIMU(calib, rgbImage->noDims, depthImage->noDims, true);
					resizeStaging(depthImage->noDims);
				}
				static_cast<ITMViewIMU*>(*view_ptr)->imu->SetFrom(imuMeasurement);
				UpdateView(view_ptr, rgbImage, depthImage, timestamp, useBilateralFilter);
			}

		private:
			void resizeStaging(Vector2i size)
			{
				delete shortImage; shortImage = new ITMShortImage(size, true, true);
				delete floatImage; floatImage = new ITMFloatImage(size, true, true);
			}

		public:
			explicit ITMViewBuilder_B200(const ITMRGBDCalib *calib, void *cudaStream = NULL) : ITMViewBuilder(calib), stream(cudaStream) { }
			~ITMViewBuilder_B200(void) { }
		};
	}
}
