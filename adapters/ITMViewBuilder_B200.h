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

			void ComputeNormalAndWeights(ITMFloat4Image *, ITMFloatImage *, const ITMFloatImage *, Vector4f)
			{
				// only the weighted-ICP tracker (TRACKER_WICP, ITMLibSettings.cpp:47) asks for sensor-noise modelling; DenseSLAM never does
				throw std::runtime_error("ITMViewBuilder_B200: modelSensorNoise is not supported");
			}

			// raw int16 depth + RGBA from the host: upload, convert with the view's calibration, optionally filter
			void UpdateView(ITMView **view_ptr, ITMUChar4Image *rgbImage, ITMShortImage *rawDepthImage, double timestamp, bool useBilateralFilter,
				bool modelSensorNoise = false)
			{
				if (modelSensorNoise) throw std::runtime_error("ITMViewBuilder_B200: modelSensorNoise is not supported");
				const bool fresh = (*view_ptr == NULL);
				if (fresh) *view_ptr = new ITMView(calib, rgbImage->noDims, rawDepthImage->noDims, true);
				if (fresh) resizeStaging(rawDepthImage->noDims);
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
