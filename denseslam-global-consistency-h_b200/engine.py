"""Host-side mirror of the reference's engine call surface, on top of the C ABI (libdsb200.so).

Class / method names follow the reference so call sites read the same:
  ITMSceneReconstructionEngine  (ITMLib/Engine/ITMSceneReconstructionEngine.h:45-96)
  ITMVisualisationEngine        (ITMLib/Engine/ITMVisualisationEngine.h:71-124)
  ITMDepthTracker               (ITMLib/Engine/ITMDepthTracker.h:26-70)
  ITMScene / ITMRenderState_VH / ITMView / ITMTrackingState  (ITMLib/Objects/*.h)
All heavy state lives in device memory owned through the C ABI helpers; numpy is used only to
move data in and out.  Errors surface as abi.Dsb200Error carrying the reference's messages.
"""
import ctypes as C

import numpy as np

from . import abi


def _lib():
    return abi.load()


class DeviceBuffer(object):
    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        abi.check(_lib().dsb200_device_alloc(C.byref(p), max(self.nbytes, 4)))
        self.ptr = p.value

    def upload(self, arr, stream=None):
        a = np.ascontiguousarray(arr)
        assert a.nbytes <= self.nbytes
        abi.check(_lib().dsb200_memcpy_h2d(self.ptr, a.ctypes.data, a.nbytes, stream))
        abi.check(_lib().dsb200_stream_synchronize(stream))

    def download(self, dtype, count, stream=None):
        out = np.empty(count, dtype=dtype)
        abi.check(_lib().dsb200_memcpy_d2h(out.ctypes.data, self.ptr, out.nbytes, stream))
        abi.check(_lib().dsb200_stream_synchronize(stream))
        return out

    def free(self):
        if self.ptr:
            _lib().dsb200_device_free(self.ptr)
            self.ptr = None


def _download(ptr, dtype, count):
    out = np.empty(count, dtype=dtype)
    if count:
        abi.check(_lib().dsb200_memcpy_d2h(out.ctypes.data, ptr, out.nbytes, None))
        abi.check(_lib().dsb200_stream_synchronize(None))
    return out


class ITMScene(object):
    """Device-resident voxel-block-hash scene (Objects/ITMScene.h)."""

    def __init__(self, params, num_blocks=abi.DEFAULT_LOCAL_BLOCKS, bucket_num=abi.DEFAULT_BUCKETS,
                 excess_size=abi.DEFAULT_EXCESS, use_swapping=False):
        self.c = abi.Scene()
        abi.check(_lib().dsb200_scene_alloc(C.byref(self.c), num_blocks, bucket_num, excess_size))
        self.c.params = params
        self.c.use_swapping = 1 if use_swapping else 0

    def free(self):
        _lib().dsb200_scene_free(C.byref(self.c))

    # state downloads (tests / checkpoints)
    def hash(self):
        return _download(self.c.hash, abi.HASH_DTYPE, self.c.total_entries)

    def voxels(self, block_ptrs=None):
        if block_ptrs is None:
            return _download(self.c.voxels, abi.VOXEL_DTYPE, self.c.num_blocks * abi.BLOCK_SIZE3)
        out = np.empty((len(block_ptrs), abi.BLOCK_SIZE3), dtype=abi.VOXEL_DTYPE)
        for i, p in enumerate(block_ptrs):
            out[i] = _download(self.c.voxels + int(p) * abi.BLOCK_SIZE3 * 8, abi.VOXEL_DTYPE, abi.BLOCK_SIZE3)
        return out

    def alloc_list(self):
        return _download(self.c.alloc_list, np.int32, self.c.num_blocks)

    def excess_list(self):
        return _download(self.c.excess_list, np.int32, self.c.excess_size)


class ITMRenderState_VH(object):
    def __init__(self, total_entries, width, height, visible_capacity=abi.DEFAULT_LOCAL_BLOCKS):
        self.c = abi.RenderState()
        self.total_entries = total_entries
        abi.check(_lib().dsb200_render_state_alloc(C.byref(self.c), total_entries, visible_capacity, width, height))

    def free(self):
        _lib().dsb200_render_state_free(C.byref(self.c))

    def visible(self):
        return _download(self.c.visible_ids, np.int32, min(self.c.no_visible, self.c.visible_capacity))

    def visible_type(self):
        return _download(self.c.visible_type, np.uint8, self.total_entries)

    def range_image(self):
        return _download(self.c.range_image, np.float32, self.c.width * self.c.height * 2).reshape(self.c.height, self.c.width, 2)

    def raycast_result(self):
        return _download(self.c.raycast_result, np.float32, self.c.width * self.c.height * 4).reshape(self.c.height, self.c.width, 4)

    def raycast_image(self):
        return _download(self.c.raycast_image, np.uint8, self.c.width * self.c.height * 4).reshape(self.c.height, self.c.width, 4)


class ITMView(object):
    """Depth + rgb on the device with calibration (Objects/ITMView.h)."""

    def __init__(self, width, height, intr_d, intr_rgb=None, rgb_to_depth_inv=None):
        self.W, self.H = width, height
        self.depth = DeviceBuffer(width * height * 4)
        self.rgb = DeviceBuffer(width * height * 4)
        self.c = abi.View()
        self.c.depth, self.c.rgb = self.depth.ptr, self.rgb.ptr
        self.c.depth_w, self.c.depth_h, self.c.rgb_w, self.c.rgb_h = width, height, width, height
        self.c.intr_d = abi.f4(intr_d)
        self.c.intr_rgb = abi.f4(intr_rgb if intr_rgb is not None else intr_d)
        self.c.rgb_to_depth_inv = abi.f16(rgb_to_depth_inv if rgb_to_depth_inv is not None else abi.IDENTITY16)

    def update(self, depth, rgb, timestamp=0.0):
        self.depth.upload(np.asarray(depth, np.float32))
        self.rgb.upload(np.asarray(rgb, np.uint8))
        self.c.timestamp = timestamp

    def update_from_raw(self, depth_i16, bgr, timestamp=0.0, scale=0.001, offset=0.0, stream=None, useBilateralFilter=False):
        """ITMViewBuilder::UpdateView (affine calibration) + the driver's BGR -> RGBA conversion, on the device"""
        d = np.ascontiguousarray(depth_i16, np.int16)
        if not hasattr(self, "raw_depth"):
            self.raw_depth, self.raw_bgr = DeviceBuffer(self.W * self.H * 2), DeviceBuffer(self.W * self.H * 3)
        self.raw_depth.upload(d, stream)
        if bgr is not None:
            self.raw_bgr.upload(np.ascontiguousarray(bgr, np.uint8), stream)
        abi.check(_lib().dsb200_view_from_raw(self.raw_depth.ptr, self.raw_bgr.ptr if bgr is not None else None, scale, offset,
                                              C.byref(self.c), stream))
        if useBilateralFilter:
            self.filter_depth(stream)
        abi.check(_lib().dsb200_stream_synchronize(stream))
        self.c.timestamp = timestamp

    def filter_depth(self, stream=None):
        """UpdateView's useBilateralFilter branch: five 5x5 bilateral passes over the depth image, in place"""
        if not hasattr(self, "filter_scratch"):
            self.filter_scratch = DeviceBuffer(self.W * self.H * 4)
        abi.check(_lib().dsb200_view_filter_depth(C.byref(self.c), self.filter_scratch.ptr, stream))
        abi.check(_lib().dsb200_stream_synchronize(stream))

    def free(self):
        self.depth.free()
        self.rgb.free()
        if hasattr(self, "raw_depth"):
            self.raw_depth.free(); self.raw_bgr.free()
        if hasattr(self, "filter_scratch"):
            self.filter_scratch.free()


class ITMTrackingState(object):
    def __init__(self, width, height):
        self.W, self.H = width, height
        self.points = DeviceBuffer(width * height * 16)
        self.normals = DeviceBuffer(width * height * 16)
        self.c = abi.Tracking()
        self.c.icp_points, self.c.icp_normals = self.points.ptr, self.normals.ptr
        self.c.requires_full_rendering = 1
        self.set_pose(abi.IDENTITY16)
        # pose_pointCloud->GetM(): pose the ICP maps were raycast from (set by CreateICPMaps, ITMVisualisationEngine_CPU.cpp:430)
        self.pose_pointCloud = np.array(abi.IDENTITY16, np.float32)

    def pose(self):
        """(M_d, invM_d) as float32[16] column-major"""
        return np.array(self.c.M_d[:], np.float32), np.array(self.c.invM_d[:], np.float32)

    def set_pose(self, M, invM=None):
        """pose_d->SetM: M is world->camera, column-major. invM defaults to the cofactor inverse."""
        self.c.M_d = abi.f16(M)
        if invM is None:
            out = (C.c_float * 16)()
            if not _lib().dsb200_mat4_inverse(self.c.M_d, out):
                raise ValueError("singular pose matrix")
            self.c.invM_d = out
        else:
            self.c.invM_d = abi.f16(invM)

    def icp_points(self):
        return self.points.download(np.float32, self.W * self.H * 4).reshape(self.H, self.W, 4)

    def icp_normals(self):
        return self.normals.download(np.float32, self.W * self.H * 4).reshape(self.H, self.W, 4)

    def free(self):
        self.points.free()
        self.normals.free()


class ITMSceneReconstructionEngine(object):
    def __init__(self, num_blocks=abi.DEFAULT_LOCAL_BLOCKS, bucket_num=abi.DEFAULT_BUCKETS, excess_size=abi.DEFAULT_EXCESS,
                 stream=None):
        self.h = C.c_void_p()
        abi.check(_lib().dsb200_recon_create(C.byref(self.h), num_blocks, bucket_num, excess_size, stream))

    def close(self):
        if self.h:
            _lib().dsb200_recon_destroy(self.h)
            self.h = None

    def SetAsync(self, flag):
        abi.check(_lib().dsb200_recon_set_async(self.h, 1 if flag else 0))

    def Sync(self, scene, renderState):
        abi.check(_lib().dsb200_recon_sync(self.h, C.byref(scene.c), C.byref(renderState.c)))

    def ResetScene(self, scene):
        abi.check(_lib().dsb200_reset_scene(self.h, C.byref(scene.c)))

    def AllocateSceneFromDepth(self, scene, view, trackingState, renderState, onlyUpdateVisibleList=False, isDefusion=False):
        abi.check(_lib().dsb200_allocate_from_depth(self.h, C.byref(scene.c), C.byref(view.c), C.byref(trackingState.c),
                                                    C.byref(renderState.c), int(onlyUpdateVisibleList), int(isDefusion)))

    def IntegrateIntoScene(self, scene, view, trackingState, renderState):
        abi.check(_lib().dsb200_integrate(self.h, C.byref(scene.c), C.byref(view.c), C.byref(trackingState.c), C.byref(renderState.c)))

    def DeIntegrateIntoScene(self, scene, view, trackingState):
        abi.check(_lib().dsb200_deintegrate(self.h, C.byref(scene.c), C.byref(view.c), C.byref(trackingState.c)))

    def Decay(self, scene, renderState, maxWeight, minAge, forceAllVoxels=False):
        abi.check(_lib().dsb200_decay(self.h, C.byref(scene.c), C.byref(renderState.c), maxWeight, minAge, int(forceAllVoxels)))

    def DecayDefusionPart(self, scene, renderState, maxWeight, minAge, forceAllVoxels=False):
        abi.check(_lib().dsb200_decay_defusion(self.h, C.byref(scene.c), C.byref(renderState.c), maxWeight, minAge, int(forceAllVoxels)))

    def SlideWindow(self, scene, renderState, maxAge):
        abi.check(_lib().dsb200_slide_window(self.h, C.byref(scene.c), C.byref(renderState.c), maxAge))

    def SlideWindowDefusionPart(self, scene, renderState, maxAge, maxSize):
        abi.check(_lib().dsb200_slide_window_defusion(self.h, C.byref(scene.c), C.byref(renderState.c), maxAge, maxSize))

    def GetDecayedBlockCount(self):
        return int(_lib().dsb200_decayed_block_count(self.h))

    def SetFusionWeightParams(self, depthWeighting, maxNewW=10, maxDistance=100):
        wp = abi.WeightParams(int(depthWeighting), maxNewW, maxDistance)
        abi.check(_lib().dsb200_set_fusion_weight_params(self.h, C.byref(wp)))

    def SetShard(self, rank, count, brick_log2=2, halo=True):
        """sharded scene (include/dsb200.h, dsb200_shard): this engine allocates only the blocks shard `rank` owns"""
        sh = abi.Shard(rank, count, brick_log2, 1 if halo else 0)
        abi.check(_lib().dsb200_recon_set_shard(self.h, C.byref(sh)))

    def HistorySizes(self):
        v = [C.c_int() for _ in range(5)]
        _lib().dsb200_history_sizes(self.h, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def SaveCheckpoint(self, scene, renderState, directory):
        """volume checkpoint: scene arrays (+ visible list / flags) in MemoryBlockPersister's file layout"""
        abi.check(_lib().dsb200_recon_abandon_prefetch(self.h))      # a pending look-ahead's entries never go into a checkpoint
        abi.check(_lib().dsb200_scene_save(C.byref(scene.c), C.byref(renderState.c) if renderState is not None else None,
                                           str(directory).encode()))

    def LoadCheckpoint(self, scene, renderState, directory):
        abi.check(_lib().dsb200_scene_load(self.h, C.byref(scene.c), C.byref(renderState.c) if renderState is not None else None,
                                           str(directory).encode()))

    def PrefetchAllocation(self, scene, nextView, nextTrackingState, renderState, waitEvent=None):
        """Look-ahead (asynchronous mode): the request pass of the NEXT frame's allocation, overlapping this frame's raycast.
        nextView / nextTrackingState may be engine objects or raw abi.View / abi.Tracking structs."""
        v = nextView.c if hasattr(nextView, "c") else nextView
        t = nextTrackingState.c if hasattr(nextTrackingState, "c") else nextTrackingState
        abi.check(_lib().dsb200_recon_prefetch_allocation(self.h, C.byref(scene.c), C.byref(v), C.byref(t), C.byref(renderState.c), waitEvent))

    def ProcessFrame(self, vis, scene, view, trackingState, renderState, stages=abi.FRAME_ALL):
        """ITMDenseMapper::ProcessFrame + ITMTrackingController::Prepare as one submission."""
        abi.check(_lib().dsb200_process_frame(self.h, vis.h if vis is not None else None, C.byref(scene.c), C.byref(view.c),
                                              C.byref(trackingState.c), C.byref(renderState.c), stages))
        if stages & abi.FRAME_ICP_MAPS:
            trackingState.pose_pointCloud = np.array(trackingState.c.M_d[:], np.float32)


class ITMVisualisationEngine(object):
    def __init__(self, width, height, max_visible=abi.DEFAULT_LOCAL_BLOCKS, stream=None):
        self.h = C.c_void_p()
        abi.check(_lib().dsb200_vis_create(C.byref(self.h), max_visible, width, height, stream))

    def close(self):
        if self.h:
            _lib().dsb200_vis_destroy(self.h)
            self.h = None

    def CreateRenderState(self, scene, imgSize, visible_capacity=None):
        return ITMRenderState_VH(scene.c.total_entries, imgSize[0], imgSize[1],
                                 visible_capacity if visible_capacity else scene.c.num_blocks)

    def FindVisibleBlocks(self, scene, M, intrinsics, renderState):
        abi.check(_lib().dsb200_find_visible_blocks(self.h, C.byref(scene.c), abi.f16(M), abi.f4(intrinsics), C.byref(renderState.c)))

    def CountVisibleBlocks(self, scene, renderState, minBlockId, maxBlockId):
        n = C.c_int()
        abi.check(_lib().dsb200_count_visible_blocks(self.h, C.byref(scene.c), C.byref(renderState.c), minBlockId, maxBlockId, C.byref(n)))
        return n.value

    def CreateExpectedDepths(self, scene, M, intrinsics, renderState):
        abi.check(_lib().dsb200_create_expected_depths(self.h, C.byref(scene.c), abi.f16(M), abi.f4(intrinsics), C.byref(renderState.c)))

    def FindSurface(self, scene, invM, intrinsics, renderState):
        abi.check(_lib().dsb200_find_surface(self.h, C.byref(scene.c), abi.f16(invM), abi.f4(intrinsics), C.byref(renderState.c)))

    def CreateICPMaps(self, scene, view, trackingState, renderState):
        abi.check(_lib().dsb200_create_icp_maps(self.h, C.byref(scene.c), C.byref(view.c), C.byref(trackingState.c), C.byref(renderState.c)))
        trackingState.pose_pointCloud = np.array(trackingState.c.M_d[:], np.float32)

    def RenderImage(self, scene, M, invM, intrinsics, renderState, outImage, outFloat, renderType=abi.RENDER_SHADED_GREYSCALE):
        abi.check(_lib().dsb200_render_image(self.h, C.byref(scene.c), abi.f16(M), abi.f16(invM), abi.f4(intrinsics),
                                             C.byref(renderState.c), outImage.ptr, outFloat.ptr if outFloat is not None else None,
                                             renderType))

    def ForwardRender(self, scene, view, trackingState, renderState):
        abi.check(_lib().dsb200_forward_render(self.h, C.byref(scene.c), C.byref(view.c), C.byref(trackingState.c), C.byref(renderState.c)))

    def ForwardProjection(self, width, height):
        """(forwardProjection image, noFwdProjMissingPoints) of the last ForwardRender"""
        p, n = C.c_void_p(), C.c_int()
        abi.check(_lib().dsb200_forward_projection(self.h, C.byref(p), C.byref(n)))
        return _download(p.value, np.float32, width * height * 4).reshape(height, width, 4), n.value

    def CreatePointCloud(self):
        abi.check(_lib().dsb200_create_point_cloud(self.h))


class ITMDepthTracker(object):
    """ICP depth tracker (ITMLib/Engine/ITMDepthTracker.h:26-70): constructor arguments and TrackCamera as in the
    reference; ComputeGandH / FilterSubsampleWithHoles run on the GPU, the Levenberg-Marquardt loop on the host
    (csrc/tracker_host.cpp).  Defaults are ITMLibSettings.cpp:13-16,62-73."""

    DEFAULT_REGIME = (abi.TRACKER_ITERATION_BOTH, abi.TRACKER_ITERATION_BOTH, abi.TRACKER_ITERATION_ROTATION,
                      abi.TRACKER_ITERATION_ROTATION, abi.TRACKER_ITERATION_ROTATION)

    def __init__(self, imgSize, trackingRegime=DEFAULT_REGIME, noHierarchyLevels=None, noICPRunTillLevel=0,
                 distThresh=0.1 * 0.1, terminationThreshold=1e-3, stream=None):
        levels = len(trackingRegime) if noHierarchyLevels is None else noHierarchyLevels
        if not 1 <= levels <= abi.TRACKER_MAX_LEVELS or levels > len(trackingRegime):
            raise ValueError("noHierarchyLevels must be 1..%d and covered by trackingRegime" % abi.TRACKER_MAX_LEVELS)
        self.W, self.H = int(imgSize[0]), int(imgSize[1])
        self.cfg = abi.TrackerConfig()
        self.cfg.levels = levels
        for i in range(levels):
            self.cfg.regime[i] = int(trackingRegime[i])
        self.cfg.no_icp_run_till_level = noICPRunTillLevel
        self.cfg.dist_thresh = distThresh
        self.cfg.termination_threshold = terminationThreshold
        self.report = abi.TrackerReport()
        self.h = C.c_void_p()
        abi.check(_lib().dsb200_tracker_create(C.byref(self.h), self.W, self.H, levels, stream))

    def close(self):
        if self.h:
            _lib().dsb200_tracker_destroy(self.h)
            self.h = None

    def TrackCamera(self, trackingState, view):
        """Refines trackingState's pose_d against its point cloud (the last CreateICPMaps) using view.depth."""
        abi.check(_lib().dsb200_tracker_track_camera(self.h, C.byref(self.cfg), C.byref(view.c), C.byref(trackingState.c),
                                                     abi.f16(trackingState.pose_pointCloud), trackingState.W, trackingState.H,
                                                     C.byref(self.report)))

    # ---- the device half on its own (what a C++ subclass of the reference's ITMDepthTracker forwards to) ----
    def SetEvaluationData(self, trackingState, view):
        abi.check(_lib().dsb200_tracker_set_evaluation_data(self.h, view.c.depth, view.c.intr_d, trackingState.c.icp_points,
                                                            trackingState.c.icp_normals, trackingState.W, trackingState.H,
                                                            view.c.intr_d, abi.f16(trackingState.pose_pointCloud)))

    def ComputeGandH(self, levelId, iterationType, distThresh, approxInvPose):
        """-> (noValidPoints, f, nabla[6], hessian[36], sums[32] float64)"""
        f, n = C.c_float(), C.c_int32()
        nabla, hessian, sums = (C.c_float * 6)(), (C.c_float * 36)(), (C.c_double * 32)()
        abi.check(_lib().dsb200_tracker_compute_g_and_h(self.h, levelId, iterationType, distThresh, abi.f16(approxInvPose),
                                                        C.byref(f), nabla, hessian, C.byref(n)))
        abi.check(_lib().dsb200_tracker_last_sums(self.h, sums))
        return n.value, f.value, np.array(nabla[:], np.float32), np.array(hessian[:], np.float32), np.array(sums[:], np.float64)

    def ComputeGandHOn(self, view, trackingState, iterationType, distThresh, approxInvPose):
        """the same evaluation on caller-owned images (dsb200_tracker_compute_g_and_h_on), here the view's full-size depth"""
        f, n = C.c_float(), C.c_int32()
        nabla, hessian, sums = (C.c_float * 6)(), (C.c_float * 36)(), (C.c_double * 32)()
        abi.check(_lib().dsb200_tracker_compute_g_and_h_on(
            self.h, view.c.depth, view.c.depth_w, view.c.depth_h, view.c.intr_d, trackingState.c.icp_points,
            trackingState.c.icp_normals, trackingState.W, trackingState.H, view.c.intr_d, abi.f16(trackingState.pose_pointCloud),
            iterationType, distThresh, abi.f16(approxInvPose), C.byref(f), nabla, hessian, C.byref(n)))
        abi.check(_lib().dsb200_tracker_last_sums(self.h, sums))
        return n.value, f.value, np.array(nabla[:], np.float32), np.array(hessian[:], np.float32), np.array(sums[:], np.float64)

    def Level(self, levelId):
        """(depth image copy, intrinsics) of a pyramid level after SetEvaluationData / TrackCamera"""
        p, w, h = C.c_void_p(), C.c_int32(), C.c_int32()
        intr = (C.c_float * 4)()
        abi.check(_lib().dsb200_tracker_level(self.h, levelId, C.byref(p), C.byref(w), C.byref(h), intr))
        return _download(p.value, np.float32, w.value * h.value).reshape(h.value, w.value), np.array(intr[:], np.float32)


class ITMSwappingEngine(object):
    """ITMSwappingEngine + the scene's ITMGlobalCache (Engine/ITMSwappingEngine.h:16-31, Objects/ITMGlobalCache.h): swap states on
    the device, block store in host memory.  `recon` (optional) is told after every call that changes the volume."""

    def __init__(self, scene, transfer_blocks=0x1000, stream=None, recon=None):
        self.h = C.c_void_p()
        abi.check(_lib().dsb200_swap_create(C.byref(self.h), scene.c.total_entries, transfer_blocks, stream))
        self.c = abi.SwapCache()
        abi.check(_lib().dsb200_swap_cache_of(self.h, C.byref(self.c)))
        self.recon = recon
        if recon is not None:      # the library then synchronises the engine before and notifies it after every swap call
            abi.check(_lib().dsb200_swap_attach_recon(self.h, recon.h))

    def close(self):
        if self.h:
            _lib().dsb200_swap_destroy(self.h)
            self.h = None

    def _changed(self):
        pass        # since dsb200_swap_attach_recon the library notifies the engine itself

    def MarkVisible(self, renderState):
        """the swap-state pass the reference's AllocateSceneFromDepth runs when useSwapping is on"""
        abi.check(_lib().dsb200_swap_mark_visible(self.h, C.byref(renderState.c)))

    def IntegrateGlobalIntoLocal(self, scene, renderState=None):
        """with a render state (the reference's signature) the allocator's swap-state pass runs first, as in the C++ adapter"""
        if renderState is not None:
            self.MarkVisible(renderState)
        n = C.c_int32()
        abi.check(_lib().dsb200_swap_integrate_global_into_local(self.h, C.byref(scene.c), C.byref(n)))
        self._changed()
        return n.value

    def SaveToGlobalMemory(self, scene, renderState=None):
        n = C.c_int32()
        abi.check(_lib().dsb200_swap_save_to_global_memory(self.h, C.byref(scene.c), C.byref(renderState.c) if renderState is not None else None,
                                                           C.byref(n)))
        self._changed()
        return n.value

    def SetStoreBudget(self, max_blocks):
        """bound on the pinned host store, in voxel blocks (0 = none); no counterpart in the reference (its store is pageable)"""
        abi.check(_lib().dsb200_swap_set_store_budget(self.h, int(max_blocks)))

    def SaveToFile(self, path):
        abi.check(_lib().dsb200_swap_save_file(self.h, path.encode()))

    def ReadFromFile(self, path):
        abi.check(_lib().dsb200_swap_load_file(self.h, path.encode()))

    # state (tests)
    def states(self):
        return _download(self.c.states, np.uint8, self.c.total_entries)

    def has_stored(self):
        return np.frombuffer((C.c_char * self.c.total_entries).from_address(self.c.has_stored), np.uint8).copy()

    def stored_block(self, entry):
        p = _lib().dsb200_swap_stored_block(self.h, int(entry))
        if not p:
            return None
        return np.frombuffer((C.c_char * (8 * abi.BLOCK_SIZE3)).from_address(p), abi.VOXEL_DTYPE).copy()


def convert_disparity_to_depth(disparity_i16, param_x, param_y, fx_depth):
    """ITMViewBuilder::ConvertDisparityToDepth on the device (host array in, host array out)"""
    d = np.ascontiguousarray(disparity_i16, np.int16)
    h, w = d.shape
    src, dst = DeviceBuffer(d.nbytes), DeviceBuffer(d.size * 4)
    src.upload(d)
    abi.check(_lib().dsb200_convert_disparity_to_depth(src.ptr, dst.ptr, w, h, param_x, param_y, fx_depth, None))
    out = dst.download(np.float32, d.size).reshape(h, w)
    src.free(); dst.free()
    return out


def depth_filter_pass(depth):
    """ITMViewBuilder::DepthFiltering, one pass (host array in, host array out)"""
    d = np.ascontiguousarray(depth, np.float32)
    h, w = d.shape
    src, dst = DeviceBuffer(d.nbytes), DeviceBuffer(d.nbytes)
    src.upload(d)
    abi.check(_lib().dsb200_depth_filter_pass(src.ptr, dst.ptr, w, h, None))
    out = dst.download(np.float32, d.size).reshape(h, w)
    src.free(); dst.free()
    return out


def compute_normal_and_weights(depth, intrinsics, normal4, sigma_z):
    """ITMViewBuilder::ComputeNormalAndWeights on the device (host arrays; normal4 (H, W, 4) and sigma_z (H, W) hold what the
    images held before the call and come back updated, since the reference leaves part of them untouched)"""
    d = np.ascontiguousarray(depth, np.float32)
    h, w = d.shape
    n4 = np.ascontiguousarray(normal4, np.float32); sz = np.ascontiguousarray(sigma_z, np.float32)
    assert n4.shape == (h, w, 4) and sz.shape == (h, w)
    src, dn, ds = DeviceBuffer(d.nbytes), DeviceBuffer(n4.nbytes), DeviceBuffer(sz.nbytes)
    src.upload(d); dn.upload(n4); ds.upload(sz)
    k = (C.c_float * 4)(*[float(np.float32(v)) for v in intrinsics])
    abi.check(_lib().dsb200_compute_normal_and_weights(src.ptr, dn.ptr, ds.ptr, w, h, k, None))
    out_n = dn.download(np.float32, n4.size).reshape(h, w, 4)
    out_s = ds.download(np.float32, sz.size).reshape(h, w)
    src.free(); dn.free(); ds.free()
    return out_n, out_s


def pose_coerce(M):
    """ITMPose::SetM + Coerce -> (M float32[16], params (tx, ty, tz, rx, ry, rz))"""
    m, p = abi.f16(M), (C.c_float * 6)()
    abi.check(_lib().dsb200_pose_coerce(m, p))
    return np.array(m[:], np.float32), np.array(p[:], np.float32)


def pose_from_params(params):
    m = (C.c_float * 16)()
    abi.check(_lib().dsb200_pose_from_params((C.c_float * 6)(*[float(x) for x in params]), m))
    return np.array(m[:], np.float32)


class ITMMeshingEngine(object):
    """ITMMeshingEngine<ITMVoxel,ITMVoxelBlockHash> (ITMLib/Engine/ITMMeshingEngine.h:19-26): marching cubes over the allocated
    blocks, triangles in the CPU engine's order."""

    def __init__(self, num_blocks=abi.DEFAULT_LOCAL_BLOCKS, stream=None):
        self.h = C.c_void_p()
        abi.check(_lib().dsb200_mesher_create(C.byref(self.h), num_blocks, stream))

    def close(self):
        if self.h:
            _lib().dsb200_mesher_destroy(self.h)
            self.h = None

    def MeshScene(self, scene, max_triangles):
        """-> (triangles[:n + 1] as abi.TRIANGLE_DTYPE (the slot past the count holds the overflow triangle when capped), n)"""
        buf = DeviceBuffer(max_triangles * abi.TRIANGLE_DTYPE.itemsize)
        n = C.c_uint32()
        try:
            abi.check(_lib().dsb200_mesh_scene(self.h, C.byref(scene.c), buf.ptr, max_triangles, C.byref(n)))
            out = buf.download(abi.TRIANGLE_DTYPE, min(n.value + 1, max_triangles))
        finally:
            buf.free()
        return out, int(n.value)
