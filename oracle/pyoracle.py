"""oracle/pyoracle.py -- TEST INFRASTRUCTURE ONLY.

ctypes wrappers for the two CPU checkers:
  * Oracle    -> oracle/librestate.so   (our CPU restatement, oracle/restate.cpp)
  * RefEngine -> oracle/_ref/libitmref.so (the unmodified reference CPU engines, oracle/ref_shim.cpp)
  * RefCudaEngine -> oracle/_ref/libitmref_cuda[_fast].so (the unmodified reference CUDA engines recompiled for sm_100,
                 oracle/refcuda_shim.cu): set-level pin of decay / slide window / de-integration, and the GPU comparator
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
abi = importlib.import_module("denseslam-global-consistency-h_b200.abi")

RESTATE_PATH = os.path.join(HERE, "librestate.so")
REF_PATH = os.path.join(HERE, "_ref", "libitmref.so")
REF_FAST_PATH = os.path.join(HERE, "_ref", "libitmref_fast.so")
REF_OMP_PATH = os.path.join(HERE, "_ref", "libitmref_omp.so")              # reference CPU engines, WITH_OPENMP (timing only)
REF_CUDA_PATH = os.path.join(HERE, "_ref", "libitmref_cuda.so")            # reference CUDA engines, IEEE build (set-level pin)
REF_CUDA_FAST_PATH = os.path.join(HERE, "_ref", "libitmref_cuda_fast.so")  # reference CUDA engines, --use_fast_math (comparator)


def build(force=False):
    """Compile librestate.so (always possible) and oracle/_ref (only where /root/reference exists)."""
    if force and os.path.exists(RESTATE_PATH):
        os.remove(RESTATE_PATH)
    subprocess.check_call(["make", "-C", HERE, "librestate.so"])   # make tracks restate.cpp / restate.h / dsb200.h
    subprocess.check_call(["make", "-C", HERE, "ref"])


def ptr(a):
    return a.ctypes.data_as(C.c_void_p).value


class HostScene(object):
    """Host-memory twin of ITMScene (numpy buffers behind a dsb200_scene struct)."""

    def __init__(self, params, num_blocks=abi.DEFAULT_LOCAL_BLOCKS, bucket_num=abi.DEFAULT_BUCKETS,
                 excess_size=abi.DEFAULT_EXCESS, use_swapping=False):
        self.voxels = np.zeros(num_blocks * abi.BLOCK_SIZE3, dtype=abi.VOXEL_DTYPE)
        self.hash = np.zeros(bucket_num + excess_size, dtype=abi.HASH_DTYPE)
        self.alloc_list = np.zeros(num_blocks, dtype=np.int32)
        self.excess_list = np.zeros(excess_size, dtype=np.int32)
        self.c = abi.Scene(ptr(self.voxels), ptr(self.hash), ptr(self.alloc_list), ptr(self.excess_list),
                           num_blocks, bucket_num, excess_size, num_blocks - 1, excess_size - 1, params,
                           1 if use_swapping else 0)


class HostSwapCache(object):
    """Host-memory twin of ITMGlobalCache; the block store is a dict of touched entries behind a sparse numpy buffer."""

    def __init__(self, total_entries, transfer_blocks=0x1000):
        self.states = np.zeros(total_entries, np.uint8)
        self.has_stored = np.zeros(total_entries, np.uint8)
        self.stored = np.zeros(total_entries * abi.BLOCK_SIZE3, dtype=abi.VOXEL_DTYPE)     # untouched pages are never committed
        self.c = abi.SwapCache(ptr(self.states), ptr(self.has_stored), ptr(self.stored), total_entries, transfer_blocks)

    def stored_block(self, entry):
        return self.stored[entry * abi.BLOCK_SIZE3:(entry + 1) * abi.BLOCK_SIZE3]


class HostRenderState(object):
    def __init__(self, total_entries, width, height, visible_capacity=abi.DEFAULT_LOCAL_BLOCKS):
        self.visible_ids = np.zeros(visible_capacity, dtype=np.int32)
        self.visible_type = np.zeros(total_entries, dtype=np.uint8)
        self.range_image = np.zeros((height, width, 2), dtype=np.float32)
        self.raycast_result = np.zeros((height, width, 4), dtype=np.float32)
        self.raycast_image = np.zeros((height, width, 4), dtype=np.uint8)
        self.c = abi.RenderState(ptr(self.visible_ids), ptr(self.visible_type), visible_capacity, 0,
                                 ptr(self.range_image), ptr(self.raycast_result), ptr(self.raycast_image),
                                 width, height)

    def visible(self):
        return self.visible_ids[:self.c.no_visible].copy()


def make_view(depth, rgb, intr_d, intr_rgb=None, rgb_to_depth_inv=None, timestamp=0.0, depth_ptr=None, rgb_ptr=None):
    """dsb200_view over host arrays (or explicit device pointers)."""
    h, w = depth.shape[:2] if depth is not None else (0, 0)
    v = abi.View()
    v.depth = depth_ptr if depth_ptr is not None else ptr(depth)
    v.rgb = rgb_ptr if rgb_ptr is not None else ptr(rgb)
    v.depth_w, v.depth_h, v.rgb_w, v.rgb_h = w, h, w, h
    v.intr_d = abi.f4(intr_d)
    v.intr_rgb = abi.f4(intr_rgb if intr_rgb is not None else intr_d)
    v.rgb_to_depth_inv = abi.f16(rgb_to_depth_inv if rgb_to_depth_inv is not None else abi.IDENTITY16)
    v.timestamp = timestamp
    return v


class Oracle(object):
    """librestate.so: serial CPU restatement; all pointers are host pointers."""

    def __init__(self):
        if not os.path.exists(RESTATE_PATH):
            build()
        L = C.CDLL(RESTATE_PATH)
        vp, P = C.c_void_p, C.POINTER
        L.rst_recon_create.restype = vp
        L.rst_recon_create.argtypes = [C.c_int] * 3
        L.rst_recon_destroy.argtypes = [vp]
        L.rst_set_fusion_weight_params.argtypes = [vp, P(abi.WeightParams)]
        L.rst_decayed_block_count.restype = C.c_size_t
        L.rst_decayed_block_count.argtypes = [vp]
        L.rst_history_sizes.argtypes = [vp] + [P(C.c_int)] * 5
        L.rst_reset_scene.argtypes = [vp, P(abi.Scene)]
        L.rst_allocate_from_depth.restype = C.c_int
        L.rst_allocate_from_depth.argtypes = [vp, P(abi.Scene), P(abi.View), P(abi.Tracking), P(abi.RenderState), C.c_int, C.c_int]
        L.rst_integrate.argtypes = [vp, P(abi.Scene), P(abi.View), P(abi.Tracking), P(abi.RenderState)]
        L.rst_deintegrate.argtypes = [vp, P(abi.Scene), P(abi.View), P(abi.Tracking)]
        L.rst_decay.argtypes = [vp, P(abi.Scene), P(abi.RenderState), C.c_int, C.c_int, C.c_int]
        L.rst_decay_defusion.argtypes = [vp, P(abi.Scene), P(abi.RenderState), C.c_int, C.c_int, C.c_int]
        L.rst_slide_window.argtypes = [vp, P(abi.Scene), P(abi.RenderState), C.c_int]
        L.rst_slide_window_defusion.argtypes = [vp, P(abi.Scene), P(abi.RenderState), C.c_int, C.c_int]
        fp = P(C.c_float)
        L.rst_find_visible_blocks.argtypes = [P(abi.Scene), fp, fp, P(abi.RenderState)]
        L.rst_count_visible_blocks.restype = C.c_int
        L.rst_count_visible_blocks.argtypes = [P(abi.Scene), P(abi.RenderState), C.c_int, C.c_int]
        L.rst_create_expected_depths.argtypes = [P(abi.Scene), fp, fp, P(abi.RenderState)]
        L.rst_find_surface.argtypes = [P(abi.Scene), fp, fp, P(abi.RenderState)]
        L.rst_create_icp_maps.argtypes = [P(abi.Scene), P(abi.View), P(abi.Tracking), P(abi.RenderState)]
        L.rst_render_image.argtypes = [P(abi.Scene), fp, fp, fp, P(abi.RenderState), vp, vp, C.c_int]
        L.rst_forward_render.argtypes = [P(abi.Scene), P(abi.View), P(abi.Tracking), P(abi.RenderState), vp, vp, P(C.c_int)]
        L.rst_set_shard.argtypes = [vp, P(abi.Shard)]
        L.rst_shard_owner.restype = C.c_int
        L.rst_shard_owner.argtypes = [C.c_int] * 5
        L.rst_shard_holds.restype = C.c_int
        L.rst_shard_holds.argtypes = [C.c_int] * 3 + [P(abi.Shard)]
        L.rst_raycast_sharded.argtypes = [C.c_int, P(vp), P(vp), C.c_int, fp, fp, C.c_int, C.c_int, vp, vp]
        L.rst_composite_nearest.argtypes = [C.c_int, P(vp), fp, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, vp]
        L.rst_icp_maps_from_raycast.argtypes = [P(abi.Scene), P(abi.View), P(abi.Tracking), vp, vp, C.c_int, C.c_int]
        L.rst_mesh_scene.restype = C.c_uint
        L.rst_mesh_scene.argtypes = [P(abi.Scene), vp, C.c_uint]
        L.rst_mc_tables.argtypes = [vp, vp]
        L.rst_mat4_inverse.restype = C.c_int
        L.rst_mat4_inverse.argtypes = [fp, fp]
        L.rst_mat4_multiply.argtypes = [fp, fp, fp]
        L.rst_filter_subsample_with_holes.argtypes = [vp, C.c_int, C.c_int, vp]
        L.rst_tracker_compute_g_and_h.restype = C.c_int
        L.rst_tracker_compute_g_and_h.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, C.c_int, C.c_int, vp, vp, vp, C.c_float,
                                                  C.c_int, C.c_int, vp, vp, vp, vp]
        L.rst_view_filter_depth.argtypes = [vp, C.c_int, C.c_int]
        L.rst_convert_disparity_to_depth.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float]
        L.rst_compute_normal_and_weights.argtypes = [vp, vp, vp, C.c_int, C.c_int, vp]
        L.rst_swap_mark_visible.argtypes = [P(abi.SwapCache), P(abi.RenderState)]
        L.rst_swap_integrate_global_into_local.restype = C.c_int
        L.rst_swap_integrate_global_into_local.argtypes = [P(abi.SwapCache), P(abi.Scene)]
        L.rst_swap_save_to_global_memory.restype = C.c_int
        L.rst_swap_save_to_global_memory.argtypes = [P(abi.SwapCache), P(abi.Scene), P(abi.RenderState)]
        L.rst_pose_coerce.argtypes = [vp, vp]
        L.rst_pose_from_params.argtypes = [vp, vp]
        L.rst_track_camera.argtypes = [P(abi.TrackerConfig), P(abi.View), P(abi.Tracking), vp, C.c_int, C.c_int, C.c_int,
                                       P(abi.TrackerReport)]
        self.L = L

    def swap_mark_visible(self, cache, rs):
        self.L.rst_swap_mark_visible(C.byref(cache.c), C.byref(rs.c))

    def swap_integrate_global_into_local(self, cache, scene):
        return self.L.rst_swap_integrate_global_into_local(C.byref(cache.c), C.byref(scene.c))

    def swap_save_to_global_memory(self, cache, scene, rs=None):
        return self.L.rst_swap_save_to_global_memory(C.byref(cache.c), C.byref(scene.c), C.byref(rs.c) if rs is not None else None)

    def convert_disparity_to_depth(self, disparity, param_x, param_y, fx):
        d = np.ascontiguousarray(disparity, np.int16)
        out = np.zeros(d.shape, np.float32)
        self.L.rst_convert_disparity_to_depth(ptr(d), ptr(out), d.shape[1], d.shape[0], param_x, param_y, fx)
        return out

    def compute_normal_and_weights(self, depth, intrinsics, normal4, sigma_z):
        """ComputeNormalAndWeights; normal4 (H, W, 4) and sigma_z (H, W) are updated in place (only part of them is written)"""
        d = np.ascontiguousarray(depth, np.float32)
        k = np.ascontiguousarray(intrinsics, np.float32)
        assert normal4.dtype == np.float32 and sigma_z.dtype == np.float32 and normal4.flags.c_contiguous and sigma_z.flags.c_contiguous
        self.L.rst_compute_normal_and_weights(ptr(d), ptr(normal4), ptr(sigma_z), d.shape[1], d.shape[0], ptr(k))

    def view_filter_depth(self, depth):
        """five bilateral passes (UpdateView with useBilateralFilter) -> filtered copy"""
        d = np.array(depth, np.float32).copy()
        h, w = d.shape
        self.L.rst_view_filter_depth(ptr(d), w, h)
        return d

    def pose_coerce(self, M):
        m = np.array(M, np.float32).copy()
        p = np.zeros(6, np.float32)
        self.L.rst_pose_coerce(ptr(m), ptr(p))
        return m, p

    def pose_from_params(self, params):
        p = np.ascontiguousarray(params, np.float32)
        m = np.zeros(16, np.float32)
        self.L.rst_pose_from_params(ptr(p), ptr(m))
        return m

    def track_camera(self, cfg, view, tracking, scene_pose, scene_w, scene_h, double_sums=False):
        """rst_track_camera: tracking (abi.Tracking with HOST maps) is updated in place; returns the report"""
        sp = np.ascontiguousarray(scene_pose, np.float32)
        rep = abi.TrackerReport()
        self.L.rst_track_camera(C.byref(cfg), C.byref(view), C.byref(tracking), ptr(sp), scene_w, scene_h,
                                1 if double_sums else 0, C.byref(rep))
        return rep

    def filter_subsample_with_holes(self, image):
        src = np.ascontiguousarray(image, np.float32)
        h, w = src.shape
        out = np.zeros((h // 2, w // 2), np.float32)
        self.L.rst_filter_subsample_with_holes(ptr(src), w, h, ptr(out))
        return out

    def tracker_compute_g_and_h(self, depth, view_intr, points, normals, scene_intr, scene_pose, approx_inv_pose,
                                dist_thresh, iteration_type, double_sums=False):
        """ComputeGandH restated -> (noValidPoints, f, nabla[6], hessian[36], sums[32])"""
        return _gh_call(self.L.rst_tracker_compute_g_and_h, depth, view_intr, points, normals, scene_intr, scene_pose,
                        approx_inv_pose, dist_thresh, iteration_type, 1 if double_sums else 0)

    def mesh_scene(self, scene, max_triangles):
        """ITMMeshingEngine_CPU::MeshScene restated -> (triangles[:n + 1] as TRIANGLE_DTYPE, n)"""
        buf = np.zeros(max_triangles, abi.TRIANGLE_DTYPE)
        n = int(self.L.rst_mesh_scene(C.byref(scene.c), ptr(buf), max_triangles))
        return buf[:min(n + 1, max_triangles)].copy(), n

    def mc_tables(self):
        edge, tri = np.zeros(256, np.int32), np.zeros((256, 16), np.int32)
        self.L.rst_mc_tables(ptr(edge), ptr(tri))
        return edge, tri

    def mat4_inverse(self, M):
        out = (C.c_float * 16)()
        ok = self.L.rst_mat4_inverse(abi.f16(M), out)
        assert ok
        return np.array(out[:], dtype=np.float32)

    def mat4_multiply(self, A, B):
        out = (C.c_float * 16)()
        self.L.rst_mat4_multiply(abi.f16(A), abi.f16(B), out)
        return np.array(out[:], dtype=np.float32)


def _gh_call(fn, depth, view_intr, points, normals, scene_intr, scene_pose, approx_inv_pose, dist_thresh, iteration_type,
             double_sums=None):
    d = np.ascontiguousarray(depth, np.float32)
    pts = np.ascontiguousarray(points, np.float32)
    nrm = np.ascontiguousarray(normals, np.float32)
    vh, vw = d.shape
    sh, sw = pts.shape[:2]
    vi = np.ascontiguousarray(view_intr, np.float32)
    si = np.ascontiguousarray(scene_intr, np.float32)
    sp = np.ascontiguousarray(scene_pose, np.float32)
    ap = np.ascontiguousarray(approx_inv_pose, np.float32)
    f = np.zeros(1, np.float32)
    nabla = np.zeros(6, np.float32)
    hessian = np.zeros(36, np.float32)
    sums = np.zeros(32, np.float64)
    args = [ptr(d), vw, vh, ptr(vi), ptr(pts), ptr(nrm), sw, sh, ptr(si), ptr(sp), ptr(ap), dist_thresh, iteration_type]
    if double_sums is not None:
        n = fn(*(args + [double_sums, ptr(f), ptr(nabla), ptr(hessian), ptr(sums)]))
    else:
        n = fn(*(args + [ptr(f), ptr(nabla), ptr(hessian)]))
    return int(n), float(f[0]), nabla, hessian, sums


def make_tracking(M, invM, icp_points=None, icp_normals=None, requires_full_rendering=True,
                  points_ptr=None, normals_ptr=None):
    t = abi.Tracking()
    t.M_d = abi.f16(M)
    t.invM_d = abi.f16(invM)
    t.requires_full_rendering = 1 if requires_full_rendering else 0
    t.icp_points = points_ptr if points_ptr is not None else (ptr(icp_points) if icp_points is not None else None)
    t.icp_normals = normals_ptr if normals_ptr is not None else (ptr(icp_normals) if icp_normals is not None else None)
    return t


class RefEngine(object):
    """oracle/_ref/libitmref.so: the unmodified reference CPU engines at the reference's fixed capacities."""

    def __init__(self, W, H, params, use_swapping=False, fast=False, path=None):
        path = path or (REF_FAST_PATH if fast else REF_PATH)
        if not os.path.exists(path):
            raise RuntimeError("reference build %s absent (built only where /root/reference exists)" % path)
        L = C.CDLL(path)
        vp = C.c_void_p
        L.itmref_create.restype = vp
        L.itmref_create.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int]
        for name in ("itmref_hash_table", "itmref_vba", "itmref_alloc_list", "itmref_excess_list", "itmref_visible_type",
                     "itmref_icp_points", "itmref_icp_normals", "itmref_icp_rendering", "itmref_out_image"):
            getattr(L, name).restype = vp
            getattr(L, name).argtypes = [vp]
        for name in ("itmref_visible_ids", "itmref_range_image", "itmref_raycast_result"):
            getattr(L, name).restype = vp
            getattr(L, name).argtypes = [vp, C.c_int]
        L.itmref_no_visible.argtypes = [vp, C.c_int]
        for name in ("itmref_destroy", "itmref_reset", "itmref_integrate", "itmref_create_expected_depths",
                     "itmref_create_icp_maps", "itmref_find_surface", "itmref_last_free_block", "itmref_last_free_excess"):
            getattr(L, name).argtypes = [vp]
        if hasattr(L, "itmref_update_view_raw"):
            L.itmref_update_view_raw.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_double]
            if hasattr(L, "itmref_convert_disparity_to_depth"):
                L.itmref_convert_disparity_to_depth.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float]
            if hasattr(L, "itmref_compute_normal_and_weights"):
                L.itmref_compute_normal_and_weights.argtypes = [vp, vp, vp, C.c_int, C.c_int, vp]
            if hasattr(L, "itmref_swap_states"):
                L.itmref_swap_integrate_global_into_local.argtypes = [vp]
                L.itmref_swap_save_to_global_memory.argtypes = [vp, C.c_int]
                L.itmref_swap_states.restype = vp
                L.itmref_swap_states.argtypes = [vp]
                L.itmref_swap_has_stored.argtypes = [vp, C.c_int]
                L.itmref_swap_has_stored_all.argtypes = [vp, vp]
                L.itmref_swap_stored_block.restype = vp
                L.itmref_swap_stored_block.argtypes = [vp, C.c_int]
            if hasattr(L, "itmref_update_view_raw_ex"):
                L.itmref_update_view_raw_ex.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_double, C.c_int]
            L.itmref_view_depth.restype = vp
            L.itmref_view_depth.argtypes = [vp]
        if hasattr(L, "itmref_forward_render"):
            L.itmref_forward_render.argtypes = [vp]
            L.itmref_forward_projection.restype = vp
            L.itmref_forward_projection.argtypes = [vp]
            L.itmref_no_missing_points.argtypes = [vp]
        L.itmref_set_calib.argtypes = [vp, vp, vp, vp]
        L.itmref_set_view.argtypes = [vp, vp, vp, C.c_double]
        L.itmref_set_pose.argtypes = [vp, vp]
        L.itmref_get_pose.argtypes = [vp, vp, vp]
        L.itmref_allocate.argtypes = [vp, C.c_int]
        L.itmref_render_free_view.argtypes = [vp, vp, vp, C.c_int]
        L.itmref_count_visible_blocks.argtypes = [vp, C.c_int, C.c_int]
        L.itmref_mat4_inv.argtypes = [vp, vp]
        L.itmref_mat4_mul.argtypes = [vp, vp, vp]
        if hasattr(L, "itmref_track_camera"):
            L.itmref_tracker_compute_g_and_h.restype = C.c_int
            L.itmref_tracker_compute_g_and_h.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, C.c_int, C.c_int, vp, vp, vp,
                                                         C.c_float, C.c_int, vp, vp, vp]
            L.itmref_filter_subsample_with_holes.argtypes = [vp, C.c_int, C.c_int, vp]
            L.itmref_track_camera.argtypes = [vp, C.c_int, vp, C.c_int, C.c_float, C.c_float]
            L.itmref_pose_coerce.argtypes = [vp, vp, vp]
            L.itmref_pose_from_params.argtypes = [vp, vp]
        self.L, self.W, self.H = L, W, H
        self.h = L.itmref_create(W, H, params.voxel_size, params.mu, params.max_w, params.vf_min, params.vf_max,
                                 params.stop_integrating_at_max_w, 1 if use_swapping else 0)
        self.total = L.itmref_no_total_entries()
        self.num_blocks = L.itmref_local_block_num()
        self.excess = L.itmref_excess_list_size()

    def close(self):
        if self.h:
            self.L.itmref_destroy(self.h)
            self.h = None

    def _arr(self, p, dtype, count):
        buf = (C.c_char * (np.dtype(dtype).itemsize * count)).from_address(p)
        return np.frombuffer(buf, dtype=dtype, count=count)

    def set_calib(self, intr_d, intr_rgb=None, rgb2depth=None):
        a = np.asarray(intr_d, np.float32)
        b = np.asarray(intr_rgb if intr_rgb is not None else intr_d, np.float32)
        m = np.asarray(rgb2depth if rgb2depth is not None else abi.IDENTITY16, np.float32)
        self.L.itmref_set_calib(self.h, ptr(a), ptr(b), ptr(m))

    def set_view(self, depth, rgb, timestamp=0.0):
        d = np.ascontiguousarray(depth, np.float32)
        r = np.ascontiguousarray(rgb, np.uint8)
        self.L.itmref_set_view(self.h, ptr(d), ptr(r), timestamp)

    def set_pose(self, M):
        m = np.ascontiguousarray(M, np.float32)
        self.L.itmref_set_pose(self.h, ptr(m))

    def get_pose(self):
        M = np.zeros(16, np.float32)
        iM = np.zeros(16, np.float32)
        self.L.itmref_get_pose(self.h, ptr(M), ptr(iM))
        return M, iM

    def mat4_inv(self, M):
        m = np.ascontiguousarray(M, np.float32)
        o = np.zeros(16, np.float32)
        self.L.itmref_mat4_inv(ptr(m), ptr(o))
        return o

    def mat4_mul(self, A, B):
        a = np.ascontiguousarray(A, np.float32)
        b = np.ascontiguousarray(B, np.float32)
        o = np.zeros(16, np.float32)
        self.L.itmref_mat4_mul(ptr(a), ptr(b), ptr(o))
        return o

    def mesh_scene(self, max_out):
        """ITMMeshingEngine_CPU::MeshScene on this context's scene -> (triangles[:n + 1], n)"""
        self.L.itmref_mesh_scene.restype = C.c_uint
        self.L.itmref_mesh_scene.argtypes = [C.c_void_p, C.c_void_p, C.c_uint]
        assert self.L.itmref_sizeof_triangle() == abi.TRIANGLE_DTYPE.itemsize
        buf = np.zeros(max_out, abi.TRIANGLE_DTYPE)
        n = int(self.L.itmref_mesh_scene(self.h, ptr(buf), max_out))
        return buf[:min(n + 1, max_out)].copy(), n

    def mc_tables(self):
        edge, tri = np.zeros(256, np.int32), np.zeros((256, 16), np.int32)
        self.L.itmref_mc_tables.argtypes = [C.c_void_p, C.c_void_p]
        self.L.itmref_mc_tables(ptr(edge), ptr(tri))
        return edge, tri

    def swap_integrate_global_into_local(self):
        self.L.itmref_swap_integrate_global_into_local(self.h)

    def swap_save_to_global_memory(self, with_render_state=True):
        self.L.itmref_swap_save_to_global_memory(self.h, 1 if with_render_state else 0)

    def swap_states(self):
        return self._arr(self.L.itmref_swap_states(self.h), np.uint8, self.total).copy()

    def swap_has_stored(self):
        out = np.zeros(self.total, np.uint8)
        self.L.itmref_swap_has_stored_all(self.h, ptr(out))
        return out

    def swap_stored_block(self, entry):
        return self._arr(self.L.itmref_swap_stored_block(self.h, int(entry)), abi.VOXEL_DTYPE, abi.BLOCK_SIZE3).copy()

    def allocate(self, only_visible=False):
        self.L.itmref_allocate(self.h, 1 if only_visible else 0)

    def integrate(self):
        self.L.itmref_integrate(self.h)

    def create_expected_depths(self):
        self.L.itmref_create_expected_depths(self.h)

    def create_icp_maps(self):
        self.L.itmref_create_icp_maps(self.h)

    def convert_disparity_to_depth(self, disparity, param_x, param_y, fx):
        d = np.ascontiguousarray(disparity, np.int16)
        out = np.zeros(d.shape, np.float32)
        self.L.itmref_convert_disparity_to_depth(ptr(d), ptr(out), d.shape[1], d.shape[0], param_x, param_y, fx)
        return out

    def compute_normal_and_weights(self, depth, intrinsics, normal4, sigma_z):
        d = np.ascontiguousarray(depth, np.float32)
        k = np.ascontiguousarray(intrinsics, np.float32)
        assert normal4.dtype == np.float32 and sigma_z.dtype == np.float32 and normal4.flags.c_contiguous and sigma_z.flags.c_contiguous
        self.L.itmref_compute_normal_and_weights(ptr(d), ptr(normal4), ptr(sigma_z), d.shape[1], d.shape[0], ptr(k))

    def update_view_raw(self, raw_depth_i16, rgba, scale=0.001, offset=0.0, timestamp=0.0, use_bilateral_filter=False):
        """ITMViewBuilder_CPU::UpdateView (TRAFO_AFFINE) -> copy of view->depth"""
        d = np.ascontiguousarray(raw_depth_i16, np.int16)
        r = np.ascontiguousarray(rgba, np.uint8)
        self.L.itmref_update_view_raw_ex(self.h, ptr(d), ptr(r), scale, offset, timestamp, 1 if use_bilateral_filter else 0)
        return self._arr(self.L.itmref_view_depth(self.h), np.float32, self.W * self.H).reshape(self.H, self.W).copy()

    def tracker_compute_g_and_h(self, depth, view_intr, points, normals, scene_intr, scene_pose, approx_inv_pose,
                                dist_thresh, iteration_type):
        """ITMDepthTracker_CPU::ComputeGandH on raw arrays -> (noValidPoints, f, nabla[6], hessian[36], unused)"""
        return _gh_call(self.L.itmref_tracker_compute_g_and_h, depth, view_intr, points, normals, scene_intr, scene_pose,
                        approx_inv_pose, dist_thresh, iteration_type)

    def filter_subsample_with_holes(self, image):
        src = np.ascontiguousarray(image, np.float32)
        h, w = src.shape
        out = np.zeros((h // 2, w // 2), np.float32)
        self.L.itmref_filter_subsample_with_holes(ptr(src), w, h, ptr(out))
        return out

    def pose_coerce(self, M):
        m = np.ascontiguousarray(M, np.float32)
        o = np.zeros(16, np.float32)
        p = np.zeros(6, np.float32)
        self.L.itmref_pose_coerce(ptr(m), ptr(o), ptr(p))
        return o, p

    def pose_from_params(self, params):
        p = np.ascontiguousarray(params, np.float32)
        o = np.zeros(16, np.float32)
        self.L.itmref_pose_from_params(ptr(p), ptr(o))
        return o

    def track_camera(self, regime, no_icp_run_till_level=0, dist_thresh=0.01, termination_threshold=1e-3):
        """ITMDepthTracker::TrackCamera on the current view / point cloud; pose_d is the start and receives the result"""
        r = np.ascontiguousarray(regime, np.int32)
        self.L.itmref_track_camera(self.h, len(r), ptr(r), no_icp_run_till_level, dist_thresh, termination_threshold)
        return self.get_pose()

    def forward_render(self):
        """ForwardRender at the current view/pose -> (forwardProjection copy, noFwdProjMissingPoints)"""
        self.L.itmref_forward_render(self.h)
        fwd = self._arr(self.L.itmref_forward_projection(self.h), np.float32, self.W * self.H * 4).reshape(self.H, self.W, 4).copy()
        return fwd, int(self.L.itmref_no_missing_points(self.h))

    def render_free_view(self, M, intr, render_type=0):
        m = np.ascontiguousarray(M, np.float32)
        i = np.ascontiguousarray(intr, np.float32)
        self.L.itmref_render_free_view(self.h, ptr(m), ptr(i), render_type)

    # state views (no copies)
    def hash(self):
        return self._arr(self.L.itmref_hash_table(self.h), abi.HASH_DTYPE, self.total)

    def voxels(self):
        return self._arr(self.L.itmref_vba(self.h), abi.VOXEL_DTYPE, self.num_blocks * abi.BLOCK_SIZE3)

    def alloc_list(self):
        return self._arr(self.L.itmref_alloc_list(self.h), np.int32, self.num_blocks)

    def excess_list(self):
        return self._arr(self.L.itmref_excess_list(self.h), np.int32, self.excess)

    def last_free(self):
        return self.L.itmref_last_free_block(self.h), self.L.itmref_last_free_excess(self.h)

    def visible(self, free_view=False):
        n = self.L.itmref_no_visible(self.h, int(free_view))
        return self._arr(self.L.itmref_visible_ids(self.h, int(free_view)), np.int32, max(n, 1))[:n].copy()

    def visible_type(self):
        return self._arr(self.L.itmref_visible_type(self.h), np.uint8, self.total)

    def range_image(self, free_view=False):
        return self._arr(self.L.itmref_range_image(self.h, int(free_view)), np.float32, self.W * self.H * 2).reshape(self.H, self.W, 2)

    def raycast_result(self, free_view=False):
        return self._arr(self.L.itmref_raycast_result(self.h, int(free_view)), np.float32, self.W * self.H * 4).reshape(self.H, self.W, 4)

    def icp_points(self):
        return self._arr(self.L.itmref_icp_points(self.h), np.float32, self.W * self.H * 4).reshape(self.H, self.W, 4)

    def icp_normals(self):
        return self._arr(self.L.itmref_icp_normals(self.h), np.float32, self.W * self.H * 4).reshape(self.H, self.W, 4)

    def icp_rendering(self):
        return self._arr(self.L.itmref_icp_rendering(self.h), np.uint8, self.W * self.H * 4).reshape(self.H, self.W, 4)

    def out_image(self):
        return self._arr(self.L.itmref_out_image(self.h), np.uint8, self.W * self.H * 4).reshape(self.H, self.W, 4)


class RefCudaEngine(object):
    """oracle/_ref/libitmref_cuda.so (IEEE build) or libitmref_cuda_fast.so (fast=True, the reference's own flags): the
    unmodified reference CUDA engines at the reference's fixed capacities.  Needs a GPU.  Every call is synchronous."""

    def __init__(self, W, H, params, use_swapping=False, fast=False):
        path = REF_CUDA_FAST_PATH if fast else REF_CUDA_PATH
        if not os.path.exists(path):
            raise RuntimeError("reference CUDA build %s absent (built only where /root/reference exists)" % path)
        L = C.CDLL(path)
        vp = C.c_void_p
        L.itmcu_create.restype = vp
        L.itmcu_create.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int]
        L.itmcu_last_error.restype = C.c_char_p
        L.itmcu_last_error.argtypes = [vp]
        L.itmcu_decayed_block_count.restype = C.c_longlong
        for name in ("itmcu_destroy", "itmcu_reset", "itmcu_integrate", "itmcu_deintegrate", "itmcu_create_expected_depths",
                     "itmcu_create_icp_maps", "itmcu_find_surface", "itmcu_last_free_block", "itmcu_last_free_excess",
                     "itmcu_decayed_block_count"):
            getattr(L, name).argtypes = [vp]
        L.itmcu_set_calib.argtypes = [vp, vp, vp, vp]
        L.itmcu_set_view.argtypes = [vp, vp, vp, C.c_double, C.c_int]
        L.itmcu_set_pose.argtypes = [vp, vp]
        L.itmcu_set_requires_full_rendering.argtypes = [vp, C.c_int]
        L.itmcu_set_weights.argtypes = [vp, C.c_int, C.c_int, C.c_int]
        L.itmcu_allocate.argtypes = [vp, C.c_int, C.c_int]
        L.itmcu_decay.argtypes = [vp, C.c_int, C.c_int]
        L.itmcu_decay_defusion.argtypes = [vp, C.c_int, C.c_int]
        L.itmcu_slide_window.argtypes = [vp, C.c_int]
        L.itmcu_slide_window_defusion.argtypes = [vp, C.c_int, C.c_int]
        L.itmcu_render_free_view.argtypes = [vp, vp, vp, C.c_int]
        L.itmcu_process_frame_timed.argtypes = [vp, vp]
        L.itmcu_history_sizes.argtypes = [vp, vp]
        L.itmcu_no_visible.argtypes = [vp, C.c_int]
        for name in ("itmcu_get_hash", "itmcu_get_alloc_list", "itmcu_get_visible_type", "itmcu_get_icp_points",
                     "itmcu_get_icp_normals", "itmcu_get_icp_rendering", "itmcu_get_out_image", "itmcu_get_out_depth"):
            getattr(L, name).argtypes = [vp, vp]
        for name in ("itmcu_get_visible_ids", "itmcu_get_range_image", "itmcu_get_raycast_result"):
            getattr(L, name).argtypes = [vp, C.c_int, vp]
        L.itmcu_get_blocks.argtypes = [vp, vp, C.c_int, vp]
        if L.itmcu_device_count() < 1:
            raise RuntimeError("the reference CUDA engines need a CUDA device")
        self.L, self.W, self.H, self.fast = L, W, H, bool(L.itmcu_fast_math())
        self.h = L.itmcu_create(W, H, params.voxel_size, params.mu, params.max_w, params.vf_min, params.vf_max,
                                params.stop_integrating_at_max_w, 1 if use_swapping else 0)
        self.total = L.itmcu_no_total_entries()
        self.num_blocks = L.itmcu_local_block_num()
        self.bucket_num = L.itmcu_bucket_num()
        self.excess = L.itmcu_excess_list_size()

    def close(self):
        if self.h:
            self.L.itmcu_destroy(self.h)
            self.h = None

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError("reference CUDA engine: %s" % self.L.itmcu_last_error(self.h).decode())

    def reset(self):
        self._ck(self.L.itmcu_reset(self.h))

    def set_calib(self, intr_d, intr_rgb=None, rgb2depth=None):
        a = np.asarray(intr_d, np.float32)
        b = np.asarray(intr_rgb if intr_rgb is not None else intr_d, np.float32)
        m = np.asarray(rgb2depth if rgb2depth is not None else abi.IDENTITY16, np.float32)
        self.L.itmcu_set_calib(self.h, ptr(a), ptr(b), ptr(m))

    def set_view(self, depth, rgb, timestamp=0.0):
        d = np.ascontiguousarray(depth, np.float32)
        r = np.ascontiguousarray(rgb, np.uint8)
        self._ck(self.L.itmcu_set_view(self.h, ptr(d), ptr(r), timestamp, 0))

    def set_view_device(self, depth_ptr, rgb_ptr, timestamp=0.0):
        self._ck(self.L.itmcu_set_view(self.h, depth_ptr, rgb_ptr, timestamp, 1))

    def set_pose(self, M):
        m = np.ascontiguousarray(M, np.float32)
        self.L.itmcu_set_pose(self.h, ptr(m))

    def set_weights(self, depth_weighting, max_new_w=10, max_distance=100):
        self.L.itmcu_set_weights(self.h, int(depth_weighting), max_new_w, max_distance)

    def set_requires_full_rendering(self, on):
        self.L.itmcu_set_requires_full_rendering(self.h, 1 if on else 0)

    def allocate(self, only_visible=False, is_defusion=False):
        self._ck(self.L.itmcu_allocate(self.h, 1 if only_visible else 0, 1 if is_defusion else 0))

    def integrate(self):
        self._ck(self.L.itmcu_integrate(self.h))

    def deintegrate(self):
        self._ck(self.L.itmcu_deintegrate(self.h))

    def decay(self, max_weight, min_age):
        self._ck(self.L.itmcu_decay(self.h, max_weight, min_age))

    def decay_defusion(self, max_weight, min_age):
        self._ck(self.L.itmcu_decay_defusion(self.h, max_weight, min_age))

    def slide_window(self, max_age):
        self._ck(self.L.itmcu_slide_window(self.h, max_age))

    def slide_window_defusion(self, max_age, max_size):
        self._ck(self.L.itmcu_slide_window_defusion(self.h, max_age, max_size))

    def decayed_blocks(self):
        return int(self.L.itmcu_decayed_block_count(self.h))

    def history_sizes(self):
        """(decay queue, slide-window queue, decay de-fusion queue, slide de-fusion queue, timestamp map) sizes"""
        out = np.zeros(5, np.int32)
        self.L.itmcu_history_sizes(self.h, ptr(out))
        return tuple(int(x) for x in out)

    def create_expected_depths(self):
        self._ck(self.L.itmcu_create_expected_depths(self.h))

    def create_icp_maps(self):
        self._ck(self.L.itmcu_create_icp_maps(self.h))

    def find_surface(self):
        self._ck(self.L.itmcu_find_surface(self.h))

    def render_free_view(self, M, intr, render_type=0):
        m = np.ascontiguousarray(M, np.float32)
        i = np.ascontiguousarray(intr, np.float32)
        self._ck(self.L.itmcu_render_free_view(self.h, ptr(m), ptr(i), render_type))

    def process_frame_timed(self):
        """allocate + integrate + expected depths + raycast/ICP maps -> ms of the four stages (device) + host wall ms"""
        ms = np.zeros(5, np.float32)
        self._ck(self.L.itmcu_process_frame_timed(self.h, ptr(ms)))
        return ms

    # ---- state (copies)
    def last_free(self):
        return self.L.itmcu_last_free_block(self.h), self.L.itmcu_last_free_excess(self.h)

    def hash(self):
        out = np.zeros(self.total, abi.HASH_DTYPE)
        self._ck(self.L.itmcu_get_hash(self.h, ptr(out)))
        return out

    def alloc_list(self):
        out = np.zeros(self.num_blocks, np.int32)
        self._ck(self.L.itmcu_get_alloc_list(self.h, ptr(out)))
        return out

    def visible(self, free_view=False):
        n = self.L.itmcu_no_visible(self.h, 1 if free_view else 0)
        out = np.zeros(max(n, 1), np.int32)
        self._ck(self.L.itmcu_get_visible_ids(self.h, 1 if free_view else 0, ptr(out)))
        return out[:n]

    def visible_type(self):
        out = np.zeros(self.total, np.uint8)
        self._ck(self.L.itmcu_get_visible_type(self.h, ptr(out)))
        return out

    def blocks(self, ptrs):
        p = np.ascontiguousarray(ptrs, np.int32)
        out = np.zeros((len(p), abi.BLOCK_SIZE3), abi.VOXEL_DTYPE)
        if len(p):
            self._ck(self.L.itmcu_get_blocks(self.h, ptr(p), len(p), ptr(out)))
        return out

    def _img(self, fn, shape, dtype, *pre):
        out = np.zeros(shape, dtype)
        self._ck(fn(self.h, *(list(pre) + [ptr(out)])))
        return out

    def range_image(self, free_view=False):
        return self._img(self.L.itmcu_get_range_image, (self.H, self.W, 2), np.float32, 1 if free_view else 0)

    def raycast_result(self, free_view=False):
        return self._img(self.L.itmcu_get_raycast_result, (self.H, self.W, 4), np.float32, 1 if free_view else 0)

    def icp_points(self):
        return self._img(self.L.itmcu_get_icp_points, (self.H, self.W, 4), np.float32)

    def icp_normals(self):
        return self._img(self.L.itmcu_get_icp_normals, (self.H, self.W, 4), np.float32)

    def icp_rendering(self):
        return self._img(self.L.itmcu_get_icp_rendering, (self.H, self.W, 4), np.uint8)

    def out_image(self):
        return self._img(self.L.itmcu_get_out_image, (self.H, self.W, 4), np.uint8)

    def out_depth(self):
        return self._img(self.L.itmcu_get_out_depth, (self.H, self.W), np.float32)
