"""ctypes mirror of include/dsb200.h (POD structs + library loader).

The product library is denseslam-global-consistency-h_b200/csrc/libdsb200.so (built by
__graft_entry__.build()).  Loading fails loudly if it is missing: there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libdsb200.so")

BLOCK_SIZE = 8
BLOCK_SIZE3 = 512
DEFAULT_LOCAL_BLOCKS = 0x40000
DEFAULT_BUCKETS = 0x100000
DEFAULT_EXCESS = 0x60000

OK = 0
ERR_CUDA, ERR_INVALID, ERR_VBA_EXHAUSTED, ERR_EXCESS_EXHAUSTED, ERR_UNSUPPORTED, ERR_NO_DEVICE = -1, -2, -3, -4, -5, -6

RENDER_SHADED_GREYSCALE, RENDER_COLOUR_FROM_VOLUME, RENDER_COLOUR_FROM_NORMAL, RENDER_DEPTH_MAP, RENDER_COLOURCODED = range(5)
PROFILE_STAGES = 5
FRAME_ALLOCATE, FRAME_INTEGRATE, FRAME_EXPECTED_DEPTHS, FRAME_ICP_MAPS, FRAME_ALL = 1, 2, 4, 8, 15
FRAME_RAYCAST = 16
# TrackerIterationType, ITMLib/Utils/ITMLibDefines.h:289-295
TRACKER_ITERATION_ROTATION, TRACKER_ITERATION_TRANSLATION, TRACKER_ITERATION_BOTH, TRACKER_ITERATION_NONE = 1, 2, 3, 4
TRACKER_STATIC_MAX_PIXELS = 1 << 22          # dsb200_tracker_compute_g_and_h_static: largest view image
TRACKER_MAX_LEVELS = 8
IPC_HANDLE_BYTES = 64
MAX_SHARDS = 16

# numpy views of the two on-device record types (ITMHashEntry 20 B, ITMVoxel_s_rgb 8 B)
HASH_DTYPE = np.dtype([("pos", "<i2", (3,)), ("pad", "<i2"), ("offset", "<i4"), ("ptr", "<i4"),
                       ("allocated_time", "<i4")])
VOXEL_DTYPE = np.dtype([("sdf", "<i2"), ("w_depth", "u1"), ("clr", "u1", (3,)), ("w_color", "u1"), ("pad", "u1")])
TRIANGLE_DTYPE = np.dtype([("p0", "<f4", (3,)), ("p1", "<f4", (3,)), ("p2", "<f4", (3,)), ("c0", "<f4", (3,)), ("c1", "<f4", (3,)),
                           ("c2", "<f4", (3,))])       # ITMMesh::Triangle (Objects/ITMMesh.h:21-24)
assert HASH_DTYPE.itemsize == 20 and VOXEL_DTYPE.itemsize == 8 and TRIANGLE_DTYPE.itemsize == 72


class SceneParams(C.Structure):
    _fields_ = [("voxel_size", C.c_float), ("mu", C.c_float), ("max_w", C.c_int32), ("vf_min", C.c_float),
                ("vf_max", C.c_float), ("stop_integrating_at_max_w", C.c_int32)]


class Scene(C.Structure):
    _fields_ = [("voxels", C.c_void_p), ("hash", C.c_void_p), ("alloc_list", C.c_void_p), ("excess_list", C.c_void_p),
                ("num_blocks", C.c_int32), ("bucket_num", C.c_int32), ("excess_size", C.c_int32),
                ("last_free_block", C.c_int32), ("last_free_excess", C.c_int32), ("params", SceneParams),
                ("use_swapping", C.c_int32)]

    @property
    def total_entries(self):
        return self.bucket_num + self.excess_size


class SwapCache(C.Structure):
    """dsb200_swap_cache (ITMGlobalCache): swap states + the host-side block store"""
    _fields_ = [("states", C.c_void_p), ("has_stored", C.c_void_p), ("stored", C.c_void_p),
                ("total_entries", C.c_int32), ("transfer_blocks", C.c_int32)]


class RenderState(C.Structure):
    _fields_ = [("visible_ids", C.c_void_p), ("visible_type", C.c_void_p), ("visible_capacity", C.c_int32),
                ("no_visible", C.c_int32), ("range_image", C.c_void_p), ("raycast_result", C.c_void_p),
                ("raycast_image", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32)]


class View(C.Structure):
    _fields_ = [("depth", C.c_void_p), ("rgb", C.c_void_p), ("depth_w", C.c_int32), ("depth_h", C.c_int32),
                ("rgb_w", C.c_int32), ("rgb_h", C.c_int32), ("intr_d", C.c_float * 4), ("intr_rgb", C.c_float * 4),
                ("rgb_to_depth_inv", C.c_float * 16), ("timestamp", C.c_double)]


class Tracking(C.Structure):
    _fields_ = [("M_d", C.c_float * 16), ("invM_d", C.c_float * 16), ("requires_full_rendering", C.c_int32),
                ("icp_points", C.c_void_p), ("icp_normals", C.c_void_p)]


class WeightParams(C.Structure):
    _fields_ = [("depth_weighting", C.c_int32), ("max_new_w", C.c_int32), ("max_distance", C.c_int32)]


class Shard(C.Structure):
    _fields_ = [("rank", C.c_int32), ("count", C.c_int32), ("brick_log2", C.c_int32), ("halo", C.c_int32)]


class ShardTables(C.Structure):
    _fields_ = [("count", C.c_int32), ("rank", C.c_int32), ("brick_log2", C.c_int32), ("pad_", C.c_int32),
                ("hash", C.c_void_p * 16), ("voxels", C.c_void_p * 16), ("range_image", C.c_void_p * 16)]


class TrackerConfig(C.Structure):
    _fields_ = [("levels", C.c_int32), ("regime", C.c_int32 * 8), ("no_icp_run_till_level", C.c_int32),
                ("dist_thresh", C.c_float), ("termination_threshold", C.c_float)]


class TrackerReport(C.Structure):
    _fields_ = [("evaluations", C.c_int32), ("steps", C.c_int32), ("last_valid_points", C.c_int32), ("last_f", C.c_float)]


def f16(a):
    """float[16] column-major ctypes array from any 16-element sequence."""
    return (C.c_float * 16)(*[float(x) for x in np.asarray(a, dtype=np.float32).reshape(-1)])


def f4(a):
    return (C.c_float * 4)(*[float(x) for x in np.asarray(a, dtype=np.float32).reshape(-1)])


IDENTITY16 = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]

_P = C.POINTER
_lib = None


def load():
    """Load libdsb200.so and declare every prototype of include/dsb200.h.  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libdsb200.so not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback for the product path." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, cp = C.c_void_p, C.c_int32, C.c_char_p
    sig = {
        "dsb200_last_error": (cp, []),
        "dsb200_device_count": (C.c_int, []),
        "dsb200_set_device": (C.c_int, [C.c_int]),
        "dsb200_recon_create": (C.c_int, [_P(vp), i32, i32, i32, vp]),
        "dsb200_recon_destroy": (C.c_int, [vp]),
        "dsb200_recon_set_async": (C.c_int, [vp, C.c_int]),
        "dsb200_recon_sync": (C.c_int, [vp, _P(Scene), _P(RenderState)]),
        "dsb200_reset_scene": (C.c_int, [vp, _P(Scene)]),
        "dsb200_allocate_from_depth": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking), _P(RenderState), C.c_int, C.c_int]),
        "dsb200_integrate": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking), _P(RenderState)]),
        "dsb200_deintegrate": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking)]),
        "dsb200_decay": (C.c_int, [vp, _P(Scene), _P(RenderState), C.c_int, C.c_int, C.c_int]),
        "dsb200_decay_defusion": (C.c_int, [vp, _P(Scene), _P(RenderState), C.c_int, C.c_int, C.c_int]),
        "dsb200_slide_window": (C.c_int, [vp, _P(Scene), _P(RenderState), C.c_int]),
        "dsb200_slide_window_defusion": (C.c_int, [vp, _P(Scene), _P(RenderState), C.c_int, C.c_int]),
        "dsb200_decayed_block_count": (C.c_size_t, [vp]),
        "dsb200_set_fusion_weight_params": (C.c_int, [vp, _P(WeightParams)]),
        "dsb200_recon_set_sdf_shadow": (C.c_int, [vp, C.c_int]),
        "dsb200_recon_invalidate_sdf_shadow": (C.c_int, [vp]),
        "dsb200_history_sizes": (C.c_int, [vp] + [_P(C.c_int)] * 5),
        "dsb200_vis_create": (C.c_int, [_P(vp), i32, i32, i32, vp]),
        "dsb200_vis_destroy": (C.c_int, [vp]),
        "dsb200_find_visible_blocks": (C.c_int, [vp, _P(Scene), _P(C.c_float), _P(C.c_float), _P(RenderState)]),
        "dsb200_count_visible_blocks": (C.c_int, [vp, _P(Scene), _P(RenderState), C.c_int, C.c_int, _P(C.c_int)]),
        "dsb200_create_expected_depths": (C.c_int, [vp, _P(Scene), _P(C.c_float), _P(C.c_float), _P(RenderState)]),
        "dsb200_find_surface": (C.c_int, [vp, _P(Scene), _P(C.c_float), _P(C.c_float), _P(RenderState)]),
        "dsb200_create_icp_maps": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking), _P(RenderState)]),
        "dsb200_render_image": (C.c_int, [vp, _P(Scene), _P(C.c_float), _P(C.c_float), _P(C.c_float), _P(RenderState), vp, vp, C.c_int]),
        "dsb200_forward_render": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking), _P(RenderState)]),
        "dsb200_forward_projection": (C.c_int, [vp, _P(vp), _P(C.c_int)]),
        "dsb200_create_point_cloud": (C.c_int, [vp]),
        "dsb200_view_from_raw": (C.c_int, [vp, vp, C.c_float, C.c_float, _P(View), vp]),
        "dsb200_process_frame": (C.c_int, [vp, vp, _P(Scene), _P(View), _P(Tracking), _P(RenderState), C.c_int]),
        "dsb200_recon_set_shard": (C.c_int, [vp, _P(Shard)]),
        "dsb200_shard_owner": (C.c_int, [C.c_int] * 5),
        "dsb200_shard_holds": (C.c_int, [C.c_int] * 3 + [_P(Shard)]),
        "dsb200_raycast_sharded": (C.c_int, [vp, _P(Scene), _P(ShardTables), _P(C.c_float), _P(C.c_float), C.c_int, C.c_int, vp, vp]),
        "dsb200_composite_nearest": (C.c_int, [vp, C.c_int, _P(vp), _P(C.c_float), C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
        "dsb200_icp_maps_from_raycast": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking), vp, vp, C.c_int, C.c_int]),
        "dsb200_ipc_export": (C.c_int, [vp, C.c_char_p]),
        "dsb200_ipc_open": (C.c_int, [C.c_char_p, _P(vp)]),
        "dsb200_ipc_close": (C.c_int, [vp]),
        "dsb200_scene_alloc": (C.c_int, [_P(Scene), i32, i32, i32]),
        "dsb200_scene_free": (C.c_int, [_P(Scene)]),
        "dsb200_render_state_alloc": (C.c_int, [_P(RenderState), i32, i32, i32, i32]),
        "dsb200_render_state_free": (C.c_int, [_P(RenderState)]),
        "dsb200_device_alloc": (C.c_int, [_P(vp), C.c_size_t]),
        "dsb200_device_free": (C.c_int, [vp]),
        "dsb200_host_alloc_pinned": (C.c_int, [_P(vp), C.c_size_t]),
        "dsb200_host_free_pinned": (C.c_int, [vp]),
        "dsb200_memcpy_h2d": (C.c_int, [vp, vp, C.c_size_t, vp]),
        "dsb200_memcpy_d2h": (C.c_int, [vp, vp, C.c_size_t, vp]),
        "dsb200_stream_synchronize": (C.c_int, [vp]),
        "dsb200_mat4_inverse": (C.c_int, [_P(C.c_float), _P(C.c_float)]),
        "dsb200_mat4_multiply": (None, [_P(C.c_float), _P(C.c_float), _P(C.c_float)]),
        "dsb200_scene_save": (C.c_int, [_P(Scene), _P(RenderState), cp]),
        "dsb200_scene_load": (C.c_int, [vp, _P(Scene), _P(RenderState), cp]),
        "dsb200_depth_filter_pass": (C.c_int, [vp, vp, i32, i32, vp]),
        "dsb200_convert_disparity_to_depth": (C.c_int, [vp, vp, i32, i32, C.c_float, C.c_float, C.c_float, vp]),
        "dsb200_view_filter_depth": (C.c_int, [_P(View), vp, vp]),
        "dsb200_swap_create": (C.c_int, [_P(vp), i32, i32, vp]),
        "dsb200_swap_destroy": (C.c_int, [vp]),
        "dsb200_swap_cache_of": (C.c_int, [vp, _P(SwapCache)]),
        "dsb200_swap_stored_block": (vp, [vp, i32]),
        "dsb200_swap_mark_visible": (C.c_int, [vp, _P(RenderState)]),
        "dsb200_swap_integrate_global_into_local": (C.c_int, [vp, _P(Scene), _P(i32)]),
        "dsb200_swap_save_to_global_memory": (C.c_int, [vp, _P(Scene), _P(RenderState), _P(i32)]),
        "dsb200_swap_save_file": (C.c_int, [vp, cp]),
        "dsb200_swap_load_file": (C.c_int, [vp, cp]),
        "dsb200_recon_notify_scene_modified": (C.c_int, [vp]),
        "dsb200_recon_notify_volume_swapped": (C.c_int, [vp, _P(Scene)]),
        "dsb200_compute_normal_and_weights": (C.c_int, [vp, vp, vp, i32, i32, _P(C.c_float), vp]),
        "dsb200_recon_prefetch_allocation": (C.c_int, [vp, _P(Scene), _P(View), _P(Tracking), _P(RenderState), vp]),
        "dsb200_recon_join_prefetch": (C.c_int, [vp]),
        "dsb200_recon_abandon_prefetch": (C.c_int, [vp]),
        "dsb200_tracker_create": (C.c_int, [_P(vp), i32, i32, i32, vp]),
        "dsb200_tracker_destroy": (C.c_int, [vp]),
        "dsb200_tracker_set_evaluation_data": (C.c_int, [vp, vp, _P(C.c_float), vp, vp, i32, i32, _P(C.c_float), _P(C.c_float)]),
        "dsb200_tracker_compute_g_and_h": (C.c_int, [vp, i32, i32, C.c_float, _P(C.c_float), _P(C.c_float), _P(C.c_float),
                                                     _P(C.c_float), _P(i32)]),
        "dsb200_tracker_compute_g_and_h_on": (C.c_int, [vp, vp, i32, i32, _P(C.c_float), vp, vp, i32, i32, _P(C.c_float), _P(C.c_float),
                                                        i32, C.c_float, _P(C.c_float), _P(C.c_float), _P(C.c_float), _P(C.c_float), _P(i32)]),
        "dsb200_tracker_last_sums": (C.c_int, [vp, _P(C.c_double)]),
        "dsb200_tracker_compute_g_and_h_static": (C.c_int, [vp, i32, i32, _P(C.c_float), vp, vp, i32, i32, _P(C.c_float), _P(C.c_float),
                                                            i32, C.c_float, _P(C.c_float), vp, vp]),
        "dsb200_tracker_level": (C.c_int, [vp, i32, _P(vp), _P(i32), _P(i32), _P(C.c_float)]),
        "dsb200_filter_subsample_with_holes": (C.c_int, [vp, i32, i32, vp, vp]),
        "dsb200_tracker_track_camera": (C.c_int, [vp, _P(TrackerConfig), _P(View), _P(Tracking), _P(C.c_float), i32, i32,
                                                  _P(TrackerReport)]),
        "dsb200_pose_coerce": (C.c_int, [_P(C.c_float), _P(C.c_float)]),
        "dsb200_pose_from_params": (C.c_int, [_P(C.c_float), _P(C.c_float)]),
        "dsb200_profile_enable": (C.c_int, [vp, C.c_int]),
        "dsb200_profile_read": (C.c_int, [vp, _P(C.c_float)]),
        "dsb200_profile_read_allocate": (C.c_int, [vp, _P(C.c_float)]),
        "dsb200_debug_force_plain_division": (C.c_int, [vp, C.c_int]),
        "dsb200_debug_set_tunable": (C.c_int, [C.c_char_p, C.c_int]),
        "dsb200_selftest_division": (C.c_int, [C.c_ulonglong, C.c_ulonglong, _P(C.c_ulonglong)]),
        "dsb200_selftest_expf": (C.c_int, [vp, vp, i32]),
        "dsb200_selftest_acosf": (C.c_int, [vp, vp, i32]),
        "dsb200_swap_attach_recon": (C.c_int, [vp, vp]),
        "dsb200_swap_set_store_budget": (C.c_int, [vp, i32]),
        "dsb200_recon_refresh_mirrors": (C.c_int, [vp, _P(Scene), _P(RenderState)]),
        "dsb200_mesher_create": (C.c_int, [_P(vp), i32, vp]),
        "dsb200_mesher_destroy": (C.c_int, [vp]),
        "dsb200_mesh_scene": (C.c_int, [vp, _P(Scene), vp, C.c_uint32, _P(C.c_uint32)]),
        "dsb200_kernel_launch_count": (C.c_ulonglong, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)   # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    lib._declared = sorted(sig.keys())
    _lib = lib
    return lib


class Dsb200Error(RuntimeError):
    """Mirrors the reference's error behaviour: std::runtime_error with the reference's strings
    (CUDA/ITMSceneReconstructionEngine_CUDA.cu:350-358) or a CUDA failure (ORcudaSafeCall)."""

    def __init__(self, code, msg):
        RuntimeError.__init__(self, "dsb200 error %d: %s" % (code, msg))
        self.code = code


def check(code):
    if code != OK:
        msg = load().dsb200_last_error()
        raise Dsb200Error(code, msg.decode() if msg else "")
    return code
