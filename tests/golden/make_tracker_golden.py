#!/usr/bin/env python
"""Generates tests/golden/tiny_tracker_ref.npz from the UNMODIFIED reference depth tracker
(ITMDepthTracker_CPU + ITMLowLevelEngine_CPU + ITMPose, compiled into oracle/_ref/libitmref.so).

Run where /root/reference exists:   python tests/golden/make_tracker_golden.py
Inputs come from tiny_room_ref.npz (the last frame's depth, the reference's ICP maps of that frame and its pose);
this file stores only what the tracker adds: the perturbed start pose, the reference's ComputeGandH outputs for
every (pyramid level, iteration type), digests of the pyramid levels, pose-conversion known answers and the pose
TrackCamera ends at.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import harness as hz  # noqa: E402
from harness import abi, po  # noqa: E402

LEVELS = 3
REGIME = [abi.TRACKER_ITERATION_BOTH, abi.TRACKER_ITERATION_BOTH, abi.TRACKER_ITERATION_ROTATION]
DIST_THRESH = 0.01
TERMINATION = 1e-3


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def tracker_inputs(z):
    """(depth float32 HxW, points, normals HxWx4, intr, scene pose) of the last frame of a room fixture"""
    W, H = int(z["W"]), int(z["H"])
    mm = z["depth_mm"][-1]
    depth = np.where((mm <= 0) | (mm > 32000), np.float32(-1.0), mm.astype(np.float32) * np.float32(1.0 / 1000.0) + np.float32(0.0)).astype(np.float32)
    return (depth.reshape(H, W), z["points"].reshape(H, W, 4).astype(np.float32), z["normals"].reshape(H, W, 4).astype(np.float32),
            z["intr"].astype(np.float32), z["M"][-1].astype(np.float32))


def perturbed_start(ref_or_oracle, scene_pose):
    """scene pose moved by a small rigid increment: the start pose TrackCamera has to undo"""
    inc = ref_or_oracle.pose_from_params(np.array([0.012, -0.008, 0.010, 0.006, -0.004, 0.005], np.float32))
    return ref_or_oracle.mat4_mul(inc, scene_pose) if hasattr(ref_or_oracle, "mat4_mul") else ref_or_oracle.mat4_multiply(inc, scene_pose)


def main():
    if not hz.has_ref():
        raise SystemExit("oracle/_ref/libitmref.so absent: run `make -C oracle` where /root/reference exists")
    z = np.load(os.path.join(HERE, "tiny_room_ref.npz"))
    depth, points, normals, intr, scene_pose = tracker_inputs(z)
    H, W = depth.shape
    cfg = hz.synth.Config("fixture", W, H, intr, float(z["voxel_size"]), float(z["mu"]), int(z["max_w"]), float(z["vf_min"]),
                          float(z["vf_max"]), "fixture")
    ref = po.RefEngine(W, H, cfg.params())
    ref.set_calib(intr)
    start = perturbed_start(ref, scene_pose)
    inv_start = ref.mat4_inv(start)
    out = dict(levels=LEVELS, regime=np.array(REGIME, np.int32), dist_thresh=np.float32(DIST_THRESH),
               termination=np.float32(TERMINATION), start_M=start)
    # pyramid + ComputeGandH per (level, type)
    d, vi = depth, intr.copy()
    gh_keys, gh_n, gh_f, gh_nabla, gh_hess, level_sha = [], [], [], [], [], []
    for level in range(LEVELS):
        if level:
            d = ref.filter_subsample_with_holes(d)
            vi = vi * np.float32(0.5)
        level_sha.append(sha(d))
        for it in (abi.TRACKER_ITERATION_ROTATION, abi.TRACKER_ITERATION_TRANSLATION, abi.TRACKER_ITERATION_BOTH):
            n, f, nabla, hess, _ = ref.tracker_compute_g_and_h(d, vi, points, normals, intr, scene_pose, inv_start, DIST_THRESH, it)
            k = 3 if it != abi.TRACKER_ITERATION_BOTH else 6
            h = hess.reshape(6, 6)[:k, :k].copy()
            gh_keys.append((level, it)); gh_n.append(n); gh_f.append(f)
            gh_nabla.append(np.pad(nabla[:k], (0, 6 - k))); gh_hess.append(np.pad(h, ((0, 6 - k), (0, 6 - k))))
    out.update(gh_keys=np.array(gh_keys, np.int32), gh_n=np.array(gh_n, np.int32), gh_f=np.array(gh_f, np.float32),
               gh_nabla=np.array(gh_nabla, np.float32), gh_hessian=np.array(gh_hess, np.float32), level_sha=np.array(level_sha))
    # pose conversions: known answers over the three angle ranges and the near-pi branch
    rng = np.random.default_rng(7)
    params = np.stack([np.concatenate([rng.normal(size=3), rng.normal(size=3) * s]) for s in (1e-5, 5e-4, 0.05, 1.0, 1.8) for _ in range(4)]).astype(np.float32)
    out["pose_params"] = params
    out["pose_M"] = np.stack([ref.pose_from_params(p) for p in params])
    co = [ref.pose_coerce(m) for m in out["pose_M"]]
    out["pose_coerced_M"] = np.stack([c[0] for c in co]); out["pose_coerced_params"] = np.stack([c[1] for c in co])
    # the whole tracker
    ref.set_view(depth, np.zeros((H, W, 4), np.uint8), 0.0)
    ref.L.itmref_set_pose(ref.h, po.ptr(scene_pose))
    # the point cloud of the fixture's last frame goes into the reference's tracking state as CreateICPMaps left it
    ref.icp_points()[:] = points.reshape(ref.icp_points().shape)
    ref.icp_normals()[:] = normals.reshape(ref.icp_normals().shape)
    set_scene_pose(ref, scene_pose)
    ref.set_pose(start)
    M, iM = ref.track_camera(REGIME, 0, DIST_THRESH, TERMINATION)
    out["final_M"] = M
    ref.close()
    path = os.path.join(HERE, "tiny_tracker_ref.npz")
    np.savez_compressed(path, **out)
    print("tiny_tracker_ref.npz: %dx%d, valid points at level 0: %d, |final - scene pose| = %.2e (start %.2e), %.1f KiB"
          % (W, H, gh_n[2], np.abs(M - scene_pose).max(), np.abs(start - scene_pose).max(), os.path.getsize(path) / 1024.0))


def set_scene_pose(ref, M):
    m = np.ascontiguousarray(M, np.float32)
    ref.L.itmref_set_scene_pose.argtypes = [__import__("ctypes").c_void_p] * 2
    ref.L.itmref_set_scene_pose(ref.h, po.ptr(m))


if __name__ == "__main__":
    main()
