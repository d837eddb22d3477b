"""ICP depth tracker on the GPU (through the C ABI) against the oracle and the reference-generated fixture.

Per-pixel terms are bit-exact by construction (same operation order, no contraction); what differs from the reference
CPU engine is only the summation: libdsb200 accumulates in double along a fixed tree, the reference in float serially.
Tolerances, written where they are used:
  valid-point counts and pyramid levels   exact
  sums vs the oracle's double serial sums  |diff| <= 1e-9 * sum|terms| bound (taken as 1e-9 * count * max|H|)  -> in practice
                                           the float results are bit-identical
  float results vs the reference's float sums (fixture)   1e-3 relative to the largest entry
  TrackCamera pose vs the oracle with double sums         2e-6 absolute per matrix entry; vs the reference fixture 1e-5
"""
import importlib
import os
import sys

import numpy as np
import pytest

import harness as hz
from harness import abi, po, synth

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLDEN)
import make_tracker_golden as mt  # noqa: E402
from test_tracker import bits, load_fixture, run_oracle_tracker, tracker_config  # noqa: E402


def eng():
    return importlib.import_module(hz.PKG + ".engine")


class GpuTracker(object):
    """device copies of (depth, points, normals) + an ITMDepthTracker"""

    def __init__(self, depth, points, normals, intr, scene_pose, regime, dist_thresh=0.01, termination=1e-3):
        e = eng()
        H, W = depth.shape
        self.view = e.ITMView(W, H, intr)
        self.view.update(depth, np.zeros((H, W, 4), np.uint8), 0.0)
        self.trk = e.ITMTrackingState(W, H)
        self.trk.points.upload(np.ascontiguousarray(points, np.float32))
        self.trk.normals.upload(np.ascontiguousarray(normals, np.float32))
        self.trk.pose_pointCloud = np.array(scene_pose, np.float32)
        self.tracker = e.ITMDepthTracker((W, H), regime, distThresh=dist_thresh, terminationThreshold=termination)

    def close(self):
        self.tracker.close(); self.view.free(); self.trk.free()


def check_sums(gpu_out, ora_out, k):
    n, f, nabla, hess, sums = gpu_out
    no, fo, nablao, hesso, sumso = ora_out
    assert n == no
    nv = (k * (k + 1)) // 2 + k + 2
    scale = max(np.abs(sumso[:nv - 1]).max(), 1e-30)
    assert np.abs(sums[:nv] - sumso[:nv]).max() <= 1e-9 * scale * max(n, 1) ** 0.5 + 1e-300
    # float outputs: at most one ulp apart (double rounding of two different double sums), normally identical
    for a, b in ((nabla[:k], nablao[:k]), (hess.reshape(6, 6)[:k, :k], hesso.reshape(6, 6)[:k, :k]), (np.float32(f), np.float32(fo))):
        a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
        assert np.abs(a - b).max() <= 1.2e-7 * max(np.abs(b).max(), 1e-30)


def test_fixture_g_and_h_and_pyramid():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    o = po.Oracle()
    g = GpuTracker(depth, points, normals, intr, scene_pose, list(t["regime"]))
    g.tracker.SetEvaluationData(g.trk, g.view)
    inv_start = o.mat4_inverse(t["start_M"])
    row = 0
    d, vi = depth, intr.copy()
    for level in range(int(t["levels"])):
        if level:
            d, vi = o.filter_subsample_with_holes(d), vi * np.float32(0.5)
        gd, gi = g.tracker.Level(level)
        assert mt.sha(gd) == str(t["level_sha"][level]), "pyramid level %d differs from the reference" % level
        assert np.array_equal(bits(gi), bits(vi))
        for it in (abi.TRACKER_ITERATION_ROTATION, abi.TRACKER_ITERATION_TRANSLATION, abi.TRACKER_ITERATION_BOTH):
            k = 3 if it != abi.TRACKER_ITERATION_BOTH else 6
            out = g.tracker.ComputeGandH(level, it, float(t["dist_thresh"]), inv_start)
            ora = o.tracker_compute_g_and_h(d, vi, points, normals, intr, scene_pose, inv_start, float(t["dist_thresh"]), it, double_sums=True)
            check_sums(out, ora, k)
            # against the reference's own (float-accumulated) numbers
            assert out[0] == int(t["gh_n"][row])
            assert abs(out[1] - float(t["gh_f"][row])) <= 1e-3 * float(t["gh_f"][row])
            ref_h = t["gh_hessian"][row][:k, :k]
            assert np.abs(out[3].reshape(6, 6)[:k, :k] - ref_h).max() <= 1e-3 * np.abs(ref_h).max()
            assert np.abs(out[2][:k] - t["gh_nabla"][row][:k]).max() <= 1e-4 * max(np.abs(t["gh_nabla"][row][:k]).max(), 1.0)
            # deterministic: a second evaluation gives the same doubles
            again = g.tracker.ComputeGandH(level, it, float(t["dist_thresh"]), inv_start)
            assert np.array_equal(again[4], out[4])
            row += 1
    assert g.tracker.ComputeGandH(0, abi.TRACKER_ITERATION_NONE, 0.01, inv_start)[0] == 0
    g.close()


def test_fixture_track_camera():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    cfg = tracker_config(int(t["levels"]), t["regime"], float(t["dist_thresh"]), float(t["termination"]))
    g = GpuTracker(depth, points, normals, intr, scene_pose, list(t["regime"]), float(t["dist_thresh"]), float(t["termination"]))
    g.trk.set_pose(t["start_M"])
    g.tracker.TrackCamera(g.trk, g.view)
    M, iM = g.trk.pose()
    Md, iMd, rep = run_oracle_tracker(depth, points, normals, intr, scene_pose, t["start_M"], cfg, double_sums=True)
    assert np.abs(M - Md).max() <= 2e-6 and np.abs(iM - iMd).max() <= 2e-6
    assert (g.tracker.report.evaluations, g.tracker.report.steps, g.tracker.report.last_valid_points) == (rep.evaluations, rep.steps, rep.last_valid_points)
    assert np.abs(M - t["final_M"]).max() <= 1e-5, "pose differs from the reference tracker's"
    assert np.array_equal(bits(iM), bits(po.Oracle().mat4_inverse(M)))
    # deterministic end to end
    g.trk.set_pose(t["start_M"])
    g.tracker.TrackCamera(g.trk, g.view)
    assert np.array_equal(bits(g.trk.pose()[0]), bits(M))
    g.close()


def test_degenerate_and_odd_sizes():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    o = po.Oracle()
    inv = o.mat4_inverse(scene_pose)
    e = eng()
    # no valid depth: zero points, f = 1e5, the pose stays
    g = GpuTracker(np.full_like(depth, -1.0), points, normals, intr, scene_pose, [3, 3])
    g.tracker.SetEvaluationData(g.trk, g.view)
    n, f, nabla, hess, sums = g.tracker.ComputeGandH(0, abi.TRACKER_ITERATION_BOTH, 0.01, inv)
    assert n == 0 and f == np.float32(1e5) and not nabla.any() and not hess.any()
    g.trk.set_pose(scene_pose)
    g.tracker.TrackCamera(g.trk, g.view)
    assert np.array_equal(bits(g.trk.pose()[0]), bits(scene_pose)) and g.tracker.report.steps == 0
    with pytest.raises(abi.Dsb200Error):
        g.tracker.ComputeGandH(5, abi.TRACKER_ITERATION_BOTH, 0.01, inv)
    with pytest.raises(abi.Dsb200Error):
        g.tracker.ComputeGandH(0, 7, 0.01, inv)
    g.close()
    # odd sizes go through the scalar subsampling kernel; the view image differs in size from the scene maps
    odd = np.ascontiguousarray(depth[:119, :157])
    g = GpuTracker(odd, points[:119, :157], normals[:119, :157], intr, scene_pose, [3, 3, 1])
    g.trk.free()
    g.trk = e.ITMTrackingState(160, 120)
    g.trk.points.upload(np.ascontiguousarray(points, np.float32)); g.trk.normals.upload(np.ascontiguousarray(normals, np.float32))
    g.trk.pose_pointCloud = np.array(scene_pose, np.float32)
    g.tracker.SetEvaluationData(g.trk, g.view)
    d = odd
    for level in range(3):
        if level:
            d = o.filter_subsample_with_holes(d)
        assert np.array_equal(bits(g.tracker.Level(level)[0]), bits(d))
    vi = intr * np.float32(0.5)
    out = g.tracker.ComputeGandH(1, abi.TRACKER_ITERATION_BOTH, 0.01, inv)
    ora = o.tracker_compute_g_and_h(o.filter_subsample_with_holes(odd), vi, points, normals, intr, scene_pose, inv, 0.01,
                                    abi.TRACKER_ITERATION_BOTH, double_sums=True)
    check_sums(out, ora, 6)
    g.close()


def test_pose_helpers_bit_exact():
    z, t, _ = load_fixture()
    e = eng()
    o = po.Oracle()
    for p, M, cM, cp in zip(t["pose_params"], t["pose_M"], t["pose_coerced_M"], t["pose_coerced_params"]):
        assert np.array_equal(bits(e.pose_from_params(p)), bits(M))
        m2, p2 = e.pose_coerce(M)
        assert np.array_equal(bits(m2), bits(cM)) and np.array_equal(bits(p2), bits(cp))
    rng = np.random.default_rng(11)
    for k in range(600):
        p = np.concatenate([rng.normal(size=3), rng.normal(size=3) * [1e-5, 5e-4, 0.05, 1.0, 1.8, 3.1][k % 6]]).astype(np.float32)
        M = e.pose_from_params(p)
        assert np.array_equal(bits(M), bits(o.pose_from_params(p)))
        noisy = M + rng.normal(size=16).astype(np.float32) * np.float32(1e-4) * (np.arange(16) % 4 != 3)
        a, b = e.pose_coerce(noisy), o.pose_coerce(noisy)
        assert np.array_equal(bits(a[0]), bits(b[0])) and np.array_equal(bits(a[1]), bits(b[1]))


@pytest.mark.parametrize("cfg_name,recovers,min_points", [("C1", True, 50000), ("C2", False, 20000)])
def test_full_size_tracking_in_the_loop(cfg_name, recovers, min_points):
    """BASELINE's full sizes (C1 640x480 room, C2 1242x375 street): fuse frames at ground-truth poses, then re-track the
    last frame from a perturbed pose against the maps the GPU raycast itself produced; the GPU tracker and the oracle
    (double sums, on the GPU's maps copied to the host) must end on the same pose.  In the room that pose must also undo
    the perturbation; the street scene (ground plane + distant facades, 2 cm voxels, 10 cm gate) is degenerate for
    point-to-plane ICP, so there only agreement with the oracle and a non-increasing error are required."""
    cfg, frames = synth.make_sequence(cfg_name, 3)
    gpu = hz.GpuBackend(cfg)
    for f in frames:
        gpu.set_frame(f)
        assert gpu.process_frame() == 0
    e = eng()
    o = po.Oracle()
    regime = list(e.ITMDepthTracker.DEFAULT_REGIME)
    tracker = e.ITMDepthTracker((cfg.W, cfg.H), regime, distThresh=0.01)
    f2, scene_pose = frames[2], frames[2].M
    assert np.array_equal(bits(gpu.trk.pose_pointCloud), bits(scene_pose))
    start = mt.perturbed_start(o, scene_pose)
    gpu.trk.set_pose(start)
    points, normals = gpu.trk.icp_points(), gpu.trk.icp_normals()
    tracker.TrackCamera(gpu.trk, gpu.view)
    M, iM = gpu.trk.pose()
    tc = tracker_config(5, regime, 0.01, 1e-3)
    depth = np.ascontiguousarray(f2.depth.reshape(cfg.H, cfg.W), np.float32)
    Mo, iMo, rep = run_oracle_tracker(depth, points, normals, np.array(cfg.intr, np.float32), scene_pose, start, tc, double_sums=True)
    assert np.abs(M - Mo).max() <= 2e-6
    assert (tracker.report.evaluations, tracker.report.steps) == (rep.evaluations, rep.steps)
    assert tracker.report.last_valid_points > min_points
    err0, err1 = np.abs(start - scene_pose).max(), np.abs(M - scene_pose).max()
    if recovers:
        assert err1 < 0.1 * err0 + 1e-3, "tracking did not undo the perturbation (%g -> %g)" % (err0, err1)
    else:
        assert err1 <= err0 * 1.05
    # the stateless entry point (what the C++ adapter forwards to) gives the same sums as the level-based one
    tracker.SetEvaluationData(gpu.trk, gpu.view)
    inv = o.mat4_inverse(start)
    a = tracker.ComputeGandH(0, abi.TRACKER_ITERATION_BOTH, 0.01, inv)
    b = tracker.ComputeGandHOn(gpu.view, gpu.trk, abi.TRACKER_ITERATION_BOTH, 0.01, inv)
    assert a[0] == b[0] and np.array_equal(a[4], b[4])
    tracker.close(); gpu.close()
