"""ICP depth tracker, CPU leg: the oracle's restatement of ComputeGandH, the depth pyramid,
the pose conversions and the whole TrackCamera loop against (1) the committed fixture generated from the reference
(tests/golden/make_tracker_golden.py) and (2) the reference itself where oracle/_ref exists.  Bit-exact throughout: the
oracle's default sums are the reference CPU engine's float accumulators."""
import os
import sys

import numpy as np
import pytest

import harness as hz
from harness import abi, po, synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLDEN)
import make_tracker_golden as mt  # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def load_fixture():
    z = np.load(os.path.join(GOLDEN, "tiny_room_ref.npz"))
    t = np.load(os.path.join(GOLDEN, "tiny_tracker_ref.npz"))
    return z, t, mt.tracker_inputs(z)


def tracker_config(levels, regime, dist_thresh, termination, till=0):
    c = abi.TrackerConfig()
    c.levels = levels
    for i in range(levels):
        c.regime[i] = int(regime[i])
    c.no_icp_run_till_level, c.dist_thresh, c.termination_threshold = till, dist_thresh, termination
    return c


def test_fixture_g_and_h_pyramid_and_pose_bit_exact():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    o = po.Oracle()
    start = mt.perturbed_start(o, scene_pose)
    assert np.array_equal(bits(start), bits(t["start_M"]))
    inv_start = o.mat4_inverse(start)
    d, vi = depth, intr.copy()
    row = 0
    for level in range(int(t["levels"])):
        if level:
            d = o.filter_subsample_with_holes(d)
            vi = vi * np.float32(0.5)
        assert mt.sha(d) == str(t["level_sha"][level]), "pyramid level %d differs from the reference" % level
        for it in (abi.TRACKER_ITERATION_ROTATION, abi.TRACKER_ITERATION_TRANSLATION, abi.TRACKER_ITERATION_BOTH):
            assert tuple(t["gh_keys"][row]) == (level, it)
            n, f, nabla, hess, sums = o.tracker_compute_g_and_h(d, vi, points, normals, intr, scene_pose, inv_start, float(t["dist_thresh"]), it)
            k = 3 if it != abi.TRACKER_ITERATION_BOTH else 6
            assert n == int(t["gh_n"][row])
            assert np.array_equal(bits(np.float32(f)), bits(t["gh_f"][row]))
            assert np.array_equal(bits(nabla[:k]), bits(t["gh_nabla"][row][:k]))
            assert np.array_equal(bits(hess.reshape(6, 6)[:k, :k]), bits(t["gh_hessian"][row][:k, :k]))
            assert sums[(k * (k + 1)) // 2 + k + 1] == n
            row += 1
    assert int(t["gh_n"][2]) > 1000
    for p, M, cM, cp in zip(t["pose_params"], t["pose_M"], t["pose_coerced_M"], t["pose_coerced_params"]):
        assert np.array_equal(bits(o.pose_from_params(p)), bits(M))
        m2, p2 = o.pose_coerce(M)
        assert np.array_equal(bits(m2), bits(cM)) and np.array_equal(bits(p2), bits(cp))


def run_oracle_tracker(depth, points, normals, intr, scene_pose, start, cfg, double_sums=False):
    o = po.Oracle()
    view = po.make_view(depth, np.zeros(depth.shape + (4,), np.uint8), intr)
    trk = po.make_tracking(start, o.mat4_inverse(start), points, normals)
    rep = o.track_camera(cfg, view, trk, scene_pose, depth.shape[1], depth.shape[0], double_sums=double_sums)
    return np.array(trk.M_d[:], np.float32), np.array(trk.invM_d[:], np.float32), rep


def test_fixture_track_camera_bit_exact():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    cfg = tracker_config(int(t["levels"]), t["regime"], float(t["dist_thresh"]), float(t["termination"]))
    M, iM, rep = run_oracle_tracker(depth, points, normals, intr, scene_pose, t["start_M"], cfg)
    assert np.array_equal(bits(M), bits(t["final_M"])), "TrackCamera ends at a different pose than the reference"
    assert np.array_equal(bits(iM), bits(po.Oracle().mat4_inverse(M)))
    assert rep.steps >= 3 and rep.last_valid_points > 1000
    # the tracker undoes most of the perturbation
    assert np.abs(M - scene_pose).max() < 0.1 * np.abs(t["start_M"] - scene_pose).max()
    # double-precision sums (the libdsb200 contract) land on the same pose to float rounding
    Md, _, _ = run_oracle_tracker(depth, points, normals, intr, scene_pose, t["start_M"], cfg, double_sums=True)
    assert np.abs(Md - M).max() < 2e-6


def test_degenerate_inputs():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    o = po.Oracle()
    inv = o.mat4_inverse(scene_pose)
    # iteration type NONE: nothing evaluated (ITMDepthTracker_CPU.cpp:25)
    assert o.tracker_compute_g_and_h(depth, intr, points, normals, intr, scene_pose, inv, 0.01, abi.TRACKER_ITERATION_NONE)[0] == 0
    # no valid depth at all / all-hole maps: zero points, f = 1e5, sums zero
    for d, p in ((np.full_like(depth, -1.0), points), (depth, np.full_like(points, -1.0))):
        n, f, nabla, hess, sums = o.tracker_compute_g_and_h(d, intr, p, normals, intr, scene_pose, inv, 0.01, abi.TRACKER_ITERATION_BOTH)
        assert n == 0 and f == np.float32(1e5) and not nabla.any() and not hess.any()
    # at most 100 valid points also reports f = 1e5 (:76)
    few = np.full_like(depth, -1.0)
    ys, xs = np.nonzero(points[..., 3] >= 0)
    few[ys[:2000:25], xs[:2000:25]] = depth[ys[:2000:25], xs[:2000:25]]
    n, f, _, _, _ = o.tracker_compute_g_and_h(few, intr, points, normals, intr, scene_pose, inv, 0.01, abi.TRACKER_ITERATION_BOTH)
    assert 0 < n <= 100 and f == np.float32(1e5)
    # a tracker that never sees a valid evaluation leaves the pose alone
    cfg = tracker_config(2, [3, 3], 0.01, 1e-3)
    M, iM, rep = run_oracle_tracker(np.full_like(depth, -1.0), points, normals, intr, scene_pose, scene_pose, cfg)
    assert np.array_equal(bits(M), bits(scene_pose)) and rep.steps == 0 and rep.evaluations == 2
    # odd image sizes: the pyramid drops the last row / column (ITMLowLevelEngine_CPU.cpp:53)
    odd = depth[:119, :157]
    half = o.filter_subsample_with_holes(odd)
    assert half.shape == (59, 78)
    assert np.array_equal(bits(half), bits(o.filter_subsample_with_holes(odd[:118, :156])))


def test_g_and_h_properties():
    z, t, (depth, points, normals, intr, scene_pose) = load_fixture()
    o = po.Oracle()
    inv = o.mat4_inverse(t["start_M"])
    n, f, nabla, hess, sums = o.tracker_compute_g_and_h(depth, intr, points, normals, intr, scene_pose, inv, 0.01,
                                                        abi.TRACKER_ITERATION_BOTH, double_sums=True)
    Hm = hess.reshape(6, 6).astype(np.float64)
    assert np.array_equal(Hm, Hm.T)
    assert np.linalg.eigvalsh(Hm).min() > -1e-6 * np.abs(Hm).max()      # sum of outer products
    # the short iterations are the diagonal blocks of the full one (same pixels pass the same tests)
    nr, _, nabr, hr, _ = o.tracker_compute_g_and_h(depth, intr, points, normals, intr, scene_pose, inv, 0.01,
                                                   abi.TRACKER_ITERATION_ROTATION, double_sums=True)
    nt, _, nabt, ht, _ = o.tracker_compute_g_and_h(depth, intr, points, normals, intr, scene_pose, inv, 0.01,
                                                   abi.TRACKER_ITERATION_TRANSLATION, double_sums=True)
    assert nr == n and nt == n
    assert np.array_equal(bits(hr.reshape(6, 6)[:3, :3]), bits(hess.reshape(6, 6)[:3, :3]))
    assert np.array_equal(bits(ht.reshape(6, 6)[:3, :3]), bits(hess.reshape(6, 6)[3:, 3:]))
    assert np.array_equal(bits(nabr[:3]), bits(nabla[:3])) and np.array_equal(bits(nabt[:3]), bits(nabla[3:]))
    # float and double accumulation agree to float-summation error
    nf, ff, nabf, hf, _ = o.tracker_compute_g_and_h(depth, intr, points, normals, intr, scene_pose, inv, 0.01, abi.TRACKER_ITERATION_BOTH)
    assert nf == n and abs(ff - f) < 1e-4 * f
    assert np.abs(hf - hess).max() < 1e-3 * np.abs(hess).max()


@pytest.mark.skipif(not hz.has_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_vs_reference_c1_five_levels():
    """At 320x240 with the application's five-level regime, on maps the reference raycast itself."""
    cfg, frames = synth.make_sequence("C1", 4)
    ref = po.RefEngine(cfg.W, cfg.H, cfg.params())
    ref.set_calib(cfg.intr)
    for f in frames[:3]:
        ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
        ref.allocate(); ref.integrate(); ref.create_expected_depths(); ref.create_icp_maps()
    points = ref.icp_points().reshape(cfg.H, cfg.W, 4).copy()
    normals = ref.icp_normals().reshape(cfg.H, cfg.W, 4).copy()
    o = po.Oracle()
    f3, scene_pose = frames[3], frames[2].M
    depth = np.ascontiguousarray(f3.depth.reshape(cfg.H, cfg.W), np.float32)
    inv_prev = o.mat4_inverse(scene_pose)
    d, vi = depth, np.array(cfg.intr, np.float32)
    for level in range(5):
        if level:
            d2 = o.filter_subsample_with_holes(d)
            assert np.array_equal(bits(d2), bits(ref.filter_subsample_with_holes(d)))
            d, vi = d2, vi * np.float32(0.5)
        for it in (1, 2, 3):
            a = ref.tracker_compute_g_and_h(d, vi, points, normals, cfg.intr, scene_pose, inv_prev, 0.01, it)
            b = o.tracker_compute_g_and_h(d, vi, points, normals, cfg.intr, scene_pose, inv_prev, 0.01, it)
            k = 3 if it != 3 else 6
            assert a[0] == b[0] and np.array_equal(bits(np.float32(a[1])), bits(np.float32(b[1])))
            assert np.array_equal(bits(a[2][:k]), bits(b[2][:k]))
            assert np.array_equal(bits(a[3].reshape(6, 6)[:k, :k]), bits(b[3].reshape(6, 6)[:k, :k]))
    # the whole loop, started from the previous frame's pose like ITMMainEngine does
    ref.set_view(f3.depth, f3.rgb, f3.timestamp)
    regime = [3, 3, 1, 1, 1]
    M, iM = ref.track_camera(regime, 0, 0.01, 1e-3)
    tc = tracker_config(5, regime, 0.01, 1e-3)
    Mo, iMo, rep = run_oracle_tracker(depth, points, normals, np.array(cfg.intr, np.float32), scene_pose, scene_pose, tc)
    assert np.array_equal(bits(M), bits(Mo)) and np.array_equal(bits(iM), bits(iMo))
    assert np.abs(M - f3.M).max() < 0.05 * np.abs(scene_pose - f3.M).max() + 1e-3
    ref.close()


@pytest.mark.skipif(not hz.has_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_pose_conversions_vs_reference():
    cfg, _ = synth.make_sequence("tiny", 1)
    ref = po.RefEngine(cfg.W, cfg.H, cfg.params())
    o = po.Oracle()
    rng = np.random.default_rng(3)
    for k in range(1500):
        scale = [1e-5, 5e-4, 0.05, 1.0, 1.8, 3.1][k % 6]
        p = np.concatenate([rng.normal(size=3), rng.normal(size=3) * scale]).astype(np.float32)
        Ma, Mb = ref.pose_from_params(p), o.pose_from_params(p)
        assert np.array_equal(bits(Ma), bits(Mb))
        noisy = Ma + rng.normal(size=16).astype(np.float32) * np.float32(1e-4) * (np.arange(16) % 4 != 3)
        ca, pa = ref.pose_coerce(noisy)
        cb, pb = o.pose_coerce(noisy)
        assert np.array_equal(bits(ca), bits(cb)) and np.array_equal(bits(pa), bits(pb))
    ref.close()
