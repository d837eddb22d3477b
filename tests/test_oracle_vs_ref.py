"""Pin the CPU restatement (oracle/restate.cpp) against the UNMODIFIED reference CPU engines
compiled from /root/reference (oracle/_ref).  Bit-exact on every output.  Skipped where the
reference build is absent; the committed fixtures (test_golden.py) carry the same evidence."""
import numpy as np
import pytest

import harness as hz
from harness import abi, synth, po

pytestmark = pytest.mark.skipif(not hz.has_ref(), reason="oracle/_ref not built (needs /root/reference)")


def run_pair(cfg_name, n_frames, use_swapping=False, weights=False):
    cfg, frames = synth.make_sequence(cfg_name, n_frames)
    ref = po.RefEngine(cfg.W, cfg.H, cfg.params(), use_swapping=use_swapping)
    ref.set_calib(cfg.intr)
    ora = hz.OracleBackend(cfg, use_swapping=use_swapping)
    for i, f in enumerate(frames):
        ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
        Mr, iMr = ref.get_pose()
        assert np.array_equal(Mr, f.M)
        assert np.array_equal(iMr.view(np.uint32), ora.o.mat4_inverse(f.M).view(np.uint32))
        ref.allocate(); ref.integrate(); ref.create_expected_depths(); ref.create_icp_maps()
        ora.set_frame(f); ora.process_frame()
        s = ora.state(copy=False)
        rstate = dict(hash=ref.hash(), visible=ref.visible(), visible_type=ref.visible_type(), last_free=ref.last_free(),
                      alloc_list=ref.alloc_list(), excess_list=ref.excess_list(), voxels=ref.voxels())
        hz.assert_state_equal(rstate, s, "frame %d ref-vs-oracle" % i)
        hz.assert_images_equal(dict(range=ref.range_image(), raycast=ref.raycast_result(), points=ref.icp_points(),
                                    normals=ref.icp_normals(), rendering=ref.icp_rendering()), ora.images(), "frame %d" % i)
    return cfg, frames, ref, ora


def test_c1_sequence_bit_exact():
    cfg, frames, ref, ora = run_pair("C1", 3)
    assert ora.rs.c.no_visible > 1000
    ref.close(); ora.close()


def test_tiny_sequence_and_free_view_bit_exact():
    cfg, frames, ref, ora = run_pair("tiny", 8)
    # free-view render path (ITMMainEngine::GetImage): FindVisibleBlocks + CreateExpectedDepths + RenderImage
    M = frames[3].M
    for rtype in (abi.RENDER_SHADED_GREYSCALE, abi.RENDER_COLOUR_FROM_VOLUME, abi.RENDER_COLOUR_FROM_NORMAL):
        ref.render_free_view(M, cfg.intr, rtype)
        ora.out_img[:] = 0
        ref.out_image()[:] = 0 if rtype == abi.RENDER_COLOUR_FROM_NORMAL else ref.out_image()
        if rtype == abi.RENDER_COLOUR_FROM_NORMAL:
            ref.render_free_view(M, cfg.intr, rtype)
        ora.free_view(M, cfg.intr, rtype)
        fi = ora.free_images()
        assert np.array_equal(ref.visible(True), fi["visible"])
        hz.assert_images_equal(dict(range=ref.range_image(True), raycast=ref.raycast_result(True), image=ref.out_image()),
                               dict(range=fi["range"], raycast=fi["raycast"], image=fi["image"]), "free view type %d" % rtype)
    assert (fi["raycast"][..., 3] > 0).sum() > 1000
    ref.close(); ora.close()


def test_use_swapping_visibility_bit_exact():
    # app default useSwapping=true enlarges the visibility frustum
    cfg, frames, ref, ora = run_pair("tiny", 6, use_swapping=True)
    ref.close(); ora.close()


def test_forward_render_matches_reference():
    """ForwardRender (CPU/ITMVisualisationEngine_CPU.cpp:464-540): restatement == reference on forwardProjection,
    the missing-point count and the shaded image."""
    for name in ("tiny", "tiny_street"):
        cfg, frames = synth.make_sequence(name, 4)
        ref = po.RefEngine(cfg.W, cfg.H, cfg.params())
        ref.set_calib(cfg.intr)
        ora = hz.OracleBackend(cfg)
        for f in frames[:3]:
            ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
            ref.allocate(); ref.integrate(); ref.create_expected_depths(); ref.create_icp_maps()
            ora.set_frame(f); ora.process_frame()
        f = frames[3]
        ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
        ref.create_expected_depths()
        ora.set_frame(f); ora.expected_depths()
        fr, nr = ref.forward_render()
        fo, no, ro = ora.forward_render()
        assert nr == no and no > 0
        hz.assert_images_equal(dict(fwd=fr, rendering=ref.icp_rendering()), dict(fwd=fo, rendering=ro), name + " forward render")
        ref.close(); ora.close()


def test_view_building_matches_reference():
    """ITMViewBuilder_CPU::UpdateView (TRAFO_AFFINE, the calibration DenseSLAM sets): restatement == reference, and the synthetic
    generator's float depth is exactly what the reference would have produced from its int16 millimetres."""
    rng = np.random.default_rng(3)
    for name in ("tiny", "tiny_street"):
        cfg, frames = synth.make_sequence(name, 2)
        ref = po.RefEngine(cfg.W, cfg.H, cfg.params())
        ref.set_calib(cfg.intr)
        for f in frames:
            raw = f.depth_mm.copy()
            raw.reshape(-1)[rng.integers(0, raw.size, 50)] = rng.integers(-32768, 32767, 50).astype(np.int16)   # negatives, > 32000
            got = ref.update_view_raw(raw, f.rgb, np.float32(1.0 / 1000.0), 0.0, f.timestamp)
            want, _ = hz.view_from_raw_restated(raw, None, np.float32(1.0 / 1000.0), 0.0)
            assert np.array_equal(hz.bits(got), hz.bits(want))
            same = raw == f.depth_mm
            assert np.array_equal(hz.bits(got[same]), hz.bits(f.depth[same]))
            # useBilateralFilter: five 5x5 passes (CPU/ITMViewBuilder_CPU.cpp:49-58), restated bit for bit
            got_f = ref.update_view_raw(raw, f.rgb, np.float32(1.0 / 1000.0), 0.0, f.timestamp, use_bilateral_filter=True)
            want_f = po.Oracle().view_filter_depth(want.reshape(cfg.H, cfg.W))
            assert np.array_equal(hz.bits(got_f), hz.bits(want_f))
            assert not np.array_equal(got_f, got) and (got_f[2:-2, 2:-2] > 0).mean() > 0.2
        # Kinect-style disparity calibration (TRAFO_KINECT): zero denominators, negative and non-positive depths included
        disp = rng.integers(-200, 1200, (cfg.H, cfg.W)).astype(np.int16)
        disp[0, :8] = 1090
        a = ref.convert_disparity_to_depth(disp, 1090.0, 0.075, 573.71)
        b = po.Oracle().convert_disparity_to_depth(disp, 1090.0, 0.075, 573.71)
        assert np.array_equal(hz.bits(a), hz.bits(b)) and (a == -1.0).any() and (a > 0).any()
        ref.close()


def test_normal_and_weights_match_reference():
    """ITMViewBuilder_CPU::ComputeNormalAndWeights (UpdateView's modelSensorNoise step): restatement == reference bit for bit,
    including which elements are left untouched (border, normal.xyz of rejected pixels)."""
    o = po.Oracle()
    ref = po.RefEngine(8, 8, synth.make_sequence("tiny", 1)[0].params())
    seen_valid = seen_rejected = 0
    for name, depth, intr in hz.normal_and_weight_cases():
        H, W = depth.shape
        fill_n = np.full((H, W, 4), 7.25, np.float32); fill_s = np.full((H, W), -3.5, np.float32)
        n_ref, s_ref = fill_n.copy(), fill_s.copy()
        n_ora, s_ora = fill_n.copy(), fill_s.copy()
        ref.compute_normal_and_weights(depth, intr, n_ref, s_ref)
        o.compute_normal_and_weights(depth, intr, n_ora, s_ora)
        assert np.array_equal(hz.bits(n_ref), hz.bits(n_ora)), name
        assert np.array_equal(hz.bits(s_ref), hz.bits(s_ora)), name
        assert (s_ref[:2] == -3.5).all() and (s_ref[:, -2:] == -3.5).all() and (n_ref[-2:] == 7.25).all(), name
        seen_valid += int((n_ref[..., 3] == 1.0).sum()); seen_rejected += int((n_ref[..., 3] == -1.0).sum())
        rej = n_ref[..., 3] == -1.0
        assert (n_ref[rej][:, :3] == 7.25).all() and (s_ref[rej] == -1.0).all(), name
    assert seen_valid > 5000 and seen_rejected > 100
    ref.close()


def test_swapping_engine_matches_reference():
    """ITMSwappingEngine_CPU on ITMGlobalCache (IntegrateGlobalIntoLocal, both SaveToGlobalMemory overloads) and the
    allocator's swap-state pass: restatement == reference on the volume, the free list, the swap states and the host store."""
    cfg, frames, order, whole_map_after = hz.swap_scenario()
    ref = po.RefEngine(cfg.W, cfg.H, cfg.params(), use_swapping=True)
    ref.set_calib(cfg.intr)
    ora = hz.OracleBackend(cfg, use_swapping=True)
    cache = po.HostSwapCache(ora.scene.c.total_entries)
    moved_in = moved_out = 0
    for step, k in enumerate(order):
        f = frames[k]
        ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
        ref.allocate(); ref.integrate()
        ref.swap_integrate_global_into_local(); ref.swap_save_to_global_memory(True)
        ora.set_frame(f); ora.allocate(); ora.o.swap_mark_visible(cache, ora.rs); ora.integrate()
        moved_in += ora.o.swap_integrate_global_into_local(cache, ora.scene)
        moved_out += ora.o.swap_save_to_global_memory(cache, ora.scene, ora.rs)
        if step == whole_map_after:
            ref.swap_save_to_global_memory(False)
            n = ora.o.swap_save_to_global_memory(cache, ora.scene, None)
            assert n > 300 and (ora.scene.hash["ptr"] == -1).sum() >= n
        what = "swap step %d (frame %d)" % (step, k)
        s = ora.state(copy=False)               # compared at once: no copy of the 1 GiB voxel array
        rstate = dict(hash=ref.hash(), visible=ref.visible(), visible_type=ref.visible_type(), last_free=ref.last_free(),
                      alloc_list=ref.alloc_list(), excess_list=ref.excess_list())
        hz.assert_state_equal(rstate, s, what, voxels=False)
        # the 1 GiB voxel arrays at the steps where the swapping paths change (first frame, return of the first stretch, whole-map
        # swap-out and the frame after it, end); tables, states and store contents at every step
        if step in (0, 4, whole_map_after, whole_map_after + 1, len(order) - 1):
            assert hz.voxels_equal_but_pad(ref.voxels(), s["voxels"]), what
        assert np.array_equal(ref.swap_states(), cache.states), what
        has = ref.swap_has_stored()
        assert np.array_equal(has, cache.has_stored), what
        for e in np.nonzero(has)[0][::7]:
            assert hz.voxels_equal_but_pad(ref.swap_stored_block(e), cache.stored_block(e)), what
    assert moved_out > 300 and moved_in > 1000 and (cache.states == 0).any() and (cache.states == 2).any()
    ref.close(); ora.close()
