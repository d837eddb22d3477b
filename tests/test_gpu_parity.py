"""GPU parity tests proper: libdsb200 (through the C ABI) vs the CPU oracle on the same seeded inputs.
Bit-exact on everything (the north star only asks +-1 LSB on voxels and 1e-3 on raycast depth)."""
import numpy as np
import pytest

import harness as hz
from harness import abi, synth

pytestmark = pytest.mark.gpu


def run_sequence(cfg_name, n_frames, caps, fused=True, use_swapping=False, weights=None, check_every=1, seed=0):
    cfg, frames = synth.make_sequence(cfg_name, n_frames, seed=seed)
    gpu, ora = hz.GpuBackend(cfg, caps, use_swapping), hz.OracleBackend(cfg, caps, use_swapping)
    if weights:
        gpu.set_weights(*weights); ora.set_weights(*weights)
    for i, f in enumerate(frames):
        gpu.set_frame(f); ora.set_frame(f)
        if fused:
            rg, ro = gpu.process_frame(), ora.process_frame()
        else:
            rg, ro = gpu.allocate(), ora.allocate()
            gpu.integrate(); ora.integrate()
            gpu.expected_depths(); ora.expected_depths()
            gpu.icp_maps(); ora.icp_maps()
        assert rg == ro
        if i % check_every == 0 or i == n_frames - 1:
            hz.assert_state_equal(ora.state(), gpu.state(), "%s frame %d" % (cfg_name, i))
            hz.assert_images_equal(ora.images(), gpu.images(), "%s frame %d" % (cfg_name, i))
    return cfg, frames, gpu, ora


def test_tiny_room_fused_sequence():
    cfg, frames, gpu, ora = run_sequence("tiny", 12, hz.SMALL_CAPS)
    assert gpu.rs.c.no_visible > 200
    gpu.close(); ora.close()


def test_tiny_room_stagewise_calls():
    cfg, frames, gpu, ora = run_sequence("tiny", 4, hz.SMALL_CAPS, fused=False)
    gpu.close(); ora.close()


def test_tiny_street_depth_weighting():
    cfg, frames, gpu, ora = run_sequence("tiny_street", 8, hz.SMALL_CAPS, weights=(True, 10, 100))
    gpu.close(); ora.close()


def test_use_swapping_enlarged_frustum():
    cfg, frames, gpu, ora = run_sequence("tiny", 6, hz.SMALL_CAPS, use_swapping=True)
    gpu.close(); ora.close()


def test_heavy_hash_collisions_excess_chains():
    # 256 buckets for thousands of blocks: almost every allocation goes through the excess list
    caps = dict(num_blocks=0x4000, bucket_num=256, excess_size=0x4000)
    cfg, frames, gpu, ora = run_sequence("tiny", 6, caps)
    h = gpu.scene.hash()
    assert (h["offset"] > 0).sum() > 200
    gpu.close(); ora.close()


def test_vba_exhaustion_reports_reference_error():
    caps = dict(num_blocks=256, bucket_num=0x8000, excess_size=0x2000)
    cfg, frames = synth.make_sequence("tiny", 2)
    gpu, ora = hz.GpuBackend(cfg, caps), hz.OracleBackend(cfg, caps)
    gpu.set_frame(frames[0]); ora.set_frame(frames[0])
    rg, ro = gpu.allocate(), ora.allocate()
    assert rg == ro == abi.ERR_VBA_EXHAUSTED
    hz.assert_state_equal(ora.state(voxels=False), gpu.state(voxels=False), "exhausted", free_lists=False)
    gpu.close(); ora.close()


def test_free_view_render_all_types():
    cfg, frames, gpu, ora = run_sequence("tiny", 6, hz.SMALL_CAPS, check_every=100)
    M = frames[2].M
    for rtype in (abi.RENDER_SHADED_GREYSCALE, abi.RENDER_COLOUR_FROM_VOLUME, abi.RENDER_COLOUR_FROM_NORMAL, abi.RENDER_DEPTH_MAP):
        gpu.free_view(M, cfg.intr, rtype); ora.free_view(M, cfg.intr, rtype)
        a, b = ora.free_images(), gpu.free_images()
        assert np.array_equal(a.pop("visible"), b.pop("visible"))
        hz.assert_images_equal(a, b, "free view type %d" % rtype)
    assert gpu.count_visible(0, 1 << 30) == ora.count_visible(0, 1 << 30) > 0
    assert gpu.count_visible(100, 200) == ora.count_visible(100, 200)
    gpu.close(); ora.close()


def test_c1_full_size_frames():
    # BASELINE configs[0]: 640x480, 5 mm voxels, reference capacities
    cfg, frames, gpu, ora = run_sequence("C1", 3, None)
    assert gpu.rs.c.no_visible > 5000
    gpu.close(); ora.close()


def test_c2_kitti_shaped_frames():
    cfg, frames, gpu, ora = run_sequence("C2", 3, None, weights=(True, 10, 100))
    assert gpu.rs.c.no_visible > 5000
    gpu.close(); ora.close()


def test_forward_render_approximate_raycast():
    """ForwardRender: previous raycast re-projected into the next pose, holes raycast, shaded."""
    for name in ("tiny", "tiny_street"):
        cfg, frames = synth.make_sequence(name, 5)
        gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
        for f in frames[:4]:
            gpu.set_frame(f); ora.set_frame(f)
            gpu.process_frame(); ora.process_frame()
        f = frames[4]                      # new view + pose, nothing fused: ITMTrackingController::Prepare's approximate branch
        gpu.set_frame(f); ora.set_frame(f)
        gpu.expected_depths(); ora.expected_depths()
        fg, ng, rg = gpu.forward_render()
        fo, no, ro = ora.forward_render()
        assert ng == no and 0 < no < cfg.W * cfg.H
        hz.assert_images_equal(dict(fwd=fo, rendering=ro), dict(fwd=fg, rendering=rg), name + " forward render")
        assert (fo[..., 3] > 0).mean() > 0.2
        gpu.close(); ora.close()


def test_view_from_raw_and_fusion_from_it():
    """dsb200_view_from_raw: int16 depth + packed BGR on the device -> float depth + RGBA, bit-exact vs the restatement
    (itself pinned to ITMViewBuilder_CPU), odd sizes included; fusing from the built view equals fusing from the floats."""
    import importlib
    eng = importlib.import_module(hz.PKG + ".engine")
    rng = np.random.default_rng(9)
    for W, H in ((160, 120), (313, 97), (5, 3)):
        raw = rng.integers(-300, 33000, size=(H, W)).astype(np.int16)
        bgr = rng.integers(0, 256, size=(H, W, 3)).astype(np.uint8)
        v = eng.ITMView(W, H, (100.0, 100.0, W / 2.0, H / 2.0))
        v.update_from_raw(raw, bgr, 0.5, scale=np.float32(1.0 / 1000.0))
        want_d, want_c = hz.view_from_raw_restated(raw, bgr, np.float32(1.0 / 1000.0))
        assert np.array_equal(hz.bits(v.depth.download(np.float32, W * H).reshape(H, W)), hz.bits(want_d))
        assert np.array_equal(v.rgb.download(np.uint8, W * H * 4).reshape(H, W, 4), want_c)
        v.free()
    cfg, frames = synth.make_sequence("tiny", 3)
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    for f in frames:
        ora.set_frame(f); ora.process_frame()
        gpu.view.update_from_raw(f.depth_mm, f.rgb[..., 2::-1], f.timestamp, scale=np.float32(1.0 / 1000.0))
        gpu.trk.set_pose(f.M)
        gpu.process_frame()
    hz.assert_state_equal(ora.state(), gpu.state(), "fusion from raw view")
    hz.assert_images_equal(ora.images(), gpu.images(), "fusion from raw view")
    gpu.close(); ora.close()


def test_sdf_shadow_on_off_and_rebuild():
    """The fused frame raycasts from the engine's sdf shadow; turning it off (voxel-array reads) and on again (rebuilt from the
    voxels) mid-sequence must not change a bit, nor must handing the engine a scene it did not maintain the shadow for."""
    lib = abi.load()
    cfg, frames = synth.make_sequence("tiny", 8)
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    for i, f in enumerate(frames):
        if i == 2:
            abi.check(lib.dsb200_recon_set_sdf_shadow(gpu.recon.h, 0))
        if i == 4:
            abi.check(lib.dsb200_recon_set_sdf_shadow(gpu.recon.h, 1))
        if i == 6:
            abi.check(lib.dsb200_recon_invalidate_sdf_shadow(gpu.recon.h))
        gpu.set_frame(f); ora.set_frame(f)
        gpu.process_frame(); ora.process_frame()
        hz.assert_images_equal(ora.images(), gpu.images(), "shadow frame %d" % i)
    hz.assert_state_equal(ora.state(), gpu.state(), "shadow sequence")
    gpu.close(); ora.close()


def test_prefetched_allocation_matches_plain_sequence():
    """dsb200_recon_prefetch_allocation: the request pass of frame k+1 submitted right after frame k (asynchronous mode,
    views alternating between two device buffers) must leave every frame bit-identical to the oracle's plain sequence;
    calls that would invalidate a pending look-ahead are refused."""
    cfg, frames = synth.make_sequence("tiny", 10)
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    eng = gpu.eng
    views = [gpu.view, eng.ITMView(cfg.W, cfg.H, cfg.intr)]
    nxt = eng.ITMTrackingState(cfg.W, cfg.H)
    gpu.recon.SetAsync(True)
    views[0].update(frames[0].depth, frames[0].rgb, frames[0].timestamp)
    nxt.set_pose(frames[0].M)
    gpu.recon.PrefetchAllocation(gpu.scene, views[0], nxt, gpu.rs)      # right after ResetScene is fine too
    for i, f in enumerate(frames):
        v = views[i % 2]
        gpu.trk.set_pose(f.M)
        gpu.recon.ProcessFrame(gpu.vis, gpu.scene, v, gpu.trk, gpu.rs)
        if i + 1 < len(frames):
            g = frames[i + 1]
            views[(i + 1) % 2].update(g.depth, g.rgb, g.timestamp)
            nxt.set_pose(g.M)
            gpu.recon.PrefetchAllocation(gpu.scene, views[(i + 1) % 2], nxt, gpu.rs)
            if i == 3:
                with pytest.raises(abi.Dsb200Error):    # a second look-ahead
                    gpu.recon.PrefetchAllocation(gpu.scene, views[(i + 1) % 2], nxt, gpu.rs)
                with pytest.raises(abi.Dsb200Error):    # another scene call in between
                    gpu.recon.IntegrateIntoScene(gpu.scene, v, gpu.trk, gpu.rs)
        ora.set_frame(f); ora.process_frame()
        if i in (4, 7) or i + 1 == len(frames):     # frames 0..4 run without any host synchronisation
            # sync() only reads counters back; the pending look-ahead stays valid.  A DEEP look-ahead (the default) has already
            # put frame i+1's new entries into the table as pending ones: to read the table as frame i left it, drop it (the undo
            # is exact, and frame i+1 then allocates un-prefetched -- both ways of getting there are in this sequence)
            gpu.recon.Sync(gpu.scene, gpu.rs)
            abi.check(abi.load().dsb200_recon_abandon_prefetch(gpu.recon.h))
            s, so = gpu.state(), ora.state()
            if i + 1 < len(frames):
                s["visible_type"] = so["visible_type"]   # the look-ahead has already re-flagged the visibility types for frame i+1
            hz.assert_state_equal(so, s, "prefetched sequence, frame %d" % i)
            hz.assert_images_equal(ora.images(), gpu.images(), "prefetched sequence frame %d" % i)
    # a look-ahead consumed by the wrong frame is refused
    gpu.recon.PrefetchAllocation(gpu.scene, views[0], nxt, gpu.rs)
    nxt2 = eng.ITMTrackingState(cfg.W, cfg.H)
    nxt2.set_pose(frames[2].M)
    with pytest.raises(abi.Dsb200Error):
        gpu.recon.ProcessFrame(gpu.vis, gpu.scene, views[0], nxt2, gpu.rs)
    gpu.recon.SetAsync(False)
    views[1].free(); nxt.free(); nxt2.free()
    gpu.close(); ora.close()


def test_device_expf_equals_host_c_library():
    """the filter's exp restates glibc's expf: bit-identical on a few million arguments over the whole range it can see"""
    import ctypes
    lib = abi.load()
    libm = ctypes.CDLL("libm.so.6")
    libm.expf.restype, libm.expf.argtypes = ctypes.c_float, [ctypes.c_float]
    rng = np.random.default_rng(17)
    x = np.concatenate([-rng.random(1 << 21) * 110.0, -np.exp(rng.random(1 << 20) * 12 - 8), rng.random(1 << 18) * 2.0,
                        [0.0, -0.0, -87.9, -88.0, -88.1, -103.2, -103.3, -103.9, -103.98, -104.0, -1e30, -np.inf, 88.7, 89.0, np.nan]]).astype(np.float32)
    y = np.zeros_like(x)
    abi.check(lib.dsb200_selftest_expf(x.ctypes.data, y.ctypes.data, x.size))
    # the C library, one call per sample of a 200k subset (ctypes is slow), plus every special value
    idx = np.concatenate([rng.integers(0, x.size - 15, 200000), np.arange(x.size - 15, x.size)])
    want = np.array([libm.expf(float(v)) for v in x[idx]], np.float32)
    got = y[idx]
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    bad = np.nonzero(hz.bits(got[~nan]) != hz.bits(want[~nan]))[0]
    assert bad.size == 0, "%d of %d differ, first x=%r: %r vs %r" % (bad.size, idx.size, x[idx][~nan][bad[:3]], got[~nan][bad[:3]], want[~nan][bad[:3]])


def test_bilateral_depth_filter_vs_oracle():
    """dsb200_view_filter_depth (UpdateView's useBilateralFilter branch).  Float arithmetic is the reference's; its exp() is
    evaluated in double and rounded to float on the GPU, glibc expf on the host: the two agree except where the exact value
    sits on a float rounding boundary, so this test asks for >= 99.5 % bit-identical pixels (one stray weight spreads over 21x21 pixels in five passes) and 1e-6 relative elsewhere."""
    eng = hz.importlib.import_module(hz.PKG + ".engine")
    o = hz.po.Oracle()
    rng = np.random.default_rng(5)
    for name, n in (("tiny", 1), ("C1", 1), ("tiny_street", 1)):
        cfg, frames = synth.make_sequence(name, n)
        f = frames[-1]
        raw = f.depth_mm.copy()
        raw.reshape(-1)[rng.integers(0, raw.size, 200)] = 0                      # holes
        view = eng.ITMView(cfg.W, cfg.H, cfg.intr)
        view.update_from_raw(raw, f.rgb[..., 2::-1], f.timestamp, np.float32(1.0 / 1000.0), 0.0, useBilateralFilter=True)
        got = view.depth.download(np.float32, cfg.W * cfg.H).reshape(cfg.H, cfg.W)
        plain, _ = hz.view_from_raw_restated(raw, None, np.float32(1.0 / 1000.0), 0.0)
        want = o.view_filter_depth(plain.reshape(cfg.H, cfg.W))
        same = hz.bits(got) == hz.bits(want)
        assert same.mean() >= 0.995, "%s: only %.5f of the pixels bit-identical" % (name, same.mean())
        assert np.array_equal(got < 0, want < 0) and np.array_equal(got == 0, want == 0)
        rel = np.abs(got - want)[~same] / np.abs(want[~same]) if (~same).any() else np.zeros(1)
        assert rel.max() <= 1e-6
        assert (want[2:-2, 2:-2] > 0).mean() > 0.2 and not np.array_equal(want, plain.reshape(cfg.H, cfg.W))
        view.free()
    # the single-pass entry point (what an ITMViewBuilder subclass forwards DepthFiltering to) composes to the same result
    d = plain.reshape(cfg.H, cfg.W)
    for _ in range(5):
        d = eng.depth_filter_pass(d)
    assert np.array_equal(hz.bits(d), hz.bits(got))
    # Kinect-style disparity conversion: bit-exact (one IEEE division per pixel)
    disp = rng.integers(-200, 1200, (97, 131)).astype(np.int16)
    disp[0, :8] = 1090
    a = eng.convert_disparity_to_depth(disp, 1090.0, 0.075, 573.71)
    b = o.convert_disparity_to_depth(disp, 1090.0, 0.075, 573.71)
    assert np.array_equal(hz.bits(a), hz.bits(b)) and (a == -1.0).any() and (a > 0).any()


def test_device_acosf_equals_host_c_library():
    """the sensor-noise model's acos restates glibc's (fdlibm) acosf: bit-identical on every branch, special values included"""
    import ctypes
    lib = abi.load()
    libm = ctypes.CDLL("libm.so.6")
    libm.acosf.restype, libm.acosf.argtypes = ctypes.c_float, [ctypes.c_float]
    rng = np.random.default_rng(23)
    special = [0.0, -0.0, 1.0, -1.0, 0.5, -0.5, np.nextafter(np.float32(0.5), np.float32(0)), np.nextafter(np.float32(1), np.float32(0)),
               -np.nextafter(np.float32(1), np.float32(0)), 2.0 ** -26, 2.0 ** -27, -(2.0 ** -26), 1e-30, 1.0000001, -3.0, np.inf, np.nan]
    x = np.concatenate([rng.uniform(-1, 1, 1 << 21), 1.0 - np.exp(-rng.random(1 << 19) * 16), np.exp(-rng.random(1 << 19) * 30) - 1.0,
                        rng.uniform(-1, 1, 1 << 18) * 1e-6, special]).astype(np.float32)
    y = np.zeros_like(x)
    abi.check(lib.dsb200_selftest_acosf(x.ctypes.data, y.ctypes.data, x.size))
    idx = np.concatenate([rng.integers(0, x.size - len(special), 200000), np.arange(x.size - len(special), x.size)])
    want = np.array([libm.acosf(float(v)) for v in x[idx]], np.float32)
    got = y[idx]
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan) and nan.sum() >= 4
    bad = np.nonzero(hz.bits(got[~nan]) != hz.bits(want[~nan]))[0]
    assert bad.size == 0, "%d of %d differ, first x=%r: %r vs %r" % (bad.size, idx.size, x[idx][~nan][bad[:3]], got[~nan][bad[:3]], want[~nan][bad[:3]])


def test_normal_and_weights_vs_oracle():
    """dsb200_compute_normal_and_weights (UpdateView's modelSensorNoise step) == oracle bit for bit, untouched elements included"""
    eng = hz.importlib.import_module(hz.PKG + ".engine")
    o = hz.po.Oracle()
    valid = 0
    for name, depth, intr in hz.normal_and_weight_cases():
        H, W = depth.shape
        want_n = np.full((H, W, 4), 7.25, np.float32); want_s = np.full((H, W), -3.5, np.float32)
        got_n, got_s = eng.compute_normal_and_weights(depth, intr, want_n, want_s)
        o.compute_normal_and_weights(depth, intr, want_n, want_s)
        assert np.array_equal(hz.bits(got_n), hz.bits(want_n)), name
        assert np.array_equal(hz.bits(got_s), hz.bits(want_s)), name
        valid += int((got_n[..., 3] == 1.0).sum())
    assert valid > 5000
    # full-size frame (C1, 640 x 480) and an image with no interior
    cfg, frames = synth.make_sequence("C1", 1)
    depth = frames[0].depth.reshape(cfg.H, cfg.W)
    want_n = np.zeros((cfg.H, cfg.W, 4), np.float32); want_s = np.zeros((cfg.H, cfg.W), np.float32)
    got_n, got_s = eng.compute_normal_and_weights(depth, cfg.intr, want_n, want_s)
    o.compute_normal_and_weights(depth, cfg.intr, want_n, want_s)
    assert np.array_equal(hz.bits(got_n), hz.bits(want_n)) and np.array_equal(hz.bits(got_s), hz.bits(want_s))
    small_n, small_s = eng.compute_normal_and_weights(np.ones((4, 9), np.float32), cfg.intr, np.zeros((4, 9, 4), np.float32), np.zeros((4, 9), np.float32))
    assert not small_n.any() and not small_s.any()


def ragged_sequence():
    """157 x 119 (no multiple of the 4-pixel vectors, 8x4 ray tiles, 8-pixel range-image cells or 16x16 rendering blocks) with
    empty frames: frame 0 (nothing to allocate, empty visible list), frame 3 (every visible block falls back to the frustum
    re-test) and a single valid pixel in frame 5."""
    cfg = synth.Config("odd", 157, 119, (118.3, 118.0, 77.7, 58.9), 0.02, 0.08, 100, 0.2, 3.0, "room")
    cfg, frames = synth.make_sequence(cfg, 7)
    for i in (0, 3, 5):
        keep = frames[i].depth.reshape(cfg.H, cfg.W)[60, 80]
        frames[i].depth = np.full_like(frames[i].depth, -1.0)
        frames[i].depth_mm = np.zeros_like(frames[i].depth_mm)
        if i == 5:
            frames[i].depth.reshape(cfg.H, cfg.W)[60, 80] = keep
    return cfg, frames


@pytest.mark.parametrize("fused", [True, False])
def test_ragged_sizes_and_empty_frames(fused):
    cfg, frames = ragged_sequence()
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    for i, f in enumerate(frames):
        gpu.set_frame(f); ora.set_frame(f)
        if fused:
            assert gpu.process_frame() == ora.process_frame()
        else:
            assert gpu.allocate() == ora.allocate()
            gpu.integrate(); ora.integrate()
            gpu.expected_depths(); ora.expected_depths()
            gpu.icp_maps(); ora.icp_maps()
        hz.assert_state_equal(ora.state(), gpu.state(), "ragged frame %d" % i)
        hz.assert_images_equal(ora.images(), gpu.images(), "ragged frame %d" % i)
        if i == 0:
            assert gpu.rs.c.no_visible == 0
    assert gpu.rs.c.no_visible > 50
    # free-view render of the ragged image size
    for b in (gpu, ora):
        b.free_view(frames[4].M, cfg.intr, abi.RENDER_SHADED_GREYSCALE)
    fg, fo = gpu.free_images(), ora.free_images()
    assert np.array_equal(fg["visible"], fo["visible"])
    hz.assert_images_equal(dict(range=fo["range"], raycast=fo["raycast"], image=fo["image"]),
                           dict(range=fg["range"], raycast=fg["raycast"], image=fg["image"]), "ragged free view")
    gpu.close(); ora.close()


def test_checkpoint_round_trip(tmp_path):
    """dsb200_scene_save / dsb200_scene_load: a fresh engine + scene that loads the checkpoint continues the sequence exactly
    like the engine that wrote it (and like the oracle's uninterrupted run); files follow MemoryBlockPersister's layout."""
    import os
    cfg, frames = synth.make_sequence("tiny", 8)
    a, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    for f in frames[:4]:
        a.set_frame(f); ora.set_frame(f)
        assert a.process_frame() == ora.process_frame()
    a.recon.SaveCheckpoint(a.scene, a.rs, tmp_path)
    total = hz.SMALL_CAPS["bucket_num"] + hz.SMALL_CAPS["excess_size"]
    raw = open(os.path.join(str(tmp_path), "hash.dat"), "rb").read()
    assert int(np.frombuffer(raw[:4], np.int32)[0]) == total and len(raw) == 4 + 20 * total
    assert os.path.getsize(os.path.join(str(tmp_path), "voxel.dat")) == 4 + 8 * 512 * hz.SMALL_CAPS["num_blocks"]
    b = hz.GpuBackend(cfg, hz.SMALL_CAPS)
    b.recon.LoadCheckpoint(b.scene, b.rs, tmp_path)
    hz.assert_state_equal(ora.state(), b.state(), "state right after loading")
    for i, f in enumerate(frames[4:]):
        for x in (a, b, ora):
            x.set_frame(f)
        ra, rb, ro = a.process_frame(), b.process_frame(), ora.process_frame()
        assert ra == rb == ro
        hz.assert_state_equal(ora.state(), b.state(), "resumed frame %d" % i)
        hz.assert_images_equal(ora.images(), b.images(), "resumed frame %d" % i)
        hz.assert_state_equal(a.state(), b.state(), "resumed vs uninterrupted frame %d" % i)
    # a scene with other capacities refuses the checkpoint
    other = dict(hz.SMALL_CAPS); other["excess_size"] //= 2
    c = hz.GpuBackend(cfg, other)
    with pytest.raises(abi.Dsb200Error):
        c.recon.LoadCheckpoint(c.scene, c.rs, tmp_path)
    for x in (a, b, c, ora):
        x.close()


def test_abandoned_lookahead_is_undone(tmp_path):
    """A look-ahead that is dropped -- dsb200_recon_notify_scene_modified, or the next allocation arriving with another view /
    pose -- must leave no trace: its request keys, request bitmaps and visibility flags are undone, and the frame allocated
    next equals the oracle's plain sequence bit for bit (stale order keys could win the next allocation)."""
    cfg, frames = synth.make_sequence("tiny", 6)
    _, far = synth.make_sequence("tiny", 2, start=45)          # a view of another part of the room: many requests of its own
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    eng, lib = gpu.eng, abi.load()
    other = eng.ITMView(cfg.W, cfg.H, cfg.intr)
    nxt = eng.ITMTrackingState(cfg.W, cfg.H)
    gpu.recon.SetAsync(True)
    for i in range(3):
        gpu.set_frame(frames[i]); ora.set_frame(frames[i])
        gpu.process_frame(); ora.process_frame()
    # (1) look ahead to the far view, then tell the engine the scene was modified: the look-ahead is dropped
    other.update(far[0].depth, far[0].rgb, far[0].timestamp); nxt.set_pose(far[0].M)
    gpu.recon.PrefetchAllocation(gpu.scene, other, nxt, gpu.rs)
    gpu.recon.Sync(gpu.scene, gpu.rs)          # the host mirrors become authoritative again after the notification
    # the DEEP look-ahead has published the far view's new blocks as pending entries (ptr <= -3: unallocated to every reader);
    # the counters and the visible list are still frame 2's, and a checkpoint of this table is refused
    import ctypes as C
    abi.check(lib.dsb200_recon_join_prefetch(gpu.recon.h)); abi.check(lib.dsb200_stream_synchronize(None))
    so = ora.state()
    assert (gpu.scene.hash()["ptr"] <= -3).sum() > 20
    assert gpu.scene.c.last_free_block == so["last_free"][0] and np.array_equal(gpu.rs.visible(), so["visible"])
    assert lib.dsb200_scene_save(C.byref(gpu.scene.c), None, str(tmp_path).encode()) == abi.ERR_INVALID
    abi.check(lib.dsb200_recon_notify_scene_modified(gpu.recon.h))
    abi.check(lib.dsb200_stream_synchronize(None))
    assert (gpu.scene.hash()["ptr"] <= -3).sum() == 0
    s = gpu.state(); s["visible_type"] = so["visible_type"]
    hz.assert_state_equal(so, s, "table after the undo")
    gpu.set_frame(frames[3]); ora.set_frame(frames[3])
    assert gpu.process_frame() == 0
    ora.process_frame()
    gpu.recon.Sync(gpu.scene, gpu.rs)
    hz.assert_state_equal(ora.state(), gpu.state(), "after a look-ahead dropped by notify_scene_modified")
    hz.assert_images_equal(ora.images(), gpu.images(), "after a look-ahead dropped by notify_scene_modified")
    # (2) look ahead to the far view, then submit a different frame: refused once (and undone), then accepted
    other.update(far[1].depth, far[1].rgb, far[1].timestamp); nxt.set_pose(far[1].M)
    gpu.recon.PrefetchAllocation(gpu.scene, other, nxt, gpu.rs)
    gpu.set_frame(frames[4]); ora.set_frame(frames[4])
    with pytest.raises(abi.Dsb200Error):
        gpu.recon.ProcessFrame(gpu.vis, gpu.scene, gpu.view, gpu.trk, gpu.rs)
    assert gpu.process_frame() == 0
    ora.process_frame()
    gpu.recon.Sync(gpu.scene, gpu.rs)
    hz.assert_state_equal(ora.state(), gpu.state(), "after a look-ahead dropped by a mismatching allocation")
    hz.assert_images_equal(ora.images(), gpu.images(), "after a look-ahead dropped by a mismatching allocation")
    gpu.set_frame(frames[5]); ora.set_frame(frames[5])
    gpu.process_frame(); ora.process_frame()
    gpu.recon.Sync(gpu.scene, gpu.rs)
    hz.assert_state_equal(ora.state(), gpu.state(), "frame after the dropped look-aheads")
    gpu.recon.SetAsync(False)
    other.free(); nxt.free()
    gpu.close(); ora.close()


def test_two_pass_raycast_is_bit_identical():
    """the bounded first raycast pass + the group-speculative finishing pass (csrc/vis.cu: k_raycast_finish; off by default)
    produce the single-pass image bit for bit, for every group size, on the fused frame and on the free-view path"""
    lib = abi.load()
    cfg, frames = synth.make_sequence("tiny_street", 5)
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    gpu.set_weights(True); ora.set_weights(True)
    try:
        for i, f in enumerate(frames):
            abi.check(lib.dsb200_debug_set_tunable(b"ray_steps", (0, 3, 6, 10, 16)[i]))
            abi.check(lib.dsb200_debug_set_tunable(b"ray_group", (8, 2, 4, 16, 32)[i]))
            gpu.set_frame(f); ora.set_frame(f)
            gpu.process_frame(); ora.process_frame()
            hz.assert_images_equal(ora.images(), gpu.images(), "two-pass raycast, frame %d" % i)
        abi.check(lib.dsb200_debug_set_tunable(b"ray_steps", 4))
        gpu.free_view(frames[2].M, cfg.intr, 0); ora.free_view(frames[2].M, cfg.intr, 0)
        hz.assert_images_equal(ora.free_images(), gpu.free_images(), "two-pass raycast, free view")
    finally:
        abi.check(lib.dsb200_debug_set_tunable(b"ray_steps", 0))
        abi.check(lib.dsb200_debug_set_tunable(b"ray_group", 8))
    gpu.close(); ora.close()
