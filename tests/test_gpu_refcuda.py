"""libdsb200 vs the UNMODIFIED reference CUDA engines recompiled for sm_100 (oracle/_ref/libitmref_cuda.so, IEEE build;
oracle/refcuda_shim.cu, oracle/Makefile).  This is the pin for the entry points the reference implements ONLY in CUDA:
DeIntegrateIntoScene (CUDA .cu:426-489), Decay / DecayDefusionPart (.cu:769-802, 880-999), SlideWindow /
SlideWindowDefusionPart (.cu:816-877, 1001-1077), GetDecayedBlockCount (.cu:1080) and the history queues (.cu:257-317).

The reference allocator is racy (atomicSub free-list pops, last-writer-wins requests, per-CTA atomicAdd compaction), so slot and
pointer assignment differ by schedule and the comparison is at SET level (harness.assert_sets_match): the set of allocated block
coordinates, the 512 voxels of every block both engines allocated in the same frame (bit for bit: both sides compute in IEEE
float without contraction), the set of visible blocks, free-list accounting, decayed-block counts and queue sizes.  Blocks born
a frame apart (same-frame bucket collisions) or kept by the reference's try-lock skip are counted and bounded, not compared."""
import numpy as np
import pytest

import harness as hz
from harness import abi, synth

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not hz.has_refcuda(), reason="oracle/_ref/libitmref_cuda.so not built")]

REF_CAPS = dict(num_blocks=abi.DEFAULT_LOCAL_BLOCKS, bucket_num=abi.DEFAULT_BUCKETS, excess_size=abi.DEFAULT_EXCESS)
# ours with a smaller voxel array (the block SET does not depend on it) but the reference's table: same buckets, same collisions
CAPS = dict(num_blocks=0x8000, bucket_num=abi.DEFAULT_BUCKETS, excess_size=abi.DEFAULT_EXCESS)


def pair(cfg, caps=CAPS, weights=False, use_swapping=False):
    gpu, ref = hz.GpuBackend(cfg, caps, use_swapping=use_swapping), hz.RefCudaBackend(cfg, use_swapping=use_swapping)
    if weights:
        gpu.set_weights(True, 10, 100); ref.set_weights(True, 10, 100)
    return gpu, ref


def step(backs, f, decay=None, slide=None):
    # DenseSlam::ProcessFrame order: fuse, SlideWindow, Decay (DS/DenseSlam.cpp:213-228)
    for b in backs:
        b.set_frame(f)
        assert b.allocate() == 0
        b.integrate()
        if slide is not None:
            b.slide_window(slide)
        if decay is not None:
            b.decay(*decay)


def compare(tr, gpu, ref, what, voxels=True, **kw):
    return hz.assert_sets_match(tr, hz.gpu_block_table(gpu, voxels), ref.block_table(voxels), what, voxels=voxels, **kw)


def assert_raycast_close(a, b, what, min_agree=0.999, rel=1e-3, min_close=0.995):
    """hit-mask agreement and relative depth of the common hits (>= 99.9 % / 1e-3)"""
    ha, hb = a["raycast"][..., 3] > 0, b["raycast"][..., 3] > 0
    agree = float((ha == hb).mean())
    assert agree >= min_agree, "%s: hit masks agree on %.4f of the pixels" % (what, agree)
    m = ha & hb
    d = np.linalg.norm(a["raycast"][..., :3][m] - b["raycast"][..., :3][m], axis=-1)
    scale = np.maximum(np.linalg.norm(a["raycast"][..., :3][m], axis=-1), 1.0)
    close = float((d <= rel * scale).mean()) if m.any() else 1.0
    assert close >= min_close, "%s: %.4f of the common hits within %g relative" % (what, close, rel)
    return agree, close


def test_fusion_sequence_matches_reference_cuda():
    """allocate + integrate + expected depths + raycast/ICP maps, room and street (depth weighting on, which the reference
    applies only in its CUDA integrator)"""
    for name, weights in (("tiny", False), ("tiny_street", True)):
        cfg, frames = synth.make_sequence(name, 8)
        gpu, ref = pair(cfg, weights=weights)
        tr = hz.SetTracker()
        for i, f in enumerate(frames):
            for b in (gpu, ref):
                b.set_frame(f)
                assert b.process_frame() == 0
            compare(tr, gpu, ref, "%s frame %d" % (name, i))
            gi, ri = gpu.images(), ref.images()
            assert_raycast_close(gi, ri, "%s frame %d" % (name, i))
        gpu.close(); ref.close()


def test_decay_and_slide_window_pinned_to_reference_cuda():
    cfg, frames = synth.make_sequence("tiny", 30)
    gpu, ref = pair(cfg)
    tr = hz.SetTracker()
    odd = 0
    for i, f in enumerate(frames):
        step((gpu, ref), f, decay=(3, 4), slide=9)
        odd = compare(tr, gpu, ref, "frame %d" % i)
        assert gpu.history_sizes() == ref.history_sizes(), "frame %d: history queues %s vs %s" % (i, gpu.history_sizes(), ref.history_sizes())
        assert abs(gpu.decayed_blocks() - ref.decayed_blocks()) <= max(8, 0.03 * gpu.decayed_blocks()), "frame %d: decayed %d vs %d" % (
            i, gpu.decayed_blocks(), ref.decayed_blocks())
    assert gpu.decayed_blocks() > 500
    print("decay+slide: %d blocks freed (reference %d), %d non-comparable at the end" % (gpu.decayed_blocks(), ref.decayed_blocks(), odd))
    gpu.close(); ref.close()


def test_decay_only_pinned_to_reference_cuda():
    cfg, frames = synth.make_sequence("tiny_street", 20)
    gpu, ref = pair(cfg, weights=True)
    tr = hz.SetTracker()
    for i, f in enumerate(frames):
        step((gpu, ref), f, decay=(2, 3))
        compare(tr, gpu, ref, "frame %d" % i)
        assert gpu.history_sizes() == ref.history_sizes()
    assert gpu.decayed_blocks() > 20
    assert abs(gpu.decayed_blocks() - ref.decayed_blocks()) <= max(8, 0.03 * gpu.decayed_blocks())
    gpu.close(); ref.close()


def test_slide_window_only_pinned_to_reference_cuda():
    cfg, frames = synth.make_sequence("tiny", 16)
    gpu, ref = pair(cfg)
    tr = hz.SetTracker()
    for i, f in enumerate(frames):
        step((gpu, ref), f, slide=5)
        compare(tr, gpu, ref, "frame %d" % i)
        assert gpu.history_sizes() == ref.history_sizes()
    assert gpu.decayed_blocks() > 500
    assert abs(gpu.decayed_blocks() - ref.decayed_blocks()) <= max(8, 0.03 * gpu.decayed_blocks())
    gpu.close(); ref.close()


def test_deintegrate_reintegrate_pinned_to_reference_cuda():
    """online correction (DS/DenseSlam.cpp:298-432): de-fuse a keyframe with its old pose, re-fuse with the corrected one,
    then the de-fusion parts of the pruning passes"""
    cfg, frames = synth.make_sequence("tiny", 6)
    gpu, ref = pair(cfg)
    tr = hz.SetTracker()
    for i, f in enumerate(frames):
        step((gpu, ref), f)
        compare(tr, gpu, ref, "frame %d" % i)
    k, newM = frames[2], frames[3].M
    for b in (gpu, ref):
        b.set_frame(k); b.deintegrate()
    compare(tr, gpu, ref, "after de-integration")
    for b in (gpu, ref):
        b.set_frame(k, M=newM)
        assert b.allocate(is_defusion=True) == 0
        b.integrate()
    compare(tr, gpu, ref, "after re-integration")
    assert gpu.history_sizes() == ref.history_sizes()
    for b in (gpu, ref):
        b.slide_window_defusion(80, 0); b.decay_defusion(2, 0)
    compare(tr, gpu, ref, "after the de-fusion parts of slide window and decay")
    assert gpu.history_sizes() == ref.history_sizes()
    assert abs(gpu.decayed_blocks() - ref.decayed_blocks()) <= 8
    # a timestamp never fused is a silent no-op in both (CUDA .cu:437-440)
    ghost = synth.make_sequence("tiny", 1, start=40)[1][0]
    before = hz.gpu_block_table(gpu)
    for b in (gpu, ref):
        b.set_frame(ghost); b.deintegrate()
    after = hz.gpu_block_table(gpu)
    assert np.array_equal(before["voxels"].view(np.uint64), after["voxels"].view(np.uint64))
    compare(tr, gpu, ref, "after de-integrating an unknown timestamp")
    gpu.close(); ref.close()


def test_integrate_deintegrate_round_trip_on_reference_cuda():
    """the property tests/test_gpu_prune.py checks on ours holds on the reference too: fuse one frame, de-fuse it with the
    same pose -> every depth weight back to 0, every sdf back to 32767"""
    cfg, frames = synth.make_sequence("tiny", 1)
    gpu, ref = pair(cfg)
    for b in (gpu, ref):
        b.set_frame(frames[0]); b.allocate(); b.integrate(); b.deintegrate()
    tr = hz.SetTracker()
    compare(tr, gpu, ref, "after fuse + de-fuse")
    v = ref.block_table()["voxels"]
    assert (v["w_depth"] == 0).all() and (v["sdf"] == 32767).all()
    gpu.close(); ref.close()


def test_c3_full_size_decay_slide_window_vs_reference_cuda():
    """BASELINE configs[2] at full size (640x480, 5 mm voxels, max_decay_weight 3, min_decay_age 40, max_age 80 as bench.py
    --workload C3 runs it) for 100 frames: block sets / free lists / queues every frame, voxel payloads at four checkpoints."""
    cfg, frames = synth.make_sequence("C3", 100)
    gpu, ref = pair(cfg, caps=REF_CAPS)
    tr = hz.SetTracker()
    for i, f in enumerate(frames):
        step((gpu, ref), f, decay=(3, 40), slide=80)
        full = i in (20, 60, 85, 99)
        compare(tr, gpu, ref, "C3 frame %d" % i, voxels=full)
        assert gpu.history_sizes() == ref.history_sizes()
    assert gpu.decayed_blocks() > 1000
    assert abs(gpu.decayed_blocks() - ref.decayed_blocks()) <= max(8, 0.03 * gpu.decayed_blocks()), "decayed %d vs %d" % (
        gpu.decayed_blocks(), ref.decayed_blocks())
    gpu.close(); ref.close()
