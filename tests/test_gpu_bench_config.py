"""The configuration bench.py TIMES, parity-checked: same capacities (bench.WORKLOADS), asynchronous mode, the allocation
look-ahead and the sdf shadow on, the same frames (warm-up + timed), compared with the CPU checker run live on the same frames
at several checkpoints (state digests of denseslam-global-consistency-h_b200/digest.py: hash, free lists, visible list, every
allocated voxel block, the frame's images) and, at the end, with the committed golden digest bench.py itself checks
(tests/golden/bench_digests.json)."""
import importlib
import json
import os

import numpy as np
import pytest

import harness as hz
from harness import abi, synth

pytestmark = pytest.mark.gpu
digest = importlib.import_module(hz.PKG + ".digest")
import bench  # noqa: E402


def oracle_digest(ora):
    s = ora.state(voxels=False)
    vox = ora.scene.voxels
    d = digest.state_digest(s["hash"], s["last_free"], s["visible"], lambda lo, hi: vox[lo * 512:hi * 512], ora.scene.c.num_blocks, ora.images())
    d["no_visible"] = int(len(s["visible"]))
    return d


def run_like_bench(workload, n_frames, checkpoints):
    spec = bench.WORKLOADS[workload]
    caps, prune = spec["caps"], spec["prune"]
    cfg, frames = synth.make_sequence(workload, n_frames + 1)
    gpu, ora = hz.GpuBackend(cfg, caps), hz.OracleBackend(cfg, caps)
    gpu.set_weights(cfg.depth_weighting, 10, 100); ora.set_weights(cfg.depth_weighting, 10, 100)
    eng, lib = gpu.eng, abi.load()
    views = []
    for f in frames:                      # frames resident in HBM, as in bench.py's first leg
        v = eng.ITMView(cfg.W, cfg.H, cfg.intr)
        v.update(f.depth, f.rgb, f.timestamp)
        views.append(v)
    nxt = eng.ITMTrackingState(cfg.W, cfg.H)
    gpu.recon.SetAsync(True)
    for i in range(n_frames):
        f = frames[i]
        gpu.trk.set_pose(f.M)
        gpu.recon.ProcessFrame(gpu.vis, gpu.scene, views[i], gpu.trk, gpu.rs)
        if prune:
            gpu.recon.SlideWindow(gpu.scene, gpu.rs, prune["max_age"])
            gpu.recon.Decay(gpu.scene, gpu.rs, prune["max_weight"], prune["min_age"])
        nxt.set_pose(frames[i + 1].M)
        gpu.recon.PrefetchAllocation(gpu.scene, views[i + 1], nxt, gpu.rs)
        ora.set_frame(f)
        assert ora.process_frame() == 0
        if prune:
            ora.slide_window(prune["max_age"]); ora.decay(prune["max_weight"], prune["min_age"])
        if (i + 1) in checkpoints:
            gpu.recon.Sync(gpu.scene, gpu.rs)
            abi.check(lib.dsb200_recon_abandon_prefetch(gpu.recon.h))     # the deep look-ahead's pending entries of frame i+1 are undone (exactly)
            abi.check(lib.dsb200_stream_synchronize(None))
            got = bench.state_digest_of(digest, eng, gpu.scene, gpu.rs, gpu.trk, gpu.recon)
            want = oracle_digest(ora)
            assert digest.compare(got, want) == "ok", "%s after %d frames: %s (%d vs %d visible, %d vs %d blocks)" % (
                workload, i + 1, digest.compare(got, want), got["no_visible"], want["no_visible"], got["blocks_allocated"], want["blocks_allocated"])
            last = got
    if prune:
        assert gpu.decayed_blocks() == ora.decayed_blocks() and (n_frames <= prune["min_age"] or gpu.decayed_blocks() > 0)
    gpu.recon.SetAsync(False)
    for v in views:
        v.free()
    nxt.free()
    gpu.close(); ora.close()
    return last


def golden(key):
    table = json.load(open(bench.DIGESTS))
    return table.get(key)


def test_c2_bench_configuration_25_frames():
    """BASELINE configs[1] exactly as bench.py --steps 20 --warmup 5 runs it (2 Mi blocks / 4 Mi buckets, async + look-ahead +
    shadow), checked every 5th frame"""
    last = run_like_bench("C2", 25, {5, 10, 15, 20, 25})
    g = golden("C2:25")
    assert g is not None and digest.compare(last, g) == "ok"


def test_c1_bench_configuration_25_frames():
    last = run_like_bench("C1", 25, {5, 15, 25})
    g = golden("C1:25")
    assert g is not None and digest.compare(last, g) == "ok"


def test_c3_bench_configuration_full_size_105_frames():
    """BASELINE configs[2] at 640x480: SlideWindow(max_age 80) + Decay(max_weight 3, min_age 40) after every frame for 105 frames
    (pruning starts at frame 41, the window slides from frame 81)"""
    last = run_like_bench("C3", 105, {40, 45, 85, 105})
    g = golden("C3:105")
    assert g is not None and digest.compare(last, g) == "ok"
