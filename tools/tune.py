#!/usr/bin/env python
"""tools/tune.py -- sweep launch-shape knobs (csrc/kernels.cuh: Tunables) of ONE stage on the GPU, on a map built from the
bench workload.  Every setting must produce bit-identical output (checked here against the first setting); the point is only
the time.  Prints one JSON line per setting and a summary.

  python tools/tune.py --stage raycast --workload C2 --frames 40 --reps 20 \
      --sweep ray_steps=0,16,24,32,48 --sweep ray_group=8,16,32

stage raycast:   times dsb200_process_frame(stages = RAYCAST) at the last fused frame's pose (sdf shadow, expected depths valid);
                 COLD: the L2 flush before it evicts the blocks integration has just written, unlike a real frame -- use `frame`
stage integrate: times dsb200_process_frame(stages = INTEGRATE) of the last frame again and again (same visible list, same
                 voxels touched every time; only the weights grow)
stage frame:     times whole frames k..k+reps from a fixed prefix (scene reset + prefix re-fused untimed for each setting)
"""
import argparse
import ctypes as C
import importlib
import itertools
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
PKG = "denseslam-global-consistency-h_b200"
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stage", default="raycast", choices=["raycast", "integrate", "frame"])
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--frames", type=int, default=40)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--sweep", action="append", default=[], help="name=v1,v2,...")
    args = ap.parse_args()
    import torch
    abi = importlib.import_module(PKG + ".abi")
    eng = importlib.import_module(PKG + ".engine")
    synth = importlib.import_module(PKG + ".synth")
    lib = abi.load()
    abi.check(lib.dsb200_set_device(0))
    dev = torch.device("cuda", 0)
    caps = bench.WORKLOADS[args.workload]["caps"]
    n_frames = args.frames + (args.reps if args.stage == "frame" else 0)
    cfg, frames = synth.make_sequence(args.workload, n_frames + 1)
    W, H = cfg.W, cfg.H
    scene = eng.ITMScene(cfg.params(), caps["num_blocks"], caps["bucket_num"], caps["excess_size"])
    recon = eng.ITMSceneReconstructionEngine(caps["num_blocks"], caps["bucket_num"], caps["excess_size"])
    recon.SetFusionWeightParams(cfg.depth_weighting, 10, 100)
    vis = eng.ITMVisualisationEngine(W, H)
    rs = vis.CreateRenderState(scene, (W, H))
    trk = eng.ITMTrackingState(W, H)
    views = []
    for f in frames:
        v = eng.ITMView(W, H, cfg.intr)
        v.update(f.depth, f.rgb, f.timestamp)
        views.append(v)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    zeros_vt = np.zeros(scene.c.total_entries, np.uint8)

    def build(upto):
        recon.SetAsync(False)
        recon.ResetScene(scene)
        rs.c.no_visible = 0
        abi.check(lib.dsb200_memcpy_h2d(rs.c.visible_type, zeros_vt.ctypes.data, scene.c.total_entries, None))
        abi.check(lib.dsb200_stream_synchronize(None))
        recon.SetAsync(True)
        for i in range(upto):
            trk.set_pose(frames[i].M)
            recon.ProcessFrame(vis, scene, views[i], trk, rs)
        recon.Sync(scene, rs)

    names = [s.split("=")[0] for s in args.sweep]
    values = [[int(x) for x in s.split("=")[1].split(",")] for s in args.sweep]
    combos = list(itertools.product(*values)) if values else [()]
    build(args.frames)
    last = args.frames - 1
    ref_out, results = None, []
    for combo in combos:
        for n, v in zip(names, combo):
            abi.check(lib.dsb200_debug_set_tunable(n.encode(), v))
        times = []
        if args.stage == "frame":
            build(args.frames)
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.reps)]
        for r in range(args.reps + 2):
            timed = r >= 2
            torch.cuda._sleep(600000)       # ~0.3 ms of GPU idling that the host uses to enqueue this repetition: no launch gaps inside the event pair
            flush.zero_()
            if timed:
                evs[r - 2][0].record()
            if args.stage == "raycast":
                trk.set_pose(frames[last].M)
                recon.ProcessFrame(vis, scene, views[last], trk, rs, abi.FRAME_RAYCAST)
            elif args.stage == "integrate":
                trk.set_pose(frames[last].M)
                recon.ProcessFrame(vis, scene, views[last], trk, rs, abi.FRAME_INTEGRATE)
            else:
                i = args.frames + max(r - 2, 0) if timed else last
                if timed:
                    trk.set_pose(frames[i].M)
                    recon.ProcessFrame(vis, scene, views[i], trk, rs)
                else:
                    trk.set_pose(frames[last].M)
                    recon.ProcessFrame(vis, scene, views[last], trk, rs, abi.FRAME_RAYCAST)
            if timed:
                evs[r - 2][1].record()
        torch.cuda.synchronize()
        recon.Sync(scene, rs)
        times = np.array([a.elapsed_time(b) for a, b in evs])
        out = rs.raycast_result() if args.stage != "integrate" else None
        same = None
        if out is not None:
            if ref_out is None:
                ref_out = out
            same = bool(np.array_equal(out.view(np.uint32), ref_out.view(np.uint32)))
        res = dict(zip(names, combo))
        res.update(stage=args.stage, ms_median=float(np.median(times)), ms_min=float(times.min()), ms_mean=float(times.mean()),
                   identical_to_first=same, visible=int(rs.c.no_visible))
        results.append(res)
        print(json.dumps(res), flush=True)
    best = min(results, key=lambda r: r["ms_median"])
    print(json.dumps({"best": best}))


if __name__ == "__main__":
    main()
