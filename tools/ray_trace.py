#!/usr/bin/env python
"""tools/ray_trace.py -- per-tile timing of k_raycast inside one fused frame (diagnostic build of libdsb200 only:
`make -C denseslam-global-consistency-h_b200/csrc clean && make -C ... EXTRA=-DDSB_RAY_TRACE`; the product build has no trace
code).  Builds a C2 map of --frames frames, fuses one more frame with the trace buffer set, and prints when each raycast tile
started and ended (globaltimer), how many loop trips its longest ray took, and the time per trip of the longest tiles.

    python tools/ray_trace.py --frames 40 [--set ray_lpt=0 --set ray_pf_after=-1]
"""
import argparse
import ctypes as C
import importlib
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
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--frames", type=int, default=40)
    ap.add_argument("--set", action="append", default=[])
    args = ap.parse_args()
    import torch
    abi = importlib.import_module(PKG + ".abi")
    eng = importlib.import_module(PKG + ".engine")
    synth = importlib.import_module(PKG + ".synth")
    lib = abi.load()
    raw = C.CDLL(abi.LIB_PATH)
    if not hasattr(raw, "dsb200_debug_ray_trace"):
        raise SystemExit("libdsb200.so was not built with EXTRA=-DDSB_RAY_TRACE")
    raw.dsb200_debug_ray_trace.argtypes = [C.c_void_p]
    abi.check(lib.dsb200_set_device(0))
    for kv in args.set:
        k, v = kv.split("=")
        abi.check(lib.dsb200_debug_set_tunable(k.encode(), int(v)))
    caps = bench.WORKLOADS[args.workload]["caps"]
    cfg, frames = synth.make_sequence(args.workload, args.frames + 4)
    W, H = cfg.W, cfg.H
    scene = eng.ITMScene(cfg.params(), caps["num_blocks"], caps["bucket_num"], caps["excess_size"])
    recon = eng.ITMSceneReconstructionEngine(caps["num_blocks"], caps["bucket_num"], caps["excess_size"])
    recon.SetFusionWeightParams(cfg.depth_weighting, 10, 100)
    vis = eng.ITMVisualisationEngine(W, H)
    rs = vis.CreateRenderState(scene, (W, H))
    trk = eng.ITMTrackingState(W, H)
    recon.ResetScene(scene)
    recon.SetAsync(True)
    view = eng.ITMView(W, H, cfg.intr)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    tilesX, tilesY = (W + 7) // 8, (H + 3) // 4
    n = tilesX * tilesY
    trace = torch.zeros(n * 17 * 4, dtype=torch.int64, device="cuda:0")     # tiles, then up to 16 group-marched items per tile
    for rep in range(4):
        i = args.frames + rep
        if rep == 0:
            for j in range(args.frames):
                view.update(frames[j].depth, frames[j].rgb, frames[j].timestamp); trk.set_pose(frames[j].M)
                recon.ProcessFrame(vis, scene, view, trk, rs)
        view.update(frames[i].depth, frames[i].rgb, frames[i].timestamp); trk.set_pose(frames[i].M)
        recon.Sync(scene, rs); torch.cuda.synchronize()
        flush.zero_(); trace.zero_(); torch.cuda.synchronize()
        assert raw.dsb200_debug_ray_trace(C.c_void_p(trace.data_ptr())) == 0
        recon.ProcessFrame(vis, scene, view, trk, rs)
        recon.Sync(scene, rs); torch.cuda.synchronize()
        assert raw.dsb200_debug_ray_trace(None) == 0
        full = trace.cpu().numpy().reshape(n * 17, 4)
        items = full[n:][full[n:, 0] > 0]
        ids = np.nonzero(full[:n, 0] > 0)[0]
        t = full[:n][ids]
        t0 = min(t[:, 0].min(), items[:, 0].min()) if len(items) else t[:, 0].min()
        if len(items):
            big = items[np.argsort(items[:, 1] - items[:, 0])[::-1][:5]]
            print("   slowest items (us, longest ray's trips, rounds):", [(round((b[1] - b[0]) / 1e3, 1), int(b[2]), int(b[3])) for b in big])
            print("   all items: trips of the longest ray mean %.1f, rounds mean %.1f -> %.2f trips per round; %.2f us per round" % (
                items[:, 2].mean(), items[:, 3].mean(), items[:, 2].sum() / max(items[:, 3].sum(), 1), (items[:, 1] - items[:, 0]).sum() / 1e3 / max(items[:, 3].sum(), 1)))
            ist, ien = (items[:, 0] - t0) / 1e3, (items[:, 1] - t0) / 1e3
            print("frame %d: %d group-marched items: start mean %.1f max %.1f us, end mean %.1f max %.1f us, duration mean %.1f p90 %.1f max %.1f us, sum %.0f us" % (
                i, len(items), ist.mean(), ist.max(), ien.mean(), ien.max(), (ien - ist).mean(), np.percentile(ien - ist, 90), (ien - ist).max(), (ien - ist).sum()))
        start, end, trips = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3, t[:, 2]
        dur = end - start
        order = np.argsort(end)[::-1]
        n_t = len(t)
        print("frame %d: kernel span %.1f us; tiles %d; tickets all handed out by %.1f us; sum of tile time %.0f us" % (i, end.max(), n_t, start.max(), dur.sum()))
        for k in order[:6]:
            print("   tile (%3d,%3d): start %6.1f  end %6.1f  dur %6.1f us  trips %4d  -> %.2f us/trip  sm %d" % (
                ids[k] // tilesX, ids[k] % tilesX, start[k], end[k], dur[k], trips[k], dur[k] / max(trips[k], 1), t[k, 3]))
        for lo, hi in ((1, 8), (8, 16), (16, 32), (32, 64), (64, 1000)):
            m = (trips >= lo) & (trips < hi)
            if m.any():
                print("   tiles with %3d..%3d trips: %5d, mean %.2f us/trip, mean start %.1f us" % (lo, hi, m.sum(), (dur[m] / trips[m]).mean(), start[m].mean()))
    recon.SetAsync(False)


if __name__ == "__main__":
    main()
