#!/usr/bin/env python
"""tools/ray_trip_stats.py -- what the raycast's march loop does per ray, counted on the CPU checker (oracle/librestate.so's
analysis hooks rst_debug_ray_steps / rst_debug_ray_kinds; TEST INFRASTRUCTURE, no GPU): trips per ray and how many of them found
no block (8-voxel step), took the minimum one-voxel step, or sampled tri-linearly.  Explains where k_raycast's tail comes from.

    python tools/ray_trip_stats.py --workload C2 --frames 30
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import harness as hz  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--frames", type=int, default=30)
    args = ap.parse_args()
    spec = bench.WORKLOADS[args.workload]
    cfg, frames = hz.synth.make_sequence(args.workload, args.frames, seed=0)
    ora = hz.OracleBackend(cfg, spec["caps"])
    ora.set_weights(cfg.depth_weighting, 10, 100)
    lib = C.CDLL(hz.po.RESTATE_PATH)      # the same mapped library: the hooks are process-wide
    lib.rst_debug_ray_steps.argtypes = [C.c_void_p, C.c_int]; lib.rst_debug_ray_kinds.argtypes = [C.c_void_p]
    n = cfg.W * cfg.H
    steps = np.zeros(n, np.int32); kinds = np.zeros(6 * n, np.int32)
    for i, f in enumerate(frames):
        ora.set_frame(f)
        last = i + 1 == len(frames)
        if last:
            lib.rst_debug_ray_steps(steps.ctypes.data_as(C.c_void_p), cfg.W)
            lib.rst_debug_ray_kinds(kinds.ctypes.data_as(C.c_void_p))
        assert ora.process_frame() == 0
    lib.rst_debug_ray_steps(None, 0); lib.rst_debug_ray_kinds(None)
    kinds = kinds.reshape(n, 6)
    marching = steps > 0
    print("%s frame %d: %d of %d rays march; trips mean %.1f, p50 %d, p90 %d, p99 %d, max %d" % (
        args.workload, args.frames - 1, marching.sum(), n, steps[marching].mean(), np.percentile(steps[marching], 50),
        np.percentile(steps[marching], 90), np.percentile(steps[marching], 99), steps.max()))
    tot = steps.sum()
    print("all trips: %.1f %% found no block (8-voxel step), %.1f %% minimum one-voxel steps, %.1f %% sampled tri-linearly" % (
        100.0 * kinds[:, 0].sum() / tot, 100.0 * kinds[:, 1].sum() / tot, 100.0 * kinds[:, 2].sum() / tot))
    for lo in (32, 64, 100):
        m = steps >= lo
        if m.any():
            t = steps[m].sum()
            print("rays with >= %3d trips: %6d rays (%.2f %%), %.1f %% of all trips; of theirs %.1f %% no block, %.1f %% one-voxel, %.1f %% tri-linear" % (
                lo, m.sum(), 100.0 * m.sum() / marching.sum(), 100.0 * t / tot, 100.0 * kinds[m, 0].sum() / t, 100.0 * kinds[m, 1].sum() / t,
                100.0 * kinds[m, 2].sum() / t))
    for lo in (1, 32, 64):
        m = steps >= lo
        nf, nfc, em, emc = kinds[m, 0].sum(), kinds[m, 3].sum(), kinds[m, 4].sum(), kinds[m, 5].sum()
        print("rays with >= %3d trips: no-block trips come in runs of %.2f on average, empty-block (sdf = 1) steps are %.1f %% of the trips, in runs of %.2f" % (
            lo, nf / max(nf - nfc, 1), 100.0 * em / steps[m].sum(), em / max(em - emc, 1)))
    # per 8x4 tile (one warp): the warp runs as long as its longest ray
    H8, W8 = cfg.H // 4, cfg.W // 8
    tiles = steps.reshape(cfg.H, cfg.W)[:H8 * 4, :W8 * 8].reshape(H8, 4, W8, 8).max(axis=(1, 3))
    rows = np.argsort(tiles.ravel())[::-1][:5]
    print("longest tiles (row of 4-pixel rows, column of 8-pixel columns, trips):", [(int(r // W8), int(r % W8), int(tiles.ravel()[r])) for r in rows])
    print("sum over tiles of the longest ray = %d warp-iterations; sum of all trips / 32 = %d" % (tiles.sum(), tot // 32))
    ora.close()


if __name__ == "__main__":
    main()
