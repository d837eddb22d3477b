"""torchrun worker of tests/test_sharded.py::test_gloo_world2_sharded_matches_single_process (CPU, gloo).
Rank r restates shard r with the oracle; rank 0 owns the camera stream and broadcasts every frame; the raycast
results are all-gathered and composited on every rank -- the same sequence DistShardedScene drives on GPUs."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import conftest  # noqa: F401,E402  (path set-up)
import harness as hz  # noqa: E402


def main():
    out_dir, mode = sys.argv[1], sys.argv[2]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    cfg, frames = hz.synth.make_sequence("tiny", 3)
    if rank != 0:
        frames = [None] * len(frames)          # only the root has the camera stream
    sc = hz.OracleShardedScene(cfg, hz.SMALL_CAPS, world, ranks=[rank], mode=mode)
    npx = cfg.W * cfg.H
    for i in range(len(frames)):
        buf = torch.zeros(npx * 2 + 18, dtype=torch.float32)
        if rank == 0:
            f = frames[i]
            h = buf.numpy()
            h[:npx] = f.depth.reshape(-1)
            h[npx:2 * npx] = np.ascontiguousarray(f.rgb, np.uint8).reshape(-1).view(np.float32)
            h[2 * npx:2 * npx + 16] = np.asarray(f.M, np.float32).reshape(-1)
            h[2 * npx + 16:] = np.array([f.timestamp], np.float64).view(np.float32)
        dist.broadcast(buf, src=0)
        h = buf.numpy()
        f = hz.synth.Frame()
        f.depth = h[:npx].reshape(cfg.H, cfg.W).copy()
        f.rgb = h[npx:2 * npx].copy().view(np.uint8).reshape(cfg.H, cfg.W, 4)
        f.M = h[2 * npx:2 * npx + 16].copy()
        f.timestamp = float(h[2 * npx + 16:].copy().view(np.float64)[0])
        def gather(arr):
            mine = torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1))
            parts = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(parts, mine)
            return [p.numpy().view(arr.dtype).reshape(arr.shape) for p in parts]
        if mode == "composite":
            sc.fuse_and_raycast(f)
            sc.composite_and_icp(gather(sc.members[0].rs.raycast_result))
        else:
            # the GPU path reads the owners' tables in place over NVLink; over gloo the tables are gathered instead
            sc.fuse(f)
            m = sc.members[0]
            hashes, voxels, ranges = gather(m.scene.hash), gather(m.scene.voxels), gather(m.rs.range_image)
            scenes = []
            for r in range(world):
                c = hz.abi.Scene()
                C.memmove(C.byref(c), C.byref(m.scene.c), C.sizeof(c))
                c.hash, c.voxels = hz.po.ptr(hashes[r]), hz.po.ptr(voxels[r])
                scenes.append(c)
            sc.raycast_lookup_and_icp(scenes, ranges)
    img = sc.images()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), hash_ptr=sc.members[0].scene.hash["ptr"], **img)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
