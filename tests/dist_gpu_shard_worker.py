"""torchrun worker of tests/test_gpu_sharded.py::test_two_gpu_dist_sharded_scene: one shard per GPU (NCCL broadcast of the
frame, CUDA IPC peer reads), every rank checks its own shard against the CPU restatement, the root checks the images."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import conftest  # noqa: F401,E402
import harness as hz  # noqa: E402


def main():
    mode = sys.argv[1]
    local = int(os.environ.get("LOCAL_RANK", "0"))
    hz.abi.check(hz.abi.load().dsb200_set_device(local))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    sharded = importlib.import_module(hz.PKG + ".sharded")
    cfg, frames = hz.synth.make_sequence("tiny", 4)
    sc = sharded.DistShardedScene(cfg, hz.SMALL_CAPS, brick_log2=1, mode=mode)
    ora = hz.OracleShardedScene(cfg, hz.SMALL_CAPS, world, brick_log2=1, mode=mode)
    for i, f in enumerate(frames):
        packed = sc.pack_frame(f).to(sc.dev) if rank == 0 else None
        sc.process_frame(packed)              # pose and timestamp travel inside the broadcast buffer
        ora.process_frame(f)
        torch.cuda.synchronize()
        hz.abi.check(hz.abi.load().dsb200_recon_sync(sc.m.recon.h, sc.m.scene.c, sc.m.rs.c))
        hz.assert_state_equal(ora.members[rank].state(), hz.gpu_member_state(sc.m), "%s rank %d frame %d" % (mode, rank, i))
        if rank == 0:
            got = dict(composite=sc.m.composite_image(), points=sc.m.trk.icp_points(), normals=sc.m.trk.icp_normals(),
                       rendering=sc.m.rs.raycast_image())
            hz.assert_images_equal(ora.images(), got, "%s frame %d" % (mode, i))
        dist.barrier()
    sc.close()
    dist.barrier()
    if rank == 0:
        print("SHARDED_OK", mode)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
