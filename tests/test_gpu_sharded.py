"""Sharded scene on the GPU, through the C ABI, against the CPU restatement of the sharded pipeline.
LocalShardedScene keeps all shards on one B200 (the transport is the only thing missing); the 2-GPU test runs
DistShardedScene under torchrun with NCCL + CUDA IPC peer reads and skips on a 1-GPU box."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

import harness as hz
from harness import abi, synth

pytestmark = pytest.mark.gpu
sharded = importlib.import_module(hz.PKG + ".sharded")


@pytest.mark.parametrize("mode,count,brick", [("lookup", 3, 1), ("lookup", 4, 2), ("composite", 3, 2)])
def test_local_sharded_scene_bit_exact_vs_oracle(mode, count, brick):
    cfg, frames = synth.make_sequence("tiny", 5)
    gpu = sharded.LocalShardedScene(cfg, hz.SMALL_CAPS, count, brick_log2=brick, mode=mode)
    ora = hz.OracleShardedScene(cfg, hz.SMALL_CAPS, count, brick_log2=brick, mode=mode)
    for i, f in enumerate(frames):
        gpu.process_frame(f)
        ora.process_frame(f)
        for r in range(count):
            abi.check(abi.load().dsb200_recon_sync(gpu.members[r].recon.h, gpu.members[r].scene.c, gpu.members[r].rs.c))
            hz.assert_state_equal(ora.members[r].state(), hz.gpu_member_state(gpu.members[r]), "%s shard %d frame %d" % (mode, r, i))
        root = gpu.members[0]
        got = dict(composite=root.composite_image(), points=root.trk.icp_points(), normals=root.trk.icp_normals(),
                   rendering=root.rs.raycast_image())
        hz.assert_images_equal(ora.images(), got, "%s frame %d" % (mode, i))
    if mode == "lookup":
        merged = root.merged_range.download(np.float32, cfg.W * cfg.H * 2).reshape(cfg.H, cfg.W, 2)
        assert np.array_equal(hz.bits(merged[:cfg.H // 8 + 1]), hz.bits(ora.merged_range[:cfg.H // 8 + 1]))
    gpu.close(); ora.close()


def test_lookup_raycast_of_one_shard_is_the_plain_raycast():
    """count = 1: the sharded entry point must reproduce dsb200_create_icp_maps' raycast bit for bit"""
    cfg, frames = synth.make_sequence("tiny_street", 3)
    one = sharded.LocalShardedScene(cfg, hz.SMALL_CAPS, 1)
    ref = hz.GpuBackend(cfg, hz.SMALL_CAPS)
    for f in frames:
        one.process_frame(f)
        ref.set_frame(f); ref.process_frame()
    hz.assert_images_equal(dict(r=ref.rs.raycast_result(), p=ref.trk.icp_points()),
                           dict(r=one.members[0].composite_image(), p=one.members[0].trk.icp_points()), "count=1")
    one.close(); ref.close()


def test_two_gpu_dist_sharded_scene():
    if abi.load().dsb200_device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dist_gpu_shard_worker.py")
    for mode, port in (("lookup", "29741"), ("composite", "29742")):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
               "--master-port", port, worker, mode]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-4000:]
        assert "SHARDED_OK" in r.stdout, r.stdout[-4000:]
