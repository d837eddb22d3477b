"""Sharded scene: ownership function, CPU restatement of the sharded pipeline, its size-independent
properties against the unsharded scene, and the world_size-2 gloo run of the N > 1 host logic (all on CPU)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import harness as hz
from harness import abi, synth

BIG_TABLE = dict(num_blocks=0x4000, bucket_num=1 << 20, excess_size=0x2000)     # collisions are rare at this table size


def test_shard_owner_host_helper_matches_oracle_and_partitions():
    lib, o = abi.load(), hz.po.Oracle()
    rng = np.random.default_rng(11)
    pts = rng.integers(-2000, 2000, size=(4000, 3))
    for count, bl in ((2, 2), (3, 0), (8, 2), (8, 3), (16, 1)):
        owners = np.array([lib.dsb200_shard_owner(int(x), int(y), int(z), count, bl) for x, y, z in pts])
        ref = np.array([o.L.rst_shard_owner(int(x), int(y), int(z), count, bl) for x, y, z in pts])
        assert np.array_equal(owners, ref)
        assert owners.min() >= 0 and owners.max() == count - 1
        share = np.bincount(owners, minlength=count) / len(pts)
        assert share.min() > 0.5 / count, "ownership badly unbalanced: %s" % share
        # all blocks of one brick share an owner (negative coordinates floor, not truncate)
        base = (pts[:200] >> bl) << bl
        for dx in (0, (1 << bl) - 1):
            moved = base + dx
            a = [lib.dsb200_shard_owner(int(x), int(y), int(z), count, bl) for x, y, z in base]
            b = [lib.dsb200_shard_owner(int(x), int(y), int(z), count, bl) for x, y, z in moved]
            assert a == b
    assert lib.dsb200_shard_owner(5, -3, 9, 1, 2) == 0


def test_sharded_oracle_partitions_the_unsharded_scene():
    cfg, frames = synth.make_sequence("tiny", 5)
    single = hz.OracleBackend(cfg, BIG_TABLE)
    sharded = hz.OracleShardedScene(cfg, BIG_TABLE, 3, brick_log2=1)
    for f in frames:
        single.set_frame(f); single.process_frame()
        sharded.process_frame(f)
    whole = hz.allocated_blocks(single.scene.hash)
    parts = [hz.allocated_blocks(m.scene.hash) for m in sharded.members]
    lib = abi.load()
    for r, p in enumerate(parts):
        assert p, "shard %d owns nothing" % r
        assert all(lib.dsb200_shard_owner(x, y, z, 3, 1) == r for x, y, z in p), "shard %d allocated a foreign block" % r
    union = set().union(*parts)
    assert sum(len(p) for p in parts) == len(union)                      # disjoint
    # identical block set up to same-frame bucket collisions (which differ between one table and three)
    assert len(union ^ whole) <= 0.01 * len(whole), "block sets differ: %d vs %d, %d mismatched" % (len(union), len(whole), len(union ^ whole))
    # voxel payload of every block both runs hold: integration is per block, so bit-identical unless a collision delayed its allocation
    def payload(backend):
        h = backend.scene.hash
        live = np.nonzero(h["ptr"] >= 0)[0]
        vox = backend.scene.voxels.view(np.uint64).reshape(-1, 512)
        return {tuple(h["pos"][i].tolist()): vox[h["ptr"][i]] for i in live}
    pw = payload(single)
    same = diff = 0
    for m in sharded.members:
        for k, v in payload(m).items():
            if k in pw:
                if np.array_equal(v, pw[k]): same += 1
                else: diff += 1
    assert same > 0 and diff <= 0.01 * (same + diff), "voxel payloads differ in %d of %d common blocks" % (diff, same + diff)
    single.close(); sharded.close()


def test_lookup_raycast_equals_unsharded_raycast_when_block_sets_agree():
    """Owner-routed lookups march exactly like one big scene: wherever the shards hold the same blocks with the same
    payload as the unsharded run (always, up to same-frame bucket collisions), the images are bit-identical."""
    cfg, frames = synth.make_sequence("tiny", 4)
    single = hz.OracleBackend(cfg, BIG_TABLE)
    sharded = hz.OracleShardedScene(cfg, BIG_TABLE, 4, brick_log2=1)
    for f in frames:
        single.set_frame(f); single.process_frame()
        sharded.process_frame(f)
    whole = hz.allocated_blocks(single.scene.hash)
    union = set().union(*[hz.allocated_blocks(m.scene.hash) for m in sharded.members])
    a, b = single.images(), sharded.images()
    same = np.all(hz.bits(a["raycast"]) == hz.bits(b["composite"]), axis=-1)
    if union == whole:
        assert same.all()
        for k in ("points", "normals", "rendering"):
            assert np.array_equal(hz.bits(a[k]), hz.bits(b[k])), k
    else:   # a bucket collision delayed a block in one of the runs: only rays near it may differ
        assert same.mean() > 0.99
    assert np.array_equal(hz.bits(single.rs.range_image[:cfg.H // 8 + 1]), hz.bits(sharded.merged_range[:cfg.H // 8 + 1])) or union != whole
    single.close(); sharded.close()


def test_sharded_composite_close_to_unsharded_raycast_with_halo():
    """With the +1 halo a ray inside an owned block never reads a missing neighbour; what remains are rays that cross
    between shards' regions.  North-star tolerance for raycast depth: 1e-3 relative."""
    lib = abi.load()
    for name, bl in (("tiny", 2), ("tiny_street", 3)):
        cfg, frames = synth.make_sequence(name, 4)
        single = hz.OracleBackend(cfg, BIG_TABLE)
        sharded = hz.OracleShardedScene(cfg, BIG_TABLE, 4, brick_log2=bl, halo=True)
        for f in frames:
            single.set_frame(f); single.process_frame()
            sharded.process_frame(f)
        # held = owned or +1 neighbour of an owned block
        for r, m in enumerate(sharded.members):
            sh = abi.Shard(r, 4, bl, 1)
            blocks = hz.allocated_blocks(m.scene.hash)
            assert all(lib.dsb200_shard_holds(x, y, z, C.byref(sh)) for x, y, z in blocks)
            assert any(lib.dsb200_shard_owner(x, y, z, 4, bl) != r for x, y, z in blocks)       # the halo is populated
        a, b = single.rs.raycast_result, sharded.composite
        hit_a, hit_b = a[..., 3] > 0, b[..., 3] > 0
        assert (hit_a == hit_b).mean() > 0.995, "%s: hit masks agree on %.4f of the pixels" % (name, (hit_a == hit_b).mean())
        both = hit_a & hit_b
        cam = single.o.mat4_inverse(frames[-1].M)[12:15] / cfg.params().voxel_size
        za = np.linalg.norm(a[..., :3][both] - cam, axis=-1)
        zb = np.linalg.norm(b[..., :3][both] - cam, axis=-1)
        rel = np.abs(za - zb) / za
        assert (rel <= 1e-3).mean() > 0.90, "%s: only %.4f of the common hits within 1e-3 relative depth" % (name, (rel <= 1e-3).mean())
        assert np.median(rel) < 1e-6
        single.close(); sharded.close()


@pytest.mark.parametrize("mode", ["lookup", "composite"])
def test_gloo_world2_sharded_matches_single_process(mode):
    """N > 1 host logic on CPU: two processes, one oracle shard each, frame broadcast + exchange of the shards' tables
    (lookup) or raycast results (composite) over gloo; must equal the single-process restatement bit for bit."""
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dist_shard_worker.py")
    out = os.path.join(hz.ROOT, ".pytest_cache", "shard_world2")
    os.makedirs(out, exist_ok=True)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731" if mode == "lookup" else "29732", worker, out, mode]
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    cfg, frames = synth.make_sequence("tiny", 3)
    ref = hz.OracleShardedScene(cfg, hz.SMALL_CAPS, 2, mode=mode)
    for f in frames:
        ref.process_frame(f)
    want = ref.images()
    for rank in (0, 1):
        got = np.load(os.path.join(out, "rank%d.npz" % rank))
        for k in want:
            assert np.array_equal(hz.bits(want[k]), hz.bits(got[k])), "rank %d image %s differs" % (rank, k)
        assert np.array_equal(got["hash_ptr"], ref.members[rank].scene.hash["ptr"])
    ref.close()
