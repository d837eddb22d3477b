"""Host-side logic added in round 2, on the CPU: the state digest bench.py's parity_check uses, the periodic C4 street, the
per-shard capacity / row-slice arithmetic of the sharded scene, and the set-level comparator used against the reference CUDA
engine (exercised here on synthetic tables, since that engine needs a GPU)."""
import importlib
import json
import os

import numpy as np
import pytest

import harness as hz
from harness import abi, synth

digest = importlib.import_module(hz.PKG + ".digest")
sharded_caps = None


def _oracle_digest(ora):
    s = ora.state(voxels=False)
    vox = ora.scene.voxels
    return digest.state_digest(s["hash"], s["last_free"], s["visible"], lambda lo, hi: vox[lo * 512:hi * 512], ora.scene.c.num_blocks, ora.images())


def test_state_digest_sees_every_component():
    cfg, frames = synth.make_sequence("tiny", 3)
    ora = hz.OracleBackend(cfg, hz.SMALL_CAPS)
    for f in frames:
        ora.set_frame(f); ora.process_frame()
    d0 = _oracle_digest(ora)
    assert digest.compare(_oracle_digest(ora), d0) == "ok"
    live = np.nonzero(ora.scene.hash["ptr"] >= 0)[0]
    # one voxel of one allocated block
    p = int(ora.scene.hash["ptr"][live[7]]) * 512 + 33
    old = ora.scene.voxels["sdf"][p]
    ora.scene.voxels["sdf"][p] = old - 1
    assert digest.compare(_oracle_digest(ora), d0) == "MISMATCH in voxels"
    ora.scene.voxels["sdf"][p] = old
    # the padding byte is not part of the state
    ora.scene.voxels["pad"][p] = 77
    assert digest.compare(_oracle_digest(ora), d0) == "ok"
    ora.scene.voxels["pad"][p] = 0
    # a voxel of a block nobody references does not count either; a hash pointer does
    ora.scene.hash["offset"][live[3]] += 1
    assert "hash" in digest.compare(_oracle_digest(ora), d0)
    ora.scene.hash["offset"][live[3]] -= 1
    ora.rs.raycast_result[5, 6, 2] += 1.0
    assert digest.compare(_oracle_digest(ora), d0) == "MISMATCH in image_raycast"
    ora.close()


def test_committed_bench_digests_cover_the_bench_defaults():
    import bench
    table = json.load(open(bench.DIGESTS))
    for wl in ("C1", "C2", "C3"):
        for n in (25, 105):       # --steps 20 --warmup 5 (the driver's round-1 arguments) and the defaults
            assert "%s:%d" % (wl, n) in table
            assert set(table["%s:%d" % (wl, n)]) >= {"hash", "visible", "voxels", "image_raycast", "image_points", "blocks_allocated"}


def test_periodic_street_repeats_with_its_period():
    cfg, frames = synth.make_sequence("tiny_loop", 2 * 16 + 3)
    P = cfg.period
    for i in (0, 1, 2):
        assert frames[i].depth is frames[i + P].depth and frames[i].rgb is frames[i + 2 * P].rgb      # shared images
    # pose i + P = pose i moved P metres along the street (world -> camera, column-major: translation in m[12:15])
    M0, M1 = frames[2].M.reshape(4, 4).T, frames[2 + P].M.reshape(4, 4).T
    assert np.allclose(M0[:3, :3], M1[:3, :3], atol=1e-6)
    c0, c1 = -M0[:3, :3].T @ M0[:3, 3], -M1[:3, :3].T @ M1[:3, 3]
    assert np.allclose(c1 - c0, [0, 0, P], atol=1e-4)
    # and the world repeats: fusing period 2 allocates as many blocks as period 1 did
    ora = hz.OracleBackend(cfg, hz.SMALL_CAPS); ora.set_weights(True)
    used = []
    for i, f in enumerate(frames[:2 * P + 1]):
        ora.set_frame(f); ora.allocate(); ora.integrate()
        if i % P == 0:
            used.append(hz.SMALL_CAPS["num_blocks"] - 1 - ora.scene.c.last_free_block)
    assert abs((used[2] - used[1]) - (used[1] - used[0])) <= 0.1 * (used[1] - used[0]), used
    ora.close()


def test_shard_capacities_and_row_slices():
    sh = importlib.import_module(hz.PKG + ".sharded")
    caps = dict(num_blocks=1 << 22, bucket_num=1 << 23, excess_size=1 << 21)
    assert sh.shard_caps(caps, 1) == caps
    for n in (2, 4, 8):
        for halo in (False, True):
            c = sh.shard_caps(caps, n, brick_log2=3, halo=halo)
            rep = (9.0 / 8.0) ** 3 if halo else 1.0
            assert c["num_blocks"] * n >= caps["num_blocks"] * rep and c["num_blocks"] <= caps["num_blocks"]
            assert c["bucket_num"] & (c["bucket_num"] - 1) == 0 and c["bucket_num"] >= c["num_blocks"]
    for H, n in ((375, 8), (480, 3), (96, 2), (7, 8)):
        rows = [sh.row_slice(H, r, n) for r in range(n)]
        assert rows[0][0] == 0 and rows[-1][1] == H and all(rows[r][1] == rows[r + 1][0] for r in range(n - 1))


def _table(keys, ptrs, voxels, visible, free, cap):
    keys = np.asarray(keys, np.int64)
    order = np.argsort(keys)
    return dict(keys=keys[order], slots=np.arange(len(keys))[order], ptrs=np.asarray(ptrs)[order], free_blocks=free, num_blocks=cap,
                duplicates=0, allocated=len(keys), aliased=np.zeros(0, np.int64), visible_keys=np.sort(np.asarray(visible, np.int64)),
                visible_dead=0, voxels=voxels[order])


def test_set_level_comparator_accepts_and_rejects_what_it_should():
    rng = np.random.default_rng(3)
    pos = rng.integers(-400, 400, size=(400, 3)).astype(np.int16)
    keys = np.unique(hz.pos_key(pos))
    bk = hz.SetTracker()._buckets(keys)
    _, first = np.unique(bk, return_index=True)
    keys = np.sort(keys[first])                       # one block per bucket: nothing is "chained"
    n = len(keys)
    vox = np.zeros((n, 512), abi.VOXEL_DTYPE)
    vox["sdf"] = rng.integers(-3000, 3000, size=(n, 512))
    A = lambda v=vox, k=keys: _table(k, np.arange(len(k)), v, k[:50], 1000 - len(k), 1000)
    B = lambda v=vox, k=keys: _table(k[::-1], np.arange(len(k))[::-1] + 7, v[::-1], k[:50][::-1], 2000 - len(k), 2000)   # other order / pointers / capacity
    assert hz.assert_sets_match(hz.SetTracker(), A(), B(), "identical sets") == 0
    # the padding byte is ignored; three differing blocks outside any chain are a failure
    v2 = vox.copy(); v2["pad"][5] = 9
    assert hz.assert_sets_match(hz.SetTracker(), A(v2), B(), "padding") == 0
    v3 = vox.copy(); v3["w_depth"][[5, 9, 11], 3] = 1
    with pytest.raises(AssertionError, match="outside collision chains"):
        hz.assert_sets_match(hz.SetTracker(), A(v3), B(), "three loose differences")
    # a block born a frame later in one engine is counted, not compared
    tr = hz.SetTracker()
    hz.assert_sets_match(tr, A(vox[:-1], keys[:-1]), B(vox, keys), "frame 0: ours lost a bucket race")
    v4 = vox.copy(); v4["w_depth"][n - 1] = 1          # ... so it has seen one observation less
    assert hz.assert_sets_match(tr, A(v4), B(), "frame 1") == 1
    # leaking a voxel block is a failure of ours
    leak = A(); leak["free_blocks"] -= 1
    with pytest.raises(AssertionError, match="leaks"):
        hz.assert_sets_match(hz.SetTracker(), leak, B(), "leak")
    # a different block SET beyond the tolerance is a failure
    with pytest.raises(AssertionError):
        hz.assert_sets_match(hz.SetTracker(), A(vox[:n // 2], keys[:n // 2]), B(), "half the map missing")


def test_oracle_ray_trip_hooks_count_consistently():
    """oracle/restate.cpp's analysis hooks (tools/ray_trip_stats.py): per-pixel loop trips and their kinds are consistent, and
    switching them on changes nothing the checker computes"""
    import ctypes as C
    cfg, frames = hz.synth.make_sequence("tiny", 3)
    plain, hooked = hz.OracleBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    lib = C.CDLL(hz.po.RESTATE_PATH)
    lib.rst_debug_ray_steps.argtypes = [C.c_void_p, C.c_int]; lib.rst_debug_ray_kinds.argtypes = [C.c_void_p]
    n = cfg.W * cfg.H
    steps, kinds = np.zeros(n, np.int32), np.zeros(6 * n, np.int32)
    for i, f in enumerate(frames):
        plain.set_frame(f); plain.process_frame()
        if i + 1 == len(frames):
            lib.rst_debug_ray_steps(steps.ctypes.data_as(C.c_void_p), cfg.W); lib.rst_debug_ray_kinds(kinds.ctypes.data_as(C.c_void_p))
        hooked.set_frame(f); hooked.process_frame()
        lib.rst_debug_ray_steps(None, 0); lib.rst_debug_ray_kinds(None)
    k = kinds.reshape(n, 6)
    assert steps.max() > 8 and (steps > 0).sum() > n // 2
    assert (k[:, 0] + k[:, 1] <= steps).all() and (k[:, 3] <= k[:, 0]).all() and (k[:, 5] <= k[:, 4]).all() and (k[:, 2] <= steps).all()
    assert k[:, 0].sum() > 0 and k[:, 2].sum() > 0
    hz.assert_images_equal(plain.images(), hooked.images(), "hooks on / off")
    plain.close(); hooked.close()
