#!/usr/bin/env python
"""Generates tests/golden/tiny_aux_ref.npz from the UNMODIFIED reference CPU code (oracle/_ref/libitmref.so): known answers for
the two auxiliary paths, ITMViewBuilder_CPU::ComputeNormalAndWeights and ITMSwappingEngine_CPU on ITMGlobalCache.

Run where /root/reference exists:   python tests/golden/make_aux_golden.py
Inputs travel with the outputs (depth images and intrinsics; int16-mm depth, rgb and poses of the swapping frames); the swapping
outputs are SHA-256 digests of the state after every step (see swap_digests)."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import harness as hz  # noqa: E402
from harness import abi, po  # noqa: E402

NORMAL_CASES = ("wall 0.004", "wall -0.03", "noise", "noise 1/f")
FILL_N, FILL_S = 7.25, -3.5


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def swap_digests(hash_table, voxels, alloc_list, last_free_block, states, has_stored, stored_block):
    """layout-independent digest of what the swapping engine touches: hash pointers, the live voxel payload (pad byte masked),
    the live part of the free list, the swap states, the stored flags and the stored blocks in entry order"""
    live = np.nonzero(hash_table["ptr"] >= 0)[0]
    blocks = hash_table["ptr"][live].astype(np.int64)
    vox = voxels.reshape(-1, abi.BLOCK_SIZE3)[blocks].view(np.uint64) & np.uint64(0x00ffffffffffffff)
    stored = np.nonzero(has_stored)[0]
    h = hashlib.sha256()
    for e in stored:
        h.update((np.ascontiguousarray(stored_block(int(e))).view(np.uint64) & np.uint64(0x00ffffffffffffff)).tobytes())
    return dict(ptr=sha(hash_table["ptr"].astype(np.int32)), voxels=sha(vox), free_list=sha(alloc_list[:last_free_block + 1].astype(np.int32)),
                last_free=str(int(last_free_block)), states=sha(states.astype(np.uint8)), has_stored=sha(stored.astype(np.int32)), stored=h.hexdigest())


def main():
    out = {}
    ref = None
    # ---- ComputeNormalAndWeights
    cases = {name: (d, k) for name, d, k in hz.normal_and_weight_cases()}
    cfg, frames, order, whole_map_after = hz.swap_scenario()
    ref = po.RefEngine(cfg.W, cfg.H, cfg.params(), use_swapping=True)
    ref.set_calib(cfg.intr)
    for i, name in enumerate(NORMAL_CASES):
        depth, intr = cases[name]
        n = np.full(depth.shape + (4,), FILL_N, np.float32); s = np.full(depth.shape, FILL_S, np.float32)
        ref.compute_normal_and_weights(depth, intr, n, s)
        out["nw_depth_%d" % i], out["nw_intr_%d" % i] = depth, np.asarray(intr, np.float32)
        out["nw_normal_%d" % i], out["nw_sigma_%d" % i] = n, s
    # ---- swapping engine, the schedule of tests/test_oracle_vs_ref.py::test_swapping_engine_matches_reference
    out.update(W=cfg.W, H=cfg.H, intr=np.asarray(cfg.intr, np.float32), voxel_size=np.float32(cfg.voxel_size), mu=np.float32(cfg.mu),
               max_w=cfg.max_w, vf_min=np.float32(cfg.vf_min), vf_max=np.float32(cfg.vf_max),
               depth_mm=np.stack([f.depth_mm for f in frames]), rgb=np.stack([f.rgb[..., :3] for f in frames]),
               M=np.stack([f.M for f in frames]), timestamps=np.array([f.timestamp for f in frames]),
               order=np.array(order, np.int32), whole_map_after=whole_map_after)
    digests = []
    for step, k in enumerate(order):
        f = frames[k]
        ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
        ref.allocate(); ref.integrate()
        ref.swap_integrate_global_into_local(); ref.swap_save_to_global_memory(True)
        if step == whole_map_after:
            ref.swap_save_to_global_memory(False)
        d = swap_digests(ref.hash(), ref.voxels(), ref.alloc_list(), ref.last_free()[0], ref.swap_states(), ref.swap_has_stored(),
                         ref.swap_stored_block)
        digests.append([d[k2] for k2 in sorted(d)])
    out["swap_digest_keys"] = np.array(sorted(d))
    out["swap_digests"] = np.array(digests)
    ref.close()
    path = os.path.join(HERE, "tiny_aux_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
