#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference CPU engines (oracle/_ref/libitmref.so,
compiled from /root/reference by oracle/Makefile with the parity flags -O2 -ffp-contract=off).

Run where /root/reference exists:   python tests/golden/make_golden.py
The fixtures are self-contained: they carry the INPUTS (int16-mm depth, rgb, poses, intrinsics, scene
parameters) next to the reference's OUTPUTS, so tests/test_golden.py needs neither the reference tree
nor the synthetic generator to be reproducible.  Large outputs (voxel payload, per-frame images) are
stored as SHA-256 digests; the last frame's state is stored in full (sparse hash table, visible
list, visibility flags, range image, raycast result, ICP maps).
"""
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
from harness import abi, po, synth  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def state_digests(hash_table, voxels, visible, visible_type, last_free):
    """Digest of everything the north star wants bit-exact, in a layout-independent form."""
    live = np.nonzero(hash_table["ptr"] != -2)[0].astype(np.int32)
    ent = hash_table[live]
    blocks = ent["ptr"][ent["ptr"] >= 0].astype(np.int64)
    vox = voxels.reshape(-1, abi.BLOCK_SIZE3)[blocks]
    payload = np.stack([vox["sdf"].astype(np.int32), vox["w_depth"].astype(np.int32), vox["clr"][..., 0].astype(np.int32),
                        vox["clr"][..., 1].astype(np.int32), vox["clr"][..., 2].astype(np.int32), vox["w_color"].astype(np.int32)], -1)
    return dict(hash=sha(np.concatenate([live[:, None].astype(np.int32), ent["pos"].astype(np.int32), ent["offset"][:, None], ent["ptr"][:, None]], 1)),
                voxels=sha(payload), visible=sha(visible.astype(np.int32)),
                visible_type=sha(np.nonzero(visible_type)[0].astype(np.int32)) + sha(visible_type[np.nonzero(visible_type)[0]]),
                last_free="%d,%d" % tuple(last_free))


def image_digests(range_image, raycast, points, normals, rendering, W, H):
    # only the top-left (W/8+1)x(H/8+1) corner of the range image is ever read
    return dict(range=sha(range_image[:H // 8 + 1, :W // 8 + 1]), raycast=sha(raycast), points=sha(points), normals=sha(normals),
                rendering=sha(rendering))


def make(cfg_name, n_frames, use_swapping, out_name):
    cfg, frames = synth.make_sequence(cfg_name, n_frames)
    ref = po.RefEngine(cfg.W, cfg.H, cfg.params(), use_swapping=use_swapping)
    ref.set_calib(cfg.intr)
    out = dict(W=cfg.W, H=cfg.H, intr=cfg.intr, voxel_size=np.float32(cfg.voxel_size), mu=np.float32(cfg.mu), max_w=cfg.max_w,
               vf_min=np.float32(cfg.vf_min), vf_max=np.float32(cfg.vf_max), use_swapping=int(use_swapping), n_frames=n_frames,
               depth_mm=np.stack([f.depth_mm for f in frames]), rgb=np.stack([f.rgb[..., :3] for f in frames]),
               M=np.stack([f.M for f in frames]), timestamps=np.array([f.timestamp for f in frames]))
    sd, idg, nvis = [], [], []
    for f in frames:
        ref.set_view(f.depth, f.rgb, f.timestamp); ref.set_pose(f.M)
        ref.allocate(); ref.integrate(); ref.create_expected_depths(); ref.create_icp_maps()
        s = state_digests(ref.hash(), ref.voxels(), ref.visible(), ref.visible_type(), ref.last_free())
        i = image_digests(ref.range_image(), ref.raycast_result(), ref.icp_points(), ref.icp_normals(), ref.icp_rendering(), cfg.W, cfg.H)
        sd.append([s[k] for k in sorted(s)]); idg.append([i[k] for k in sorted(i)]); nvis.append(len(ref.visible()))
    out["state_keys"] = np.array(sorted(s)); out["image_keys"] = np.array(sorted(i))
    out["state_digests"] = np.array(sd); out["image_digests"] = np.array(idg); out["no_visible"] = np.array(nvis)
    # last frame in full
    h = ref.hash()
    live = np.nonzero(h["ptr"] != -2)[0].astype(np.int32)
    out["hash_live"] = live
    out["hash_pos"], out["hash_offset"], out["hash_ptr"] = h["pos"][live], h["offset"][live], h["ptr"][live]
    out["visible"] = ref.visible()
    vt = ref.visible_type()
    out["vt_idx"] = np.nonzero(vt)[0].astype(np.int32); out["vt_val"] = vt[out["vt_idx"]]
    out["last_free"] = np.array(ref.last_free())
    out["range"] = ref.range_image()[:cfg.H // 8 + 1, :cfg.W // 8 + 1].copy()
    out["raycast"] = ref.raycast_result().copy()
    out["points"] = ref.icp_points().copy(); out["normals"] = ref.icp_normals().copy()
    out["rendering"] = ref.icp_rendering()[..., 0].copy()
    # free-view greyscale render at the pose of frame 1 (ITMMainEngine::GetImage path)
    ref.render_free_view(frames[1].M, cfg.intr, abi.RENDER_SHADED_GREYSCALE)
    out["free_M"] = frames[1].M
    out["free_visible"] = ref.visible(True)
    out["free_raycast_sha"] = np.array(sha(ref.raycast_result(True)))
    out["free_image"] = ref.out_image()[..., 0].copy()
    ref.close()
    path = os.path.join(HERE, out_name)
    np.savez_compressed(path, **out)
    print("%s: %d frames, %d visible, %d live entries, %.0f KiB" % (out_name, n_frames, nvis[-1], live.size, os.path.getsize(path) / 1024.0))


if __name__ == "__main__":
    if not hz.has_ref():
        raise SystemExit("oracle/_ref/libitmref.so absent: run `make -C oracle` where /root/reference exists")
    make("tiny", 4, False, "tiny_room_ref.npz")
    make("tiny", 3, True, "tiny_room_swapping_ref.npz")
    make("tiny_street", 3, False, "tiny_street_ref.npz")
