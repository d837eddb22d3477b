"""Digests of a scene's observable state (numpy + hashlib only), so that a run can prove WHICH map it produced.

bench.py takes one after its timed passes and compares it with the digest committed under tests/golden/bench_digests.json
(`parity_check` in its JSON line); tests/golden/make_bench_digests.py produces those from the CPU checker on the same
synthetic frames.  What enters a digest is exactly what the parity tests compare (tests/harness.py assert_state_equal /
assert_images_equal): hash entries without allocatedTime, free-list tops, the visible list, the voxels of every
allocated block without the padding byte, and the frame's images."""
import hashlib

import numpy as np

BLOCK = 512


def _sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        a = np.ascontiguousarray(a)
        mv = memoryview(a).cast("B")
        for o in range(0, len(mv), 1 << 26):
            h.update(mv[o:o + (1 << 26)])
    return h.hexdigest()[:32]


def state_digest(hash_table, last_free, visible, fetch_blocks, num_blocks, images):
    """hash_table: structured (pos, pad, offset, ptr, allocated_time); last_free: (block, excess) tops; visible: int32 list;
    fetch_blocks(lo, hi) -> voxel records of blocks [lo, hi) (structured, 8 bytes each); images: dict name -> array."""
    out = {}
    out["hash"] = _sha(hash_table["pos"], hash_table["offset"], hash_table["ptr"], np.asarray(last_free, np.int64))
    out["visible"] = _sha(np.asarray(visible, np.int32))
    live = hash_table["ptr"][hash_table["ptr"] >= 0].astype(np.int64)
    out["blocks_allocated"] = int(live.size)
    if live.size:
        lo = int(live.min())
        vox = fetch_blocks(lo, num_blocks).view(np.uint64).reshape(-1, BLOCK)
        h = hashlib.sha256()
        mask = np.uint64(0x00ffffffffffffff)          # 8th byte of ITMVoxel_s_rgb is padding
        for o in range(0, live.size, 8192):           # slot order, bounded temporaries
            h.update(np.ascontiguousarray(vox[live[o:o + 8192] - lo] & mask).tobytes())
        out["voxels"] = h.hexdigest()[:32]
    else:
        out["voxels"] = _sha(np.zeros(0, np.uint8))
    for k in sorted(images.keys()):
        out["image_" + k] = _sha(images[k])
    return out


def compare(got, want):
    """-> 'ok' or a short description of the components that differ"""
    bad = [k for k in sorted(want.keys()) if got.get(k) != want[k]]
    return "ok" if not bad else "MISMATCH in " + ", ".join(bad)
