"""Two interchangeable back ends behind one interface, so parity tests read as
`run the same calls on both, compare the state`:

  OracleBackend  -> oracle/librestate.so   (CPU restatement; the checker)
  GpuBackend     -> libdsb200.so via the package's engine mirror (the product; needs a B200)
  RefBackend     -> oracle/_ref/libitmref.so (unmodified reference CPU engines; fixed capacities,
                    no decay / slide window / de-integration -- the reference CPU engine throws)
"""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

PKG = "denseslam-global-consistency-h_b200"
abi = importlib.import_module(PKG + ".abi")
synth = importlib.import_module(PKG + ".synth")
import pyoracle as po  # noqa: E402

SMALL_CAPS = dict(num_blocks=0x4000, bucket_num=0x8000, excess_size=0x2000)


def has_ref():
    return os.path.exists(po.REF_PATH)


def has_refcuda():
    return os.path.exists(po.REF_CUDA_PATH)


def has_gpu():
    try:
        return abi.load().dsb200_device_count() > 0
    except Exception:
        return False


class OracleBackend(object):
    name = "oracle"

    def __init__(self, cfg, caps=None, use_swapping=False):
        caps = caps or {}
        self.cfg = cfg
        self.o = po.Oracle()
        self.scene = po.HostScene(cfg.params(), use_swapping=use_swapping, **caps)
        vis_cap = caps.get("num_blocks", abi.DEFAULT_LOCAL_BLOCKS)
        self.rs = po.HostRenderState(self.scene.c.total_entries, cfg.W, cfg.H, vis_cap)
        self.rs_free = po.HostRenderState(self.scene.c.total_entries, cfg.W, cfg.H, vis_cap)
        self.e = self.o.L.rst_recon_create(self.scene.c.num_blocks, self.scene.c.bucket_num, self.scene.c.excess_size)
        self.pts = np.zeros((cfg.H, cfg.W, 4), np.float32)
        self.nrm = np.zeros((cfg.H, cfg.W, 4), np.float32)
        self.out_img = np.zeros((cfg.H, cfg.W, 4), np.uint8)
        self.out_depth = np.zeros((cfg.H, cfg.W), np.float32)
        self.view = None
        self.trk = None
        self.requires_full_rendering = True
        self.reset()

    def close(self):
        self.o.L.rst_recon_destroy(self.e)

    def reset(self):
        self.o.L.rst_reset_scene(self.e, C.byref(self.scene.c))

    def set_weights(self, depth_weighting, max_new_w=10, max_distance=100):
        wp = abi.WeightParams(int(depth_weighting), max_new_w, max_distance)
        self.o.L.rst_set_fusion_weight_params(self.e, C.byref(wp))

    def set_frame(self, f, M=None):
        self._f = f
        M = f.M if M is None else M
        self.view = po.make_view(f.depth, f.rgb, self.cfg.intr, timestamp=f.timestamp)
        self.trk = po.make_tracking(M, self.o.mat4_inverse(M), self.pts, self.nrm, self.requires_full_rendering)

    def allocate(self, only_visible=False, is_defusion=False):
        return self.o.L.rst_allocate_from_depth(self.e, C.byref(self.scene.c), C.byref(self.view), C.byref(self.trk),
                                                C.byref(self.rs.c), int(only_visible), int(is_defusion))

    def integrate(self):
        self.o.L.rst_integrate(self.e, C.byref(self.scene.c), C.byref(self.view), C.byref(self.trk), C.byref(self.rs.c))

    def deintegrate(self):
        self.o.L.rst_deintegrate(self.e, C.byref(self.scene.c), C.byref(self.view), C.byref(self.trk))

    def decay(self, max_weight, min_age):
        self.o.L.rst_decay(self.e, C.byref(self.scene.c), C.byref(self.rs.c), max_weight, min_age, 0)

    def decay_defusion(self, max_weight, min_age):
        self.o.L.rst_decay_defusion(self.e, C.byref(self.scene.c), C.byref(self.rs.c), max_weight, min_age, 0)

    def slide_window(self, max_age):
        self.o.L.rst_slide_window(self.e, C.byref(self.scene.c), C.byref(self.rs.c), max_age)

    def slide_window_defusion(self, max_age, max_size):
        self.o.L.rst_slide_window_defusion(self.e, C.byref(self.scene.c), C.byref(self.rs.c), max_age, max_size)

    def decayed_blocks(self):
        return int(self.o.L.rst_decayed_block_count(self.e))

    def history_sizes(self):
        v = [C.c_int() for _ in range(5)]
        self.o.L.rst_history_sizes(self.e, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def expected_depths(self):
        self.o.L.rst_create_expected_depths(C.byref(self.scene.c), self.trk.M_d, self.view.intr_d, C.byref(self.rs.c))

    def icp_maps(self):
        self.o.L.rst_create_icp_maps(C.byref(self.scene.c), C.byref(self.view), C.byref(self.trk), C.byref(self.rs.c))

    def process_frame(self):
        rc = self.allocate()
        self.integrate()
        self.expected_depths()
        self.icp_maps()
        return rc

    def forward_render(self):
        """ForwardRender at the current view/pose; returns (forwardProjection, noMissing, rendering)"""
        fwd = np.zeros((self.cfg.H, self.cfg.W, 4), np.float32)
        missing = np.zeros(self.cfg.H * self.cfg.W, np.int32)
        n = C.c_int()
        self.o.L.rst_forward_render(C.byref(self.scene.c), C.byref(self.view), C.byref(self.trk), C.byref(self.rs.c),
                                    po.ptr(fwd), po.ptr(missing), C.byref(n))
        return fwd, n.value, self.rs.raycast_image.copy()

    def free_view(self, M, intr, render_type=0):
        Mc, iM, ic = abi.f16(M), abi.f16(self.o.mat4_inverse(M)), abi.f4(intr)
        self.o.L.rst_find_visible_blocks(C.byref(self.scene.c), Mc, ic, C.byref(self.rs_free.c))
        self.o.L.rst_create_expected_depths(C.byref(self.scene.c), Mc, ic, C.byref(self.rs_free.c))
        self.o.L.rst_render_image(C.byref(self.scene.c), Mc, iM, ic, C.byref(self.rs_free.c), self.out_img.ctypes.data,
                                  self.out_depth.ctypes.data, render_type)

    def count_visible(self, lo, hi):
        return self.o.L.rst_count_visible_blocks(C.byref(self.scene.c), C.byref(self.rs.c), lo, hi)

    # ---- state
    def state(self, voxels=True, copy=True):
        """copy=False: views of the live arrays (the voxel array alone is 1 GiB at the reference capacities) for callers that
        compare at once and do not keep the state across a later call"""
        c = (lambda a: a.copy()) if copy else (lambda a: a)
        s = dict(hash=c(self.scene.hash), visible=self.rs.visible(), visible_type=c(self.rs.visible_type),
                 last_free=(self.scene.c.last_free_block, self.scene.c.last_free_excess),
                 alloc_list=c(self.scene.alloc_list), excess_list=c(self.scene.excess_list))
        if voxels:
            s["voxels"] = c(self.scene.voxels)
        return s

    def images(self):
        return dict(range=self.rs.range_image.copy(), raycast=self.rs.raycast_result.copy(), points=self.pts.copy(),
                    normals=self.nrm.copy(), rendering=self.rs.raycast_image.copy())

    def free_images(self):
        return dict(visible=self.rs_free.visible(), range=self.rs_free.range_image.copy(),
                    raycast=self.rs_free.raycast_result.copy(), image=self.out_img.copy(), depth=self.out_depth.copy())


class GpuBackend(object):
    name = "gpu"

    def __init__(self, cfg, caps=None, use_swapping=False):
        eng = importlib.import_module(PKG + ".engine")
        caps = caps or {}
        self.cfg, self.eng = cfg, eng
        nb = caps.get("num_blocks", abi.DEFAULT_LOCAL_BLOCKS)
        bn = caps.get("bucket_num", abi.DEFAULT_BUCKETS)
        ex = caps.get("excess_size", abi.DEFAULT_EXCESS)
        self.scene = eng.ITMScene(cfg.params(), nb, bn, ex, use_swapping)
        self.recon = eng.ITMSceneReconstructionEngine(nb, bn, ex)
        self.vis = eng.ITMVisualisationEngine(cfg.W, cfg.H)
        self.rs = self.vis.CreateRenderState(self.scene, (cfg.W, cfg.H))
        self.rs_free = self.vis.CreateRenderState(self.scene, (cfg.W, cfg.H))
        self.view = eng.ITMView(cfg.W, cfg.H, cfg.intr)
        self.trk = eng.ITMTrackingState(cfg.W, cfg.H)
        self.out_img = eng.DeviceBuffer(cfg.W * cfg.H * 4)
        self.out_depth = eng.DeviceBuffer(cfg.W * cfg.H * 4)
        self.out_img.upload(np.zeros(cfg.W * cfg.H * 4, np.uint8))
        self.out_depth.upload(np.zeros(cfg.W * cfg.H, np.float32))
        self.reset()

    def close(self):
        self.recon.close(); self.vis.close()
        for o in (self.scene, self.rs, self.rs_free, self.view, self.trk, self.out_img, self.out_depth):
            o.free()

    requires_full_rendering = True

    def reset(self):
        self.recon.ResetScene(self.scene)

    def set_weights(self, depth_weighting, max_new_w=10, max_distance=100):
        self.recon.SetFusionWeightParams(depth_weighting, max_new_w, max_distance)

    def set_frame(self, f, M=None):
        self.view.update(f.depth, f.rgb, f.timestamp)
        self.trk.set_pose(f.M if M is None else M)
        self.trk.c.requires_full_rendering = 1 if self.requires_full_rendering else 0

    def _rc(self, fn, *a, **k):
        try:
            fn(*a, **k)
            return 0
        except abi.Dsb200Error as e:
            if e.code in (abi.ERR_VBA_EXHAUSTED, abi.ERR_EXCESS_EXHAUSTED):
                return e.code
            raise

    def allocate(self, only_visible=False, is_defusion=False):
        return self._rc(self.recon.AllocateSceneFromDepth, self.scene, self.view, self.trk, self.rs, only_visible, is_defusion)

    def integrate(self):
        self.recon.IntegrateIntoScene(self.scene, self.view, self.trk, self.rs)

    def deintegrate(self):
        self.recon.DeIntegrateIntoScene(self.scene, self.view, self.trk)

    def decay(self, max_weight, min_age):
        self.recon.Decay(self.scene, self.rs, max_weight, min_age)

    def decay_defusion(self, max_weight, min_age):
        self.recon.DecayDefusionPart(self.scene, self.rs, max_weight, min_age)

    def slide_window(self, max_age):
        self.recon.SlideWindow(self.scene, self.rs, max_age)

    def slide_window_defusion(self, max_age, max_size):
        self.recon.SlideWindowDefusionPart(self.scene, self.rs, max_age, max_size)

    def decayed_blocks(self):
        return self.recon.GetDecayedBlockCount()

    def history_sizes(self):
        return self.recon.HistorySizes()

    def expected_depths(self):
        self.vis.CreateExpectedDepths(self.scene, self.trk.c.M_d, self.cfg.intr, self.rs)

    def icp_maps(self):
        self.vis.CreateICPMaps(self.scene, self.view, self.trk, self.rs)

    def process_frame(self):
        return self._rc(self.recon.ProcessFrame, self.vis, self.scene, self.view, self.trk, self.rs)

    def forward_render(self):
        self.vis.ForwardRender(self.scene, self.view, self.trk, self.rs)
        fwd, n = self.vis.ForwardProjection(self.cfg.W, self.cfg.H)
        return fwd, n, self.rs.raycast_image()

    def free_view(self, M, intr, render_type=0):
        lib = abi.load()
        Mc = abi.f16(M)
        iM = (C.c_float * 16)()
        lib.dsb200_mat4_inverse(Mc, iM)
        self.vis.FindVisibleBlocks(self.scene, M, intr, self.rs_free)
        self.vis.CreateExpectedDepths(self.scene, M, intr, self.rs_free)
        self.vis.RenderImage(self.scene, M, iM, intr, self.rs_free, self.out_img, self.out_depth, render_type)

    def count_visible(self, lo, hi):
        return self.vis.CountVisibleBlocks(self.scene, self.rs, lo, hi)

    def state(self, voxels=True):
        s = dict(hash=self.scene.hash(), visible=self.rs.visible(), visible_type=self.rs.visible_type(),
                 last_free=(self.scene.c.last_free_block, self.scene.c.last_free_excess),
                 alloc_list=self.scene.alloc_list(), excess_list=self.scene.excess_list())
        if voxels:
            s["voxels"] = self.scene.voxels()
        return s

    def images(self):
        return dict(range=self.rs.range_image(), raycast=self.rs.raycast_result(), points=self.trk.icp_points(),
                    normals=self.trk.icp_normals(), rendering=self.rs.raycast_image())

    def free_images(self):
        n = self.cfg.W * self.cfg.H
        return dict(visible=self.rs_free.visible(), range=self.rs_free.range_image(), raycast=self.rs_free.raycast_result(),
                    image=self.out_img.download(np.uint8, n * 4).reshape(self.cfg.H, self.cfg.W, 4),
                    depth=self.out_depth.download(np.float32, n).reshape(self.cfg.H, self.cfg.W))


class RefCudaBackend(object):
    """oracle/_ref/libitmref_cuda.so: the UNMODIFIED reference CUDA engines recompiled for sm_100 (IEEE build unless
    fast=True), behind the call names of the other back ends.  Fixed reference capacities.  Allocation there is racy
    (atomicSub free-list pops, last-writer-wins requests, per-CTA atomicAdd compaction), so its state is compared at SET
    level: see block_table() / assert_sets_match()."""
    name = "refcuda"

    def __init__(self, cfg, use_swapping=False, fast=False):
        self.cfg = cfg
        self.e = po.RefCudaEngine(cfg.W, cfg.H, cfg.params(), use_swapping=use_swapping, fast=fast)
        self.e.set_calib(cfg.intr)
        self.requires_full_rendering = True

    def close(self):
        self.e.close()

    def reset(self):
        self.e.reset()

    def set_weights(self, depth_weighting, max_new_w=10, max_distance=100):
        self.e.set_weights(depth_weighting, max_new_w, max_distance)

    def set_frame(self, f, M=None):
        self.e.set_view(f.depth, f.rgb, f.timestamp)
        self.e.set_pose(f.M if M is None else M)
        self.e.set_requires_full_rendering(self.requires_full_rendering)

    def allocate(self, only_visible=False, is_defusion=False):
        self.e.allocate(only_visible, is_defusion)
        return 0

    def integrate(self):
        self.e.integrate()

    def deintegrate(self):
        self.e.deintegrate()

    def decay(self, max_weight, min_age):
        self.e.decay(max_weight, min_age)

    def decay_defusion(self, max_weight, min_age):
        self.e.decay_defusion(max_weight, min_age)

    def slide_window(self, max_age):
        self.e.slide_window(max_age)

    def slide_window_defusion(self, max_age, max_size):
        self.e.slide_window_defusion(max_age, max_size)

    def decayed_blocks(self):
        return self.e.decayed_blocks()

    def history_sizes(self):
        return self.e.history_sizes()

    def expected_depths(self):
        self.e.create_expected_depths()

    def icp_maps(self):
        self.e.create_icp_maps()

    def process_frame(self):
        self.allocate(); self.integrate(); self.expected_depths(); self.icp_maps()
        return 0

    def free_view(self, M, intr, render_type=0):
        self.e.render_free_view(M, intr, render_type)

    def images(self):
        return dict(range=self.e.range_image(), raycast=self.e.raycast_result(), points=self.e.icp_points(),
                    normals=self.e.icp_normals(), rendering=self.e.icp_rendering())

    def block_table(self, voxels=True):
        h = self.e.hash()
        return make_block_table(h, self.e.visible(), (lambda ptrs: self.e.blocks(ptrs)) if voxels else None, self.e.last_free(),
                                self.e.num_blocks, allow_duplicates=True)


# ---------------------------------------------------------------- set-level state (reference CUDA engine vs ours)
def pos_key(pos):
    """(n, 3) int16 block coordinates -> sortable int64 keys"""
    p = pos.astype(np.int64) + 32768
    return (p[:, 0] << 32) | (p[:, 1] << 16) | p[:, 2]


def make_block_table(hash_table, visible_ids, fetch_blocks, last_free, num_blocks, allow_duplicates=False):
    """What is comparable between two engines whose slot / pointer assignment differs: the SET of allocated block
    coordinates, each block's 512 voxels, the SET of visible blocks, and the number of free voxel blocks.
    allow_duplicates: the reference CUDA allocator writes a request's block coordinates with four 16-bit stores
    (buildHashAllocAndVisibleType_device, CUDA .cu:1521-1531; SASS: 4 x STG.U16), so two pixels racing for one slot can leave a
    MIXED coordinate, which may coincide with a block that already exists: such duplicates are counted (`duplicates`), the first
    copy is kept.  Ours must never have one."""
    live = np.nonzero(hash_table["ptr"] >= 0)[0]
    keys = pos_key(hash_table["pos"][live])
    order = np.argsort(keys, kind="stable")
    keys, live = keys[order], live[order]
    first = np.ones(keys.size, bool)
    first[1:] = keys[1:] != keys[:-1]
    dup = int((~first).sum())
    assert allow_duplicates or dup == 0, "%d block coordinates are allocated twice" % dup
    allocated = int(keys.size)
    keys, live = keys[first], live[first]
    ptrs = hash_table["ptr"][live]
    # two live entries sharing one voxel block: the aftermath of a double free (the reference frees a block twice when a torn
    # coordinate left it with two entries for one coordinate; the block is then handed out twice).  Never in ours.
    up, cnt = np.unique(ptrs, return_counts=True)
    aliased = keys[np.isin(ptrs, up[cnt > 1])]
    assert allow_duplicates or aliased.size == 0, "%d live entries share a voxel block" % aliased.size
    t = dict(keys=keys, slots=live, ptrs=ptrs, free_blocks=last_free[0] + 1, num_blocks=num_blocks,
             duplicates=dup, allocated=allocated, aliased=aliased)
    vis = np.asarray(visible_ids)
    t["visible_keys"] = np.sort(pos_key(hash_table["pos"][vis])) if vis.size else np.zeros(0, np.int64)
    t["visible_dead"] = int((hash_table["ptr"][vis] < 0).sum()) if vis.size else 0
    if fetch_blocks is not None:
        t["voxels"] = fetch_blocks(t["ptrs"])
    return t


def gpu_block_table(gpu, voxels=True):
    """the same for a GpuBackend / OracleBackend state()"""
    s = gpu.state(voxels=voxels)
    fetch = (lambda ptrs: s["voxels"].reshape(-1, abi.BLOCK_SIZE3)[ptrs]) if voxels else None
    return make_block_table(s["hash"], s["visible"], fetch, s["last_free"], gpu.scene.c.num_blocks)


class SetTracker(object):
    """Follows two engines frame by frame and decides which blocks are COMPARABLE voxel by voxel:
      * both engines allocated the block in the same frame (its `birth`): the reference CUDA allocator resolves same-frame bucket
        collisions by a race (CUDA .cu:1521-1531 writes entriesAllocType / blockCoords last-writer-wins; the loser is picked up a
        frame later, DeviceAgnostic/ITMSceneReconstructionEngine.h:483), the serial CPU engine by scan order, so a few blocks are
        born a frame apart and have seen one observation less;
      * the block never shared its hash bucket with another live block (`chained`): which slot of the excess list a chain member
        gets, and which member moves up when the head is deleted (deleteBlock, CUDA .cu:568-586), depend on the order the
        reference's atomics happened to run in; history snapshots hold SLOT ids, so a pruning pass aimed at a slot reaches whichever
        block lives there at that time -- chain members can be pruned a pass earlier or later than in the serial order;
      * the reference's deleteBlock skips a deletion when two blocks of one bucket die in the same launch (try-lock, .cu:510-515).
    Blocks born in different frames are counted and bounded, not compared; every other block is compared voxel by voxel, and a
    difference is accepted only in a `chained` block (and only in a bounded number of them)."""

    def __init__(self, bucket_mask=abi.DEFAULT_BUCKETS - 1):
        self.birth = [dict(), dict()]
        self.chained = set()
        self.aliased = set()          # reference blocks that ever shared their voxel block with another entry (reference defect)
        self.frame = 0
        self.mask = bucket_mask

    def _buckets(self, keys):
        k = keys.astype(np.int64)
        x = ((k >> 32) & 0xffff) - 32768
        y = ((k >> 16) & 0xffff) - 32768
        z = (k & 0xffff) - 32768
        h = ((x.astype(np.uint32) * np.uint32(73856093)) ^ (y.astype(np.uint32) * np.uint32(19349669)) ^ (z.astype(np.uint32) * np.uint32(83492791)))
        return (h & np.uint32(self.mask)).astype(np.int64)

    def update(self, ta, tb):
        for side, t in enumerate((ta, tb)):
            now = set(t["keys"].tolist())
            b = self.birth[side]
            for k in list(b.keys()):
                if k not in now:
                    del b[k]
            for k in now:
                if k not in b:
                    b[k] = self.frame
        union = np.union1d(ta["keys"], tb["keys"])
        alive = set(union.tolist())
        self.chained &= alive                       # a block that died in both engines starts afresh
        self.aliased = (self.aliased | set(tb.get("aliased", np.zeros(0, np.int64)).tolist())) & set(tb["keys"].tolist())
        if union.size:
            bk = self._buckets(union)
            order = np.argsort(bk, kind="stable")
            sb = bk[order]
            dup = np.zeros(sb.size, bool)
            dup[1:] |= sb[1:] == sb[:-1]
            dup[:-1] |= sb[1:] == sb[:-1]
            self.chained |= set(union[order][dup].tolist())
        self.frame += 1

    def comparable(self, keys):
        a, b = self.birth
        return np.array([(k in a and k in b and a[k] == b[k] and k not in self.aliased) for k in keys.tolist()], bool)

    def is_chained(self, keys):
        return np.array([k in self.chained for k in keys.tolist()], bool)


def assert_sets_match(tracker, ours, ref, what, max_odd_fraction=0.03, voxels=True, max_chain_mismatch=0.01):
    """ours / ref: block tables taken after the same call sequence.  Returns the number of non-comparable blocks."""
    tracker.update(ours, ref)
    both = np.intersect1d(ours["keys"], ref["keys"])
    union = np.union1d(ours["keys"], ref["keys"])
    ok = tracker.comparable(both)
    odd = union.size - int(ok.sum()) + ref["duplicates"]
    only = np.setdiff1d(ours["keys"], ref["keys"]).size + np.setdiff1d(ref["keys"], ours["keys"]).size
    assert odd <= max(8, max_odd_fraction * union.size), "%s: %d of %d blocks not comparable (only ours %d, only reference %d)" % (
        what, odd, union.size, np.setdiff1d(ours["keys"], ref["keys"]).size, np.setdiff1d(ref["keys"], ours["keys"]).size)
    assert only <= max(8, 0.005 * union.size), "%s: the block sets differ in %d of %d blocks" % (what, only, union.size)
    # free-list accounting: allocated + free == capacity.  Exact for ours; the reference may double-free a block when a torn
    # coordinate write left it with two entries for one coordinate (see make_block_table): bounded, reported
    assert ours["allocated"] + ours["free_blocks"] == ours["num_blocks"], "%s: ours leaks voxel blocks (%d allocated + %d free != %d)" % (
        what, ours["allocated"], ours["free_blocks"], ours["num_blocks"])
    ref_leak = ref["allocated"] + ref["free_blocks"] - ref["num_blocks"]
    assert abs(ref_leak) <= 8, "%s: reference free-list accounting off by %d" % (what, ref_leak)
    # visible SET (the reference's list order is schedule-dependent)
    vdiff = np.setxor1d(ours["visible_keys"], ref["visible_keys"])
    assert vdiff.size <= odd + ours["visible_dead"] + ref["visible_dead"], "%s: visible sets differ in %d blocks (%d non-comparable)" % (what, vdiff.size, odd)
    if voxels and both.size:
        ia = np.searchsorted(ours["keys"], both[ok]); ib = np.searchsorted(ref["keys"], both[ok])
        va, vb = ours["voxels"][ia], ref["voxels"][ib]
        m = np.uint64(0x00ffffffffffffff)         # 8th byte = padding
        ne = ((va.view(np.uint64) & m) != (vb.view(np.uint64) & m)).reshape(len(ia), -1)
        bad = ne.any(axis=1)
        if bad.any():
            chained = tracker.is_chained(both[ok])
            loose = np.nonzero(bad & ~chained)[0]       # a differing block that never shared a bucket: a real disagreement
            # (the reference's races are real races: in roughly one run out of three a single block outside any chain comes out
            # pruned one pass apart; two such blocks are tolerated, reported through tracker.loose_mismatches)
            tracker.loose_mismatches = max(getattr(tracker, "loose_mismatches", 0), int(loose.size))
            if loose.size > 2 or int(bad.sum()) > max(4, max_chain_mismatch * len(ia)):
                blk = loose if loose.size else np.nonzero(bad)[0]
                d = np.abs(va["sdf"].astype(np.int32) - vb["sdf"].astype(np.int32)).max()
                k0 = int(both[ok][blk[0]])
                raise AssertionError("%s: %d of %d comparable blocks differ (%d of them outside collision chains), %d voxels, max |dsdf| %d LSB; "
                                     "first block key %d (bucket %d, slot ours %d / reference %d, born frame %s, %d voxels): %s vs %s" % (
                                         what, int(bad.sum()), len(ia), loose.size, int(ne.sum()), int(d), k0,
                                         int(tracker._buckets(np.array([k0]))[0]), int(ours["slots"][ia[blk[0]]]), int(ref["slots"][ib[blk[0]]]),
                                         tracker.birth[0].get(k0), int(ne[blk[0]].sum()), va[blk[0]][ne[blk[0]]][:3], vb[blk[0]][ne[blk[0]]][:3]))
            tracker.chain_mismatches = max(getattr(tracker, "chain_mismatches", 0), int(bad.sum()))
    return odd


# ---------------------------------------------------------------- comparators
def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_state_equal(a, b, what="", voxels=True, free_lists=True):
    """Bit-exact: hash (pos, offset, ptr; allocated_time excluded per SURVEY A9), visible list,
    visibility flags, free-list tops (+ the live part of the free lists), voxels."""
    for k in ("pos", "offset", "ptr"):
        bad = np.nonzero(np.any(a["hash"][k] != b["hash"][k], axis=-1) if k == "pos" else (a["hash"][k] != b["hash"][k]))[0]
        assert bad.size == 0, "%s hash.%s differs at %d slots, first %s: %s vs %s" % (
            what, k, bad.size, bad[:5], a["hash"][k][bad[:5]], b["hash"][k][bad[:5]])
    assert a["last_free"] == b["last_free"], "%s free-list tops %s vs %s" % (what, a["last_free"], b["last_free"])
    assert np.array_equal(a["visible"], b["visible"]), "%s visible list differs (%d vs %d)" % (what, len(a["visible"]), len(b["visible"]))
    assert np.array_equal(a["visible_type"], b["visible_type"]), "%s visible_type differs at %d slots" % (
        what, int((a["visible_type"] != b["visible_type"]).sum()))
    if free_lists:
        n = a["last_free"][0] + 1
        assert np.array_equal(a["alloc_list"][:max(n, 0)], b["alloc_list"][:max(n, 0)]), "%s VBA free list differs" % what
    if voxels and "voxels" in a and "voxels" in b:
        va, vb = a["voxels"].view(np.uint64), b["voxels"].view(np.uint64)
        if not np.array_equal(va, vb):
            bad = np.nonzero(va != vb)[0]
            d = np.abs(a["voxels"]["sdf"][bad].astype(np.int32) - b["voxels"]["sdf"][bad].astype(np.int32))
            raise AssertionError("%s voxels differ at %d of %d (max |dsdf| %d LSB), first idx %s: %s vs %s" % (
                what, bad.size, va.size, int(d.max()), bad[:3], a["voxels"][bad[:3]], b["voxels"][bad[:3]]))


def assert_images_equal(a, b, what=""):
    for k in sorted(a.keys()):
        x, y = bits(a[k]), bits(b[k])
        if not np.array_equal(x, y):
            bad = np.nonzero(x.reshape(-1) != y.reshape(-1))[0]
            raise AssertionError("%s image '%s' differs at %d of %d elements; first %s: %s vs %s" % (
                what, k, bad.size, x.size, bad[:4], a[k].reshape(-1)[bad[:4]], b[k].reshape(-1)[bad[:4]]))


# ---------------------------------------------------------------- sharded scene
class OracleShardedScene(object):
    """CPU restatement of the sharded scene: `count` oracle shards with the ownership filter, each running the
    unchanged serial pipeline on its own table; nearest-hit composite; ICP maps on the composite."""

    def __init__(self, cfg, caps, count, brick_log2=2, ranks=None, halo=False, mode=None):
        mode = mode or ("composite" if halo else "lookup")
        halo = halo or mode == "composite"
        self.cfg, self.count, self.mode, self.brick_log2 = cfg, count, mode, brick_log2
        self.ranks = list(range(count)) if ranks is None else list(ranks)     # the shards THIS process restates
        self.members = []
        for r in self.ranks:
            m = OracleBackend(cfg, caps)
            sh = abi.Shard(r, count, brick_log2, 1 if halo else 0)
            m.o.L.rst_set_shard(m.e, C.byref(sh))
            self.members.append(m)
        self.composite = np.zeros((cfg.H, cfg.W, 4), np.float32)
        self.merged_range = np.zeros((cfg.H, cfg.W, 2), np.float32)

    def fuse(self, f):
        for m in self.members:
            m.set_frame(f)
            m.allocate(); m.integrate(); m.expected_depths()

    def fuse_and_raycast(self, f):
        self.fuse(f)
        for m in self.members:
            m.o.L.rst_find_surface(C.byref(m.scene.c), m.trk.invM_d, m.view.intr_d, C.byref(m.rs.c))

    def raycast_lookup_and_icp(self, scenes=None, ranges=None):
        """whole-scene raycast with owner-routed reads; scenes / ranges default to this process's members (all shards)"""
        m0 = self.members[0]
        scenes = [m.scene.c for m in self.members] if scenes is None else scenes
        ranges = [m.rs.range_image for m in self.members] if ranges is None else ranges
        sp = (C.c_void_p * len(scenes))(*[C.addressof(s) for s in scenes])
        rp = (C.c_void_p * len(ranges))(*[po.ptr(r) for r in ranges])
        m0.o.L.rst_raycast_sharded(len(scenes), sp, rp, self.brick_log2, m0.trk.invM_d, m0.view.intr_d, self.cfg.W, self.cfg.H,
                                   po.ptr(self.merged_range), po.ptr(self.composite))
        m0.o.L.rst_icp_maps_from_raycast(C.byref(m0.scene.c), C.byref(m0.view), C.byref(m0.trk), po.ptr(self.composite),
                                         po.ptr(m0.rs.raycast_image), self.cfg.W, self.cfg.H)

    def composite_and_icp(self, results=None):
        """results: list of (H, W, 4) float32 raycast images in shard order (defaults to this process's members)"""
        m0 = self.members[0]
        results = [np.ascontiguousarray(m.rs.raycast_result) for m in self.members] if results is None else results
        arr = (C.c_void_p * len(results))(*[po.ptr(r) for r in results])
        m0.o.L.rst_composite_nearest(len(results), arr, m0.trk.invM_d, self.cfg.params().voxel_size, self.cfg.W, self.cfg.H,
                                     0, self.cfg.H, po.ptr(self.composite))
        m0.o.L.rst_icp_maps_from_raycast(C.byref(m0.scene.c), C.byref(m0.view), C.byref(m0.trk), po.ptr(self.composite),
                                         po.ptr(m0.rs.raycast_image), self.cfg.W, self.cfg.H)

    def process_frame(self, f):
        if self.mode == "composite":
            self.fuse_and_raycast(f)
            self.composite_and_icp()
        else:
            self.fuse(f)
            self.raycast_lookup_and_icp()

    def images(self):
        m0 = self.members[0]
        return dict(composite=self.composite.copy(), points=m0.pts.copy(), normals=m0.nrm.copy(), rendering=m0.rs.raycast_image.copy())

    def close(self):
        for m in self.members:
            m.close()


def gpu_member_state(m):
    """state dict of a sharded.ShardMember in the layout of GpuBackend.state()"""
    return dict(hash=m.scene.hash(), visible=m.rs.visible(), visible_type=m.rs.visible_type(),
                last_free=(m.scene.c.last_free_block, m.scene.c.last_free_excess), alloc_list=m.scene.alloc_list(),
                excess_list=m.scene.excess_list(), voxels=m.scene.voxels())


def allocated_blocks(hash_table):
    """set of (x, y, z) of the live entries of a hash table download"""
    live = hash_table[hash_table["ptr"] >= 0]
    return set(map(tuple, live["pos"].tolist()))


# ---------------------------------------------------------------- view building
def view_from_raw_restated(depth_i16, bgr, scale=0.001, offset=0.0):
    """numpy restatement of convertDepthAffineToFloat (Engine/DeviceAgnostic/ITMViewBuilder.h:22-28) and of the driver's
    BGR -> RGBA loop (DS/InfiniTamDriver.cpp:84-104); float32 multiply then add, as the reference evaluates it"""
    d = np.asarray(depth_i16, np.int16)
    depth = np.where((d <= 0) | (d > 32000), np.float32(-1.0), d.astype(np.float32) * np.float32(scale) + np.float32(offset)).astype(np.float32)
    rgba = None
    if bgr is not None:
        b = np.asarray(bgr, np.uint8)
        rgba = np.concatenate([b[..., ::-1], np.full(b.shape[:-1] + (1,), 255, np.uint8)], axis=-1)
    return depth, rgba


def normal_and_weight_cases():
    """depth images and intrinsics for ComputeNormalAndWeights: rendered frames (with holes), a noisy wall seen at grazing angles and
    white noise, each with the focal length as the reference passes it (fx) and as a pinhole un-projection would (1 / fx) so that
    acos sees its three branches (|n.z| < 0.5, n.z > 0.5, n.z < -0.5)."""
    rng = np.random.default_rng(11)
    out = []
    for name in ("tiny", "tiny_street"):
        cfg, frames = synth.make_sequence(name, 1)
        d = frames[0].depth.reshape(cfg.H, cfg.W).copy()
        fx, fy, cx, cy = [np.float32(v) for v in cfg.intr]
        out.append((name, d, (fx, fy, cx, cy)))
        out.append((name + " 1/f", d, (np.float32(1) / fx, np.float32(1) / fy, cx, cy)))
    H, W = 61, 83
    u = np.arange(W, dtype=np.float32)[None, :]
    for slope in (0.0005, 0.004, 0.02, -0.03):
        wall = (np.float32(1.5) + np.float32(slope) * u + rng.normal(0, 0.002, (H, W)).astype(np.float32)).astype(np.float32)
        wall[rng.random((H, W)) < 0.02] = -1.0
        out.append(("wall %g" % slope, wall, (np.float32(1 / 500.0), np.float32(1 / 500.0), np.float32(W / 2), np.float32(H / 2))))
    noise = rng.uniform(0.3, 4.0, (H, W)).astype(np.float32)
    noise[rng.random((H, W)) < 0.05] = -1.0
    noise[10:14, 10:14] = 2.0                                   # flat patch: zero cross product -> rejected
    out.append(("noise", noise, (np.float32(525.0), np.float32(525.0), np.float32(41.5), np.float32(30.5))))
    out.append(("noise 1/f", noise, (np.float32(1 / 525.0), np.float32(1 / 525.0), np.float32(41.5), np.float32(30.5))))
    return out


def swap_scenario():
    """frames and their order for the swapping tests: two stretches of the tiny room trajectory played forwards and then
    backwards (blocks leave the frustum, are swapped out, come back and are combined with their stored copy), and the index in
    the order after which the whole map is swapped out as well (SaveToGlobalMemory without a render state, DS/DenseSlam.h:248-251)"""
    cfg, a = synth.make_sequence("tiny", 3, start=0)
    _, b = synth.make_sequence("tiny", 3, start=30)
    frames = a + b
    order = list(range(len(frames))) + list(range(len(frames) - 2, -1, -1))
    return cfg, frames, order, 7


def voxels_equal_but_pad(a, b):
    m = np.uint64(0x00ffffffffffffff)      # the 8th byte of ITMVoxel_s_rgb is compiler padding
    return np.array_equal(a.view(np.uint64) & m, b.view(np.uint64) & m)


class SwapOnOracle(object):
    """the swapping calls of the oracle behind the method names SwapOnGpu has (for checks that run on either backend)"""

    def __init__(self, back, transfer=0x1000):
        self.back = back
        self.cache = po.HostSwapCache(back.scene.c.total_entries, transfer)

    def mark(self):
        self.back.o.swap_mark_visible(self.cache, self.back.rs)

    def swap_in(self):
        return self.back.o.swap_integrate_global_into_local(self.cache, self.back.scene)

    def save(self, with_render_state):
        return self.back.o.swap_save_to_global_memory(self.cache, self.back.scene, self.back.rs if with_render_state else None)

    def has_stored(self):
        return self.cache.has_stored.copy()

    def states(self):
        return self.cache.states.copy()

    def stored_block(self, entry):
        return self.cache.stored_block(entry).copy()

    def close(self):
        pass


class SwapOnGpu(object):
    def __init__(self, back, transfer=0x1000):
        self.back = back
        self.swap = back.eng.ITMSwappingEngine(back.scene, transfer, recon=back.recon)

    def mark(self):
        self.swap.MarkVisible(self.back.rs)

    def swap_in(self):
        return self.swap.IntegrateGlobalIntoLocal(self.back.scene)

    def save(self, with_render_state):
        return self.swap.SaveToGlobalMemory(self.back.scene, self.back.rs if with_render_state else None)

    def has_stored(self):
        return self.swap.has_stored()

    def states(self):
        return self.swap.states()

    def stored_block(self, entry):
        return self.swap.stored_block(entry)

    def close(self):
        self.swap.close()


def check_swap_round_trip(back, sw, frames, min_blocks):
    """Size-independent properties of a whole-map swap-out followed by the swap-in of the visible part (any backend, any size):
    (1) every block whose entry is in state 2 reaches the host store and its hash pointer becomes -1 (entries the allocator put
        on the excess list in the last frame were never visible, are still in state 0 and stay); the freed part of the volume is
        back to its reset state, the rest is untouched, and free list + remaining pointers are a permutation of the block indices;
    (2) after the next allocation has handed blocks back to the visible entries, IntegrateGlobalIntoLocal restores each of them to
        its stored copy: combining a block of weight w with an empty one is (w * x) / w, so weights come back exactly and the
        quantised sdf / colours within 1 LSB."""
    nb = back.scene.c.num_blocks
    for f in frames:
        back.set_frame(f)
        assert back.allocate() == 0
        sw.mark(); back.integrate()
        while sw.swap_in():                  # until no entry is left in state 1 (a call moves at most the transfer limit)
            pass
    before = back.state()
    live = np.nonzero(before["hash"]["ptr"] >= 0)[0]
    moving = live[sw.states()[live] == 2]
    staying = live[sw.states()[live] != 2]
    assert moving.size >= min_blocks and staying.size < moving.size
    total_out = 0
    while True:
        n = sw.save(False)
        if n == 0:
            break
        total_out += n
    assert total_out == moving.size
    out = back.state()
    assert (out["hash"]["ptr"][moving] == -1).all()
    assert np.array_equal(out["hash"]["ptr"][staying], before["hash"]["ptr"][staying])
    assert np.array_equal(np.nonzero(out["hash"]["ptr"] >= 0)[0], staying)
    assert out["last_free"][0] == nb - 1 - staying.size
    held = np.concatenate([out["alloc_list"][:out["last_free"][0] + 1], out["hash"]["ptr"][staying]])
    assert np.array_equal(np.sort(held), np.arange(nb, dtype=np.int32))
    reset = np.ones(nb, bool); reset[before["hash"]["ptr"][staying]] = False
    out_blocks, before_blocks = out["voxels"].reshape(-1, abi.BLOCK_SIZE3), before["voxels"].reshape(-1, abi.BLOCK_SIZE3)
    assert (out_blocks[reset].view(np.uint64) == np.uint64(0x7fff)).all()
    assert np.array_equal(out_blocks[~reset].view(np.uint64), before_blocks[~reset].view(np.uint64))
    has = sw.has_stored()
    assert np.array_equal(np.nonzero(has)[0], moving) and (sw.states()[moving] == 0).all()
    for e in moving[::max(1, moving.size // 200)]:
        assert np.array_equal(sw.stored_block(e).view(np.uint64), before_blocks[before["hash"]["ptr"][e]].view(np.uint64)), "stored block of entry %d" % e
    # (2) bring the visible part back: allocation only (no integration), swap-state pass, swap-in until nothing is left in state 1
    back.set_frame(frames[-1])
    assert back.allocate() == 0
    sw.mark()
    total_in = 0
    while True:
        n = sw.swap_in()
        if n == 0:
            break
        total_in += n
    after = back.state()
    visible = after["visible"]
    assert total_in >= visible.size - staying.size and visible.size >= min_blocks // 4
    after_blocks = after["voxels"].reshape(-1, abi.BLOCK_SIZE3)
    checked = 0
    for e in visible[::max(1, visible.size // 400)]:
        if not has[e]:
            continue
        got, want = after_blocks[after["hash"]["ptr"][e]], sw.stored_block(e)
        assert np.array_equal(got["w_depth"], want["w_depth"]) and np.array_equal(got["w_color"], want["w_color"]), "weights of entry %d" % e
        assert np.abs(got["sdf"].astype(np.int32) - want["sdf"].astype(np.int32)).max() <= 1, "sdf of entry %d" % e
        assert np.abs(got["clr"].astype(np.int32) - want["clr"].astype(np.int32)).max() <= 1, "colour of entry %d" % e
        checked += 1
    assert checked >= 20
    return moving.size, visible.size

