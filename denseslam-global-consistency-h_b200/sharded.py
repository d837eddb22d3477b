"""Sharded scene: one large voxel-block-hash map partitioned by block-coordinate ownership.

The reference is single-GPU; this is the multi-GPU row of the hot path.  Every shard owns the bricks
`dsb200_shard_owner(block) == rank` and runs the unchanged per-scene pipeline on them:

    frame (depth, rgb, pose)  --broadcast-->  every shard
    shard r:  allocate (own blocks only) -> integrate -> expected depths                 [no communication]
    raycast ("lookup" mode, default): shard r marches the image tiles t % count == r through the WHOLE scene -- each
        block probe goes to the owner's hash table / voxel array (peer memory over NVLink) -- and stores its tiles into
        the root's image.  Same march as one big scene, so the image is the single-scene image bit for bit.
    raycast ("composite" mode): every shard raycasts only what it holds (+1 halo of neighbour blocks), the root keeps
        the nearest hit per pixel.  No remote probes, but replicated halo blocks and rays that cross shard borders
        deviate from the single-scene march (tests/test_sharded.py quantifies it).
    ICP maps (image-space stencil) on the root

`ShardMember` is one shard (engines + scene + double-buffered raycast result).  `LocalShardedScene` puts all
members on ONE GPU (parity tests, and a capacity-free functional check); `DistShardedScene` is one member per
process/GPU under torch.distributed: NCCL broadcasts the frame, the raycast results are exchanged by mapping the
peers' buffers with CUDA IPC and reading them in place over NVLink inside the composite kernel
(dsb200_composite_nearest) -- a stream-ordered NCCL all-reduce of one int is the only synchronisation.
"""
import ctypes as C

import numpy as np

from . import abi
from . import engine as eng


def _lib():
    return abi.load()


def shard_caps(caps, count, brick_log2=2, halo=False, margin=1.25):
    """Per-shard capacities for a scene of `caps` split over `count` shards (capacity / N + halo): the voxel block
    array and the excess list shrink with N (times the halo replication ((B + 1) / B)^3 of bricks of B blocks, times a margin for
    the imbalance of the ownership hash); the bucket count is the next power of two."""
    if count <= 1:
        return dict(caps)
    rep = ((1 << brick_log2) + 1.0) ** 3 / float(1 << brick_log2) ** 3 if halo else 1.0
    f = min(1.0, rep * margin / count)
    nb = max(1024, int(caps["num_blocks"] * f))
    # buckets: twice the proportional share, rounded up to a power of two -- hashIndex clusters for small block coordinates, and a
    # shard whose table is loaded like the unsharded one walks longer chains in every probe (measured at N = 8: 2 Mi buckets for
    # 565 k blocks cost 30 % of the frame rate against 4 Mi)
    bn = 1
    while bn < max(1024, int(2 * caps["bucket_num"] * f)):
        bn <<= 1
    return dict(num_blocks=nb, bucket_num=bn, excess_size=max(1024, int(caps["excess_size"] * f)))


def row_slice(height, rank, count):
    """rows [begin, end) of the image that shard `rank` composites (reduce-scatter shape: every shard reads its slice of all
    the shards' results and writes the nearest hits into the root's image)"""
    return (height * rank) // count, (height * (rank + 1)) // count


class ShardMember(object):
    def __init__(self, cfg, caps, rank, count, brick_log2=2, use_swapping=False, halo=True):
        self.cfg, self.rank, self.count, self.brick_log2 = cfg, rank, count, brick_log2
        nb, bn, ex = caps["num_blocks"], caps["bucket_num"], caps["excess_size"]
        self.scene = eng.ITMScene(cfg.params(), nb, bn, ex, use_swapping)
        self.recon = eng.ITMSceneReconstructionEngine(nb, bn, ex)
        self.recon.SetShard(rank, count, brick_log2, halo)
        self.vis = eng.ITMVisualisationEngine(cfg.W, cfg.H)
        self.rs = self.vis.CreateRenderState(self.scene, (cfg.W, cfg.H))
        self.view = eng.ITMView(cfg.W, cfg.H, cfg.intr)
        self.trk = eng.ITMTrackingState(cfg.W, cfg.H)
        npx = cfg.W * cfg.H
        # the render state's own raycast buffer + one more: frame f raycasts into ray[f & 1], so a peer may still be
        # reading frame f while this shard already raycasts frame f + 1
        self.ray = [self.rs.c.raycast_result, eng.DeviceBuffer(npx * 16).ptr]
        self.composite = eng.DeviceBuffer(npx * 16)       # the whole scene's raycast image (root)
        self.merged_range = eng.DeviceBuffer(npx * 8)
        self.recon.ResetScene(self.scene)

    def fuse_and_raycast(self, frame_idx, stages=abi.FRAME_ALLOCATE | abi.FRAME_INTEGRATE | abi.FRAME_EXPECTED_DEPTHS | abi.FRAME_RAYCAST):
        """own-block allocation, integration, expected depths and raycast of the current view/pose"""
        self.rs.c.raycast_result = self.ray[frame_idx & 1]
        self.recon.ProcessFrame(self.vis, self.scene, self.view, self.trk, self.rs, stages)

    def composite_rows(self, result_ptrs, rows=None, out_ptr=None):
        """nearest-hit composite of rows [r0, r1) of the shards' raycast results (device or IPC-mapped peer pointers) into
        out_ptr (default: this shard's composite image; the root's image, peer-mapped, in the distributed scene)"""
        n = len(result_ptrs)
        arr = (C.c_void_p * n)(*result_ptrs)
        r0, r1 = rows if rows is not None else (0, self.cfg.H)
        abi.check(_lib().dsb200_composite_nearest(self.vis.h, n, arr, self.trk.c.invM_d, self.cfg.params().voxel_size,
                                                  self.cfg.W, self.cfg.H, r0, r1, out_ptr if out_ptr is not None else self.composite.ptr))

    def composite_and_icp(self, result_ptrs, rows=None):
        """whole-image composite on this shard, then ICP maps"""
        self.composite_rows(result_ptrs, rows)
        self.icp_from_composite()

    def tables(self):
        """(hash, voxels, range image) device pointers of this shard, for dsb200_shard_tables"""
        return self.scene.c.hash, self.scene.c.voxels, self.rs.c.range_image

    def raycast_lookup(self, tables, out_ptr):
        """march this shard's image tiles through the whole sharded scene; tables = [(hash, voxels, range)] in shard order"""
        t = abi.ShardTables()
        t.count, t.rank, t.brick_log2 = self.count, self.rank, self.brick_log2
        for i, (h, v, r) in enumerate(tables):
            t.hash[i], t.voxels[i], t.range_image[i] = h, v, r
        abi.check(_lib().dsb200_raycast_sharded(self.vis.h, C.byref(self.scene.c), C.byref(t), self.trk.c.invM_d, self.view.c.intr_d,
                                                self.cfg.W, self.cfg.H, self.merged_range.ptr, out_ptr))

    def icp_from_composite(self):
        abi.check(_lib().dsb200_icp_maps_from_raycast(self.vis.h, C.byref(self.scene.c), C.byref(self.view.c), C.byref(self.trk.c),
                                                      self.composite.ptr, self.rs.c.raycast_image, self.cfg.W, self.cfg.H))

    def composite_image(self):
        return self.composite.download(np.float32, self.cfg.W * self.cfg.H * 4).reshape(self.cfg.H, self.cfg.W, 4)

    def close(self):
        self.rs.c.raycast_result = self.ray[0]
        self.recon.close(); self.vis.close()
        _lib().dsb200_device_free(self.ray[1])
        for o in (self.scene, self.rs, self.view, self.trk, self.composite, self.merged_range):
            o.free()


class LocalShardedScene(object):
    """All shards of one scene on the current GPU (functional twin of DistShardedScene without the transport)."""

    def __init__(self, cfg, caps, count, brick_log2=2, mode="lookup"):
        self.cfg, self.count, self.mode = cfg, count, mode
        self.members = [ShardMember(cfg, caps, r, count, brick_log2, halo=(mode == "composite")) for r in range(count)]
        self.frame_idx = 0

    def process_frame(self, f):
        local = abi.FRAME_ALLOCATE | abi.FRAME_INTEGRATE | abi.FRAME_EXPECTED_DEPTHS
        for m in self.members:
            m.view.update(f.depth, f.rgb, f.timestamp)
            m.trk.set_pose(f.M)
            m.fuse_and_raycast(self.frame_idx, local | (abi.FRAME_RAYCAST if self.mode == "composite" else 0))
        root = self.members[0]
        if self.mode == "composite":
            ptrs = [m.ray[self.frame_idx & 1] for m in self.members]
            for m in self.members:           # every shard composites its row slice into the root's image
                m.trk.set_pose(f.M)
                m.composite_rows(ptrs, row_slice(self.cfg.H, m.rank, self.count), root.composite.ptr)
            root.icp_from_composite()
        else:
            tables = [m.tables() for m in self.members]
            for m in self.members:
                m.raycast_lookup(tables, root.composite.ptr)
            root.icp_from_composite()
        abi.check(_lib().dsb200_stream_synchronize(None))
        self.frame_idx += 1

    def close(self):
        for m in self.members:
            m.close()


class DistShardedScene(object):
    """One shard per process / GPU under torch.distributed (backend nccl).  `root` holds the camera stream."""

    def __init__(self, cfg, caps, brick_log2=2, root=0, mode="lookup"):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank, self.world, self.root = dist.get_rank(), dist.get_world_size(), root
        self.cfg = cfg
        self.mode = mode
        self.m = ShardMember(cfg, caps, self.rank, self.world, brick_log2, halo=(mode == "composite"))
        self.dev = torch.device("cuda", torch.cuda.current_device())
        npx = cfg.W * cfg.H
        # one packed frame buffer = one broadcast: depth f32 | rgb u8x4 | pose M f32[16] | timestamp f64 (as 2 x f32 slots)
        self.frame_buf = torch.empty(npx * 2 + 16 + 2, dtype=torch.float32, device=self.dev)
        self.m.view.c.depth = self.frame_buf.data_ptr()
        self.m.view.c.rgb = self.frame_buf.data_ptr() + npx * 4
        self.token = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.frame_idx = 0
        # exchange CUDA IPC handles: [ray0, ray1, hash, voxels, range image, root image]; peers[r][k] = mapped pointer
        mine = [self.m.ray[0], self.m.ray[1], self.m.scene.c.hash, self.m.scene.c.voxels, self.m.rs.c.range_image,
                self.m.composite.ptr]
        handles = []
        for p in mine:
            h = C.create_string_buffer(abi.IPC_HANDLE_BYTES)
            abi.check(_lib().dsb200_ipc_export(p, h))
            handles.append(h.raw)
        gathered = [None] * self.world
        dist.all_gather_object(gathered, handles)
        self.peers, self._opened = [], []
        for r in range(self.world):
            if r == self.rank:
                self.peers.append(mine)
                continue
            row = []
            for k in range(len(mine)):
                p = C.c_void_p()
                abi.check(_lib().dsb200_ipc_open(gathered[r][k], C.byref(p)))
                row.append(p.value)
                self._opened.append(p.value)
            self.peers.append(row)
        self.peer_ray = [[self.peers[r][b] for r in range(self.world)] for b in range(2)]
        self.tables = [(self.peers[r][2], self.peers[r][3], self.peers[r][4]) for r in range(self.world)]
        self.root_image = self.peers[root][5]

    def pack_frame(self, f):
        """root only: host frame -> the packed device buffer (outside the sharded path proper)"""
        t = self.torch
        npx = self.cfg.W * self.cfg.H
        host = np.empty(npx * 2 + 18, np.float32)
        host[:npx] = np.asarray(f.depth, np.float32).reshape(-1)
        host[npx:2 * npx] = np.ascontiguousarray(f.rgb, np.uint8).reshape(-1).view(np.float32)
        host[2 * npx:2 * npx + 16] = np.asarray(f.M, np.float32).reshape(-1)
        host[2 * npx + 16:] = np.array([f.timestamp], np.float64).view(np.float32)
        return t.from_numpy(host)

    def process_frame(self, packed_dev=None, M=None, timestamp=None):
        """packed_dev: on the root, a device tensor in pack_frame layout.  M / timestamp: host copies of the pose every rank
        already knows (the bench's synthetic trajectory); when None they are read back from the broadcast buffer."""
        dist, t = self.dist, self.torch
        npx = self.cfg.W * self.cfg.H
        if self.rank == self.root and packed_dev is not None:
            self.frame_buf.copy_(packed_dev, non_blocking=True)
        if self.world > 1:
            dist.broadcast(self.frame_buf, src=self.root)                      # NCCL over NVLink
        if M is None:
            tail = self.frame_buf[2 * npx:].cpu().numpy()
            M, timestamp = tail[:16], float(tail[16:18].view(np.float64)[0])
        self.m.view.c.timestamp = timestamp
        self.m.trk.set_pose(M)
        local = abi.FRAME_ALLOCATE | abi.FRAME_INTEGRATE | abi.FRAME_EXPECTED_DEPTHS
        if self.mode == "composite":
            self.m.fuse_and_raycast(self.frame_idx, local | abi.FRAME_RAYCAST)
            self.rendezvous()                # every shard's raycast of this frame has completed
            # reduce-scatter shape: shard r composites its row slice from all N results and writes it into the root's image
            # (each GPU reads 16 W H bytes, 1/N of them locally, instead of the root pulling 16 N W H bytes through one link)
            self.m.composite_rows(self.peer_ray[self.frame_idx & 1], row_slice(self.cfg.H, self.rank, self.world), self.root_image)
            self.rendezvous()                # every slice is in the root's image
            if self.rank == self.root:
                self.m.icp_from_composite()
        else:
            self.m.fuse_and_raycast(self.frame_idx, local)
            self.rendezvous()                # every shard's tables and range image are final for this frame
            self.m.raycast_lookup(self.tables, self.root_image)
            self.rendezvous()                # all tiles are in the root's image; nobody reads my tables any more
            if self.rank == self.root:
                self.m.icp_from_composite()
        self.frame_idx += 1

    def rendezvous(self):
        """stream-ordered barrier: an NCCL all-reduce of one int completes on a rank only after every rank has reached it"""
        if self.world > 1:
            self.dist.all_reduce(self.token)

    def close(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        for p in self._opened:
            _lib().dsb200_ipc_close(p)
        self.m.view.c.depth, self.m.view.c.rgb = self.m.view.depth.ptr, self.m.view.rgb.ptr
        self.m.close()
