"""Host <-> device block swapping on the GPU (dsb200_swap_*) against the oracle, which
tests/test_oracle_vs_ref.py pins to ITMSwappingEngine_CPU: volume, hash pointers, free list, swap states and the host store
must match bit for bit after every call."""
import os
import tempfile

import numpy as np
import pytest

import harness as hz
from harness import abi, po

pytestmark = pytest.mark.gpu


def compare(gpu, ora, swap, cache, what):
    hz.assert_state_equal(gpu.state(), ora.state(), what)
    assert np.array_equal(swap.states(), cache.states), what + ": swap states"
    has = swap.has_stored()
    assert np.array_equal(has, cache.has_stored), what + ": stored flags"
    for e in np.nonzero(has)[0][::5]:
        assert np.array_equal(swap.stored_block(e).view(np.uint64), cache.stored_block(e).view(np.uint64)), what + ": stored block %d" % e


SMALL = dict(num_blocks=8192, bucket_num=16384, excess_size=4096)
WIDE = dict(num_blocks=8192, bucket_num=1 << 21, excess_size=0x20000)     # 1089 candidate tiles: the tile scan takes two trips


@pytest.mark.parametrize("transfer,caps", [(0x1000, SMALL), (96, SMALL), (0x1000, WIDE)], ids=["all", "capped", "wide-table"])
def test_swapping_engine_vs_oracle(transfer, caps):
    """the reference's per-frame schedule (allocate with useSwapping, integrate, IntegrateGlobalIntoLocal, SaveToGlobalMemory with
    the render state) over a trajectory that leaves and re-enters parts of the room, a whole-map swap-out in the middle
    (SaveToGlobalMemory without a render state), then fused frames on the swapped volume.  transfer = 96 makes every call hit the
    per-call block limit (the first 96 candidates in entry order move); the small table (16 Ki buckets) also puts a good part of
    the entries on the excess list."""
    cfg, frames, order, whole_map_after = hz.swap_scenario()
    gpu = hz.GpuBackend(cfg, caps=caps, use_swapping=True)
    ora = hz.OracleBackend(cfg, caps=caps, use_swapping=True)
    swap = gpu.eng.ITMSwappingEngine(gpu.scene, transfer, recon=gpu.recon)
    cache = po.HostSwapCache(ora.scene.c.total_entries, transfer)
    moved_in = moved_out = capped = 0
    for step, k in enumerate(order):
        f = frames[k]
        gpu.set_frame(f); ora.set_frame(f)
        assert gpu.allocate() == ora.allocate()
        swap.MarkVisible(gpu.rs); ora.o.swap_mark_visible(cache, ora.rs)
        gpu.integrate(); ora.integrate()
        a, b = swap.IntegrateGlobalIntoLocal(gpu.scene), ora.o.swap_integrate_global_into_local(cache, ora.scene)
        assert a == b
        moved_in += a; capped += int(a == transfer)
        a, b = swap.SaveToGlobalMemory(gpu.scene, gpu.rs), ora.o.swap_save_to_global_memory(cache, ora.scene, ora.rs)
        assert a == b
        moved_out += a; capped += int(a == transfer)
        if step == whole_map_after:
            a, b = swap.SaveToGlobalMemory(gpu.scene), ora.o.swap_save_to_global_memory(cache, ora.scene, None)
            assert a == b and a > 0
        compare(gpu, ora, swap, cache, "transfer %d step %d (frame %d)" % (transfer, step, k))
    assert moved_in > 300 and moved_out > 90 and (cache.states == 0).any() and (cache.states == 2).any()
    assert (capped > 4) if transfer == 96 else (capped == 0)
    # the fused frame (allocation, integration from / into the sdf shadow, raycast) on a volume the swapping engine has rewritten
    for k in (2, 3):
        f = frames[k]
        gpu.set_frame(f); ora.set_frame(f)
        assert gpu.process_frame() == ora.process_frame()
        swap.MarkVisible(gpu.rs); ora.o.swap_mark_visible(cache, ora.rs)
        assert swap.IntegrateGlobalIntoLocal(gpu.scene) == ora.o.swap_integrate_global_into_local(cache, ora.scene)
        compare(gpu, ora, swap, cache, "fused frame %d after swapping" % k)
        hz.assert_images_equal(gpu.images(), ora.images(), "fused frame %d after swapping" % k)
    gpu.set_frame(frames[3]); ora.set_frame(frames[3])
    assert gpu.process_frame() == ora.process_frame()
    hz.assert_images_equal(gpu.images(), ora.images(), "frame after combining the stored blocks")
    compare(gpu, ora, swap, cache, "frame after combining the stored blocks")
    swap.close(); gpu.close(); ora.close()


def test_swap_store_file_round_trip():
    """ITMGlobalCache::SaveToFile / ReadFromFile layout: flags then every block; a fresh cache reads the store back"""
    cfg, frames, order, _ = hz.swap_scenario()
    caps = dict(num_blocks=4096, bucket_num=8192, excess_size=1024)
    gpu = hz.GpuBackend(cfg, caps=caps, use_swapping=True)
    swap = gpu.eng.ITMSwappingEngine(gpu.scene, recon=gpu.recon)
    gpu.set_frame(frames[0]); gpu.allocate(); swap.MarkVisible(gpu.rs); gpu.integrate()
    assert swap.IntegrateGlobalIntoLocal(gpu.scene) > 100
    n = swap.SaveToGlobalMemory(gpu.scene)
    assert n > 100 and (gpu.scene.hash()["ptr"] == -1).sum() == n
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "cache.dat")
        swap.SaveToFile(path)
        total = gpu.scene.c.total_entries
        assert os.path.getsize(path) == total + total * 512 * 8
        other = gpu.eng.ITMSwappingEngine(gpu.scene)
        other.ReadFromFile(path)
        has = swap.has_stored()
        assert has.sum() == n and np.array_equal(other.has_stored(), has)
        fused = 0
        for e in np.nonzero(has)[0][::9]:
            assert np.array_equal(other.stored_block(e).view(np.uint64), swap.stored_block(e).view(np.uint64))
            fused += int((other.stored_block(e)["w_depth"] > 0).any())
        assert fused > 5
        other.close()
        # a truncated file is refused up front and leaves the cache with nothing marked stored
        with open(path, "r+b") as f:
            f.truncate(total + (total // 2) * 512 * 8)
        broken = gpu.eng.ITMSwappingEngine(gpu.scene)
        with pytest.raises(abi.Dsb200Error):
            broken.ReadFromFile(path)
        assert broken.has_stored().sum() == 0
        assert abi.load().dsb200_swap_stored_block(broken.h, int(np.nonzero(has)[0][0])) in (None, 0)
        broken.close()
    swap.close(); gpu.close()


def test_swapping_with_an_asynchronous_engine():
    """SaveToGlobalMemory used the scene's HOST mirror of the free-list top, which is stale while the
    reconstruction engine runs asynchronously.  With the engine attached (dsb200_swap_attach_recon) the swap calls refresh the
    mirrors first and notify the engine afterwards, so an asynchronous pipeline with swap-outs in between -- no dsb200_recon_sync
    by the caller anywhere -- ends in the oracle's state."""
    cfg, frames, order, _ = hz.swap_scenario()
    gpu = hz.GpuBackend(cfg, caps=SMALL, use_swapping=True)
    ora = hz.OracleBackend(cfg, caps=SMALL, use_swapping=True)
    swap = gpu.eng.ITMSwappingEngine(gpu.scene, 0x1000, recon=gpu.recon)
    cache = po.HostSwapCache(ora.scene.c.total_entries, 0x1000)
    gpu.recon.SetAsync(True)
    for step, k in enumerate(order[:8]):
        f = frames[k]
        gpu.set_frame(f); ora.set_frame(f)
        gpu.recon.ProcessFrame(gpu.vis, gpu.scene, gpu.view, gpu.trk, gpu.rs)      # asynchronous: host mirrors NOT updated
        ora.process_frame()
        swap.MarkVisible(gpu.rs); ora.o.swap_mark_visible(cache, ora.rs)
        a, b = swap.IntegrateGlobalIntoLocal(gpu.scene), ora.o.swap_integrate_global_into_local(cache, ora.scene)
        assert a == b, step
        a, b = swap.SaveToGlobalMemory(gpu.scene, gpu.rs), ora.o.swap_save_to_global_memory(cache, ora.scene, ora.rs)
        assert a == b, step
    gpu.recon.Sync(gpu.scene, gpu.rs)
    gpu.recon.SetAsync(False)
    compare(gpu, ora, swap, cache, "asynchronous engine + swapping")
    hz.assert_images_equal(gpu.images(), ora.images(), "asynchronous engine + swapping")
    swap.close(); gpu.close(); ora.close()


def test_host_store_budget_is_enforced_without_damage():
    """ADVICE round 1 (pinned store growth): with a bound on the host store a swap-out moves what still fits, then fails cleanly
    with nothing changed; raising the bound lets it finish, and every stored block is the block that was in the volume."""
    cfg, frames, order, _ = hz.swap_scenario()
    caps = dict(num_blocks=4096, bucket_num=8192, excess_size=1024)
    gpu = hz.GpuBackend(cfg, caps=caps, use_swapping=True)
    swap = gpu.eng.ITMSwappingEngine(gpu.scene, 64, recon=gpu.recon)          # 64 blocks per call
    gpu.set_frame(frames[0]); gpu.allocate(); swap.MarkVisible(gpu.rs); gpu.integrate()
    while swap.IntegrateGlobalIntoLocal(gpu.scene):
        pass
    before_hash, before_vox = gpu.scene.hash().copy(), gpu.scene.voxels().copy()
    live = int((before_hash["ptr"] >= 0).sum())
    assert live > 300
    swap.SetStoreBudget(100)
    assert swap.SaveToGlobalMemory(gpu.scene) == 64
    assert swap.SaveToGlobalMemory(gpu.scene) == 36                            # what still fits: a shorter transfer
    h1, v1, free1 = gpu.scene.hash().copy(), gpu.scene.voxels().copy(), gpu.scene.c.last_free_block
    with pytest.raises(abi.Dsb200Error) as err:
        swap.SaveToGlobalMemory(gpu.scene)
    assert "host block store is full" in str(err.value)
    assert gpu.scene.hash().tobytes() == h1.tobytes() and np.array_equal(gpu.scene.voxels().view(np.uint64), v1.view(np.uint64))
    assert gpu.scene.c.last_free_block == free1 and swap.has_stored().sum() == 100
    swap.SetStoreBudget(0)
    moved = 100
    while True:
        n = swap.SaveToGlobalMemory(gpu.scene)
        if n == 0:
            break
        moved += n
    has = swap.has_stored()
    assert moved == has.sum() and moved >= live - 64                           # entries still in state 0 (never visible) stay
    for e in np.nonzero(has)[0][::7]:
        p = int(before_hash["ptr"][e])
        assert np.array_equal(swap.stored_block(e).view(np.uint64), before_vox[p * 512:(p + 1) * 512].view(np.uint64))
    swap.close(); gpu.close()
