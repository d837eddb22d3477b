#!/usr/bin/env python
"""bench.py -- frames/s of the fused volumetric-fusion step (allocate + integrate + expected depths +
raycast/ICP maps) on the KITTI-stereo-shaped workload of BASELINE.json configs[1]
(1242x375 synthetic disparity depth, 2 cm voxels, depth weighting on).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

ours:      one step = one depth frame through dsb200_process_frame.  `value` = frames/s with the
           frames already resident in HBM (device time, CUDA events, L2 flushed between steps);
           `e2e` = the same through the C ABI with pinned HOST buffers (H2D of depth+rgb and a
           D2H of the rendered map + counters inside the timed region).
reference: the reference's own CPU engines (oracle/_ref, compiled from the unmodified reference
           sources) on the host cores, on a bounded sample of the same workload.
N > 1:     the path shards by independent scenes ("replicas only"): every rank fuses its
           own sequence (seed = rank), no data-path collective; value = total frames / max time.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
PKG = "denseslam-global-consistency-h_b200"

WORKLOAD = "C2"
CAPS = dict(num_blocks=1 << 21, bucket_num=1 << 22, excess_size=1 << 20)   # 8 GiB VBA: 1000 street frames fit
# other BASELINE.json configs, selected with --workload (the driver's default run stays C2, the config the metric is quoted on)
WORKLOADS = {
    "C1": dict(caps=dict(num_blocks=1 << 20, bucket_num=1 << 21, excess_size=1 << 19), prune=None,
               text="C1 ICL-NUIM-shaped 640x480 synthetic depth, 5 mm voxels, mu 2 cm, orbit in a furnished room"),
    "C2": dict(caps=CAPS, prune=None,
               text="C2 KITTI-stereo-shaped 1242x375 synthetic disparity depth, 2 cm voxels, mu 8 cm, depth weighting on"),
    "C3": dict(caps=dict(num_blocks=1 << 20, bucket_num=1 << 21, excess_size=1 << 19), prune=dict(max_weight=3, min_age=40, max_age=80),
               text="C3 TUM-RGBD-shaped 640x480 synthetic depth, 5 mm voxels, voxel decay (max weight 3, min age 40) + "
                    "sliding-window block pruning (max age 80) after every frame"),
}
REF_BLOCK_RESERVE = 24000   # the reference's capacities are compile-time (262144 blocks): its arms stop before the array runs out
REF_MAX_FRAMES = 60         # ... and are bounded to this many timed frames (about 25 s of one host core at C2)
METRIC = "frames/s alloc+integrate+raycast (1242x375, 2 cm voxels)"
DIGESTS = os.path.join(ROOT, "tests", "golden", "bench_digests.json")


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, rank=0, world=1):
        # N > 1: rank 0 alone samples, all N GPUs with one nvidia-smi process (N concurrent NVML pollers perturb submission)
        self.rows, self.proc = [], None
        self.gpu = ",".join(str(i) for i in range(world)) if world > 1 else str(gpu_index)
        self.enabled = rank == 0

    def start(self):
        if not self.enabled:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", self.gpu, "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.first = threading.Event()
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            # nvidia-smi's start-up (NVML initialisation, tens of ms) holds driver locks that stall kernel submission: a 40 ms step was
            # measured 20 ms into a timed region (profiles/r02v).  Wait for its first row so that only the 100 ms polls, which do not
            # show up in the step times, fall into the timed region.
            self.first.wait(3.0)
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])
            self.first.set()

    def stop(self):
        if not self.enabled:
            return None
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 9 for i in range(4) if r[5 + i].lower().startswith("active")})
        out = {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
               "samples": len(sm)}
        gpus = sorted({r[0] for r in self.rows if len(r) >= 9})
        if len(gpus) > 1:      # one record per GPU of the job
            out["per_gpu"] = {g: {"sm_mhz": float(np.median([float(r[1]) for r in self.rows if len(r) >= 9 and r[0] == g and r[1].replace(".", "").isdigit()] or [0])),
                                  "samples": sum(1 for r in self.rows if len(r) >= 9 and r[0] == g)} for g in gpus}
        return out


def ncu_traffic(stage):
    """DRAM bytes per launch (read + write) of the stage's dominant kernel from the committed `ncu --set full` extract
    profiles/ncu_traffic.json = {"captured_at_commit": ..., "command": ..., "kernels": {name: bytes per launch}} (see
    profiles/README.md).  A capture is a property of the commit it was taken at: the stamp travels with the number."""
    kernel = {"integrate": "k_fuse_warp_flat", "raycast": "k_raycast", "allocate+visible_list": "k_alloc_requests"}.get(stage)
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if WORKLOAD != t.get("workload", "C2") or kernel not in t["kernels"]:
            return None, None
        return float(t["kernels"][kernel]), "profiles/ncu_traffic.json captured at commit %s" % t.get("captured_at_commit")
    except Exception:
        return None, None


def make_frames(n, seed):
    synth = importlib.import_module(PKG + ".synth")
    return synth.make_sequence(WORKLOAD, n, seed=seed)


# ------------------------------------------------------------------------------------------------ reference arm
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_reference_cpu(frames_all, cfg, first, count, lib_kind):
    """The reference's own CPU engines (oracle/_ref, unmodified sources) over frames [0, first) untimed and
    [first, first + count) timed, from a reset scene -- the frames our arm times.  lib_kind: 'serial' (reference default build,
    1 core) or 'omp' (its WITH_OPENMP option, all host threads; allocation requests race there, timing only).
    Stops early when the compile-time voxel block array (262144 blocks) is about to run out.  -> (frames/s, frames timed)"""
    import pyoracle as po
    path = po.REF_OMP_PATH if lib_kind == "omp" else po.REF_FAST_PATH
    eng = po.RefEngine(cfg.W, cfg.H, cfg.params(), path=path)
    eng.set_calib(cfg.intr)

    def step(f):
        eng.set_view(f.depth, f.rgb, f.timestamp); eng.set_pose(f.M)
        eng.allocate(); eng.integrate(); eng.create_expected_depths(); eng.create_icp_maps()
    for f in frames_all[:first]:
        step(f)
    total, done = 0.0, 0
    for f in frames_all[first:first + count]:
        if eng.L.itmref_last_free_block(eng.h) < REF_BLOCK_RESERVE:
            break
        t0 = time.perf_counter()
        step(f)
        total += time.perf_counter() - t0
        done += 1
    eng.close()
    return (done / total if total > 0 else 0.0), done


def run_reference(args, rank, world, frames=None, cfg=None):
    if rank != 0:
        return None
    import pyoracle as po
    abi = importlib.import_module(PKG + ".abi")
    K = min(args.steps, REF_MAX_FRAMES)
    if frames is None:
        cfg, frames = make_frames(args.warmup + K, 0)
    threads = host_threads()
    extra = {}
    if os.path.exists(po.REF_FAST_PATH):
        kind = "reference"
        fps1, done = time_reference_cpu(frames, cfg, args.warmup, K, "serial")
        extra["serial_1_core"] = {"value": fps1, "unit": "frames/s", "frames": done}
        fps, cores = fps1, 1
        if os.path.exists(po.REF_OMP_PATH) and threads > 1:
            os.environ.setdefault("OMP_NUM_THREADS", str(threads))
            fpsN, doneN = time_reference_cpu(frames, cfg, args.warmup, K, "omp")
            extra["openmp_all_threads"] = {"value": fpsN, "unit": "frames/s", "frames": doneN, "threads": threads}
            if fpsN > fps:
                fps, cores, done = fpsN, threads, doneN
        sample = ("frames %d..%d of the %s sequence from a reset scene (the frames the GPU arm times%s), frames 0..%d fused untimed; "
                  "reference CPU engines compiled from the unmodified sources, -O3 -march=x86-64-v3, serial (the reference's default "
                  "build) and with its WITH_OPENMP option on %d host threads (nproc = %d); the better of the two is the value; the "
                  "reference CPU integrator ignores the depth-weighting parameters" % (
                      args.warmup, args.warmup + done - 1, WORKLOAD, "" if done == args.steps else ", bounded to %d: the reference's "
                      "voxel block array is 262144 blocks at compile time" % done, args.warmup - 1, threads, os.cpu_count() or threads))
    else:   # the oracle port is the only CPU implementation available on this box
        import harness as hz
        kind, cores = "port", 1
        ora = hz.OracleBackend(cfg, dict(num_blocks=abi.DEFAULT_LOCAL_BLOCKS, bucket_num=abi.DEFAULT_BUCKETS, excess_size=abi.DEFAULT_EXCESS))
        ora.set_weights(cfg.depth_weighting, 10, 100)
        for f in frames[:args.warmup]:
            ora.set_frame(f); ora.process_frame()
        t0 = time.perf_counter()
        done = 0
        for f in frames[args.warmup:args.warmup + K]:
            ora.set_frame(f); ora.process_frame(); done += 1
        fps = done / (time.perf_counter() - t0)
        sample = "frames %d..%d of the %s sequence, oracle/restate.cpp port on one core" % (args.warmup, args.warmup + done - 1, WORKLOAD)
    return {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": done,
            "warmup": args.warmup, "ms_per_step": 1000.0 / fps if fps > 0 else None, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 math -> i16 sdf / u8 weights", "data": "synthetic",
            "config": {"workload": workload_text()},
            "cpu_baseline": dict({"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample}, **extra),
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def workload_text():
    return WORKLOADS[WORKLOAD]["text"] + "; one scene per GPU (replicas only; every GPU runs the same synthetic sequence)"


def time_reference_cuda(frames, cfg, d_depth, d_rgb, first, count, flush):
    """SURVEY 8(d) GPU comparator: the reference's own CUDA engines (unmodified sources recompiled for sm_100 with the
    reference's --use_fast_math, oracle/_ref/libitmref_cuda_fast.so) on the same device-resident frames, same L2 flush between
    frames.  Their calls are synchronous (legacy stream, blocking 12-byte copies), so a frame is timed both as the sum of its four
    CUDA-event stage times and by the host clock around the four calls."""
    import pyoracle as po
    import torch
    if not os.path.exists(po.REF_CUDA_FAST_PATH):
        return {"unavailable": "oracle/_ref/libitmref_cuda_fast.so not built (needs /root/reference at build time)"}
    eng = po.RefCudaEngine(cfg.W, cfg.H, cfg.params(), use_swapping=False, fast=True)
    eng.set_calib(cfg.intr)
    eng.set_weights(cfg.depth_weighting, 10, 100)
    stage = np.zeros(5)
    done = 0
    for i in range(first + count):
        if eng.last_free()[0] < REF_BLOCK_RESERVE:
            break
        eng.set_view_device(d_depth[i].data_ptr(), d_rgb[i].data_ptr(), frames[i].timestamp)
        eng.set_pose(frames[i].M)
        if i >= first:
            flush.zero_()
            torch.cuda.synchronize()
        ms = eng.process_frame_timed()
        if i >= first:
            stage += ms
            done += 1
    visible = eng.L.itmcu_no_visible(eng.h, 0)
    eng.close()
    if not done:
        return {"unavailable": "no frame fitted the reference's compile-time capacities"}
    dev = float(stage[:4].sum())
    return {"value": done / (dev * 1e-3), "unit": "frames/s", "frames": done, "ms_per_frame_device": dev / done,
            "ms_per_frame_host_clock": float(stage[4]) / done, "value_host_clock": done / (float(stage[4]) * 1e-3),
            "stages_ms": dict(zip(["allocate+visible_list", "integrate", "expected_depths", "raycast+icp_maps"], [float(x) / done for x in stage[:4]])),
            "final_visible_blocks": int(visible),
            "what": "reference CUDA engines (ITMSceneReconstructionEngine_CUDA.cu, ITMVisualisationEngine_CUDA.cu, unmodified) recompiled "
                    "-gencode arch=compute_100,code=sm_100 --use_fast_math; frames %d..%d resident in HBM, L2 flushed between frames; device "
                    "time = CUDA events around the four engine calls on the legacy stream" % (first, first + done - 1)}


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args, rank, world, local_rank):
    """the driver's line: the C2 workload (or --workload), and at N = 1 the other single-GPU BASELINE configs (C1, C3) as
    `workloads` sub-objects measured the same way (own clocks, own parity check), so the 640x480 half of the metric is in the
    same record"""
    import torch
    import torch.distributed as dist
    global WORKLOAD, CAPS, METRIC
    abi = importlib.import_module(PKG + ".abi")
    lib = abi.load()          # raises if libdsb200.so is missing: no fallback
    if lib.dsb200_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device; libdsb200 has no CPU fallback")
    abi.check(lib.dsb200_set_device(local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    out = measure_workload(args, rank, world, local_rank, full=True)
    if not args.no_sharded:
        # BASELINE configs[3] next to the replicas line: ONE ~4 M-block scene sharded over the N GPUs (N = 1: the whole map on
        # one GPU, the denominator of the driver's scaling curve), fast mode and exact mode
        init_group(world, dev)
        sh = {}
        for mode in ("composite", "lookup"):
            r = measure_sharded(args, rank, world, local_rank, mode)
            if r is not None:
                sh[mode] = r
        if out is not None:
            out["sharded"] = sh
    if world == 1 and out is not None and not args.single_workload:
        main_wl, main_caps, main_metric = WORKLOAD, CAPS, METRIC
        out["workloads"] = {}
        for wl in sorted(WORKLOADS.keys()):
            if wl == main_wl:
                continue
            WORKLOAD, CAPS, METRIC = wl, WORKLOADS[wl]["caps"], metric_text(wl)
            sub = measure_workload(args, rank, world, local_rank, full=False)
            out["workloads"][wl] = {k: sub[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "warmup", "e2e", "clocks", "roofline",
                                                        "stages", "parity_check", "gpu_launches", "config", "prune_ms_per_frame",
                                                        "decayed_blocks") if k in sub}
        WORKLOAD, CAPS, METRIC = main_wl, main_caps, main_metric
    if dist.is_initialized():
        dist.destroy_process_group()
    return out


def metric_text(wl):
    if wl == "C2":
        return "frames/s alloc+integrate+raycast (1242x375, 2 cm voxels)"
    return "frames/s alloc+integrate+raycast (640x480, 5 mm voxels%s)" % (", decay + slide window" if wl == "C3" else "")


def state_digest_of(digest, eng_mod, scene, rs, trk, recon):
    """digest of the map the timed frames left behind (untimed; downloads the allocated part of the voxel array)"""
    h = scene.hash()
    nb = scene.c.num_blocks

    def fetch(lo, hi):
        return eng_mod._download(scene.c.voxels + lo * 512 * 8, eng_mod.abi.VOXEL_DTYPE, (hi - lo) * 512)
    images = dict(range=rs.range_image(), raycast=rs.raycast_result(), points=trk.icp_points(), normals=trk.icp_normals(),
                  rendering=rs.raycast_image())
    d = digest.state_digest(h, (scene.c.last_free_block, scene.c.last_free_excess), rs.visible(), fetch, nb, images)
    d["no_visible"] = int(rs.c.no_visible)
    return d


def check_digest(got, n_frames):
    key = "%s:%d" % (WORKLOAD, n_frames)
    try:
        table = json.load(open(DIGESTS))
    except Exception:
        return "no golden digests (tests/golden/bench_digests.json missing)", key
    if key not in table:
        return "no golden digest for %s (tests/golden/make_bench_digests.py %s)" % (key, key), key
    digest = importlib.import_module(PKG + ".digest")
    return digest.compare(got, table[key]), key


def measure_workload(args, rank, world, local_rank, full):
    import torch
    import torch.distributed as dist
    abi = importlib.import_module(PKG + ".abi")
    eng = importlib.import_module(PKG + ".engine")
    digest = importlib.import_module(PKG + ".digest")
    lib = abi.load()
    dev = torch.device("cuda", local_rank)
    K, Wm = args.steps, args.warmup
    cfg, frames = make_frames(K + Wm + 1, seed=0)  # replicas: every GPU runs the same sequence (weak scaling, equal work per GPU);
    # one extra frame: the look-ahead issued during the last timed frame has a real successor, so every timed frame carries
    # exactly one request pass (its successor's) and none is left outside the timed region
    W, H = cfg.W, cfg.H
    npx = W * H

    scene = eng.ITMScene(cfg.params(), CAPS["num_blocks"], CAPS["bucket_num"], CAPS["excess_size"])
    recon = eng.ITMSceneReconstructionEngine(CAPS["num_blocks"], CAPS["bucket_num"], CAPS["excess_size"])
    recon.SetFusionWeightParams(cfg.depth_weighting, 10, 100)
    vis = eng.ITMVisualisationEngine(W, H)
    rs = vis.CreateRenderState(scene, (W, H))
    trk = eng.ITMTrackingState(W, H)
    view = eng.ITMView(W, H, cfg.intr)          # staging view for the e2e leg
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    # ---- leg 1: inputs resident in HBM
    d_depth = torch.from_numpy(np.stack([f.depth for f in frames])).to(dev)
    d_rgb = torch.from_numpy(np.stack([f.rgb for f in frames])).to(dev)
    views = []
    for i, f in enumerate(frames):
        v = abi.View()
        C.memmove(C.byref(v), C.byref(view.c), C.sizeof(abi.View))
        v.depth, v.rgb, v.timestamp = d_depth[i].data_ptr(), d_rgb[i].data_ptr(), f.timestamp
        views.append(v)

    prune = WORKLOADS[WORKLOAD]["prune"]
    prune_ev = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]
    prune_ms = [0.0]

    def prune_pass(timed=False):
        """DenseSlam::ProcessFrame's map maintenance (DS/DenseSlam.cpp:217-251): SlideWindow, then Decay"""
        if not prune:
            return
        if timed:
            prune_ev[0].record()
        abi.check(lib.dsb200_slide_window(recon.h, C.byref(scene.c), C.byref(rs.c), prune["max_age"]))
        abi.check(lib.dsb200_decay(recon.h, C.byref(scene.c), C.byref(rs.c), prune["max_weight"], prune["min_age"], 0))
        if timed:
            prune_ev[1].record()
            prune_ev[1].synchronize()
            prune_ms[0] += prune_ev[0].elapsed_time(prune_ev[1])

    lookahead = not args.no_lookahead

    def frame_async(i, timed_prune=False, profile=False):
        trk.set_pose(frames[i].M)
        abi.check(lib.dsb200_process_frame(recon.h, vis.h, C.byref(scene.c), C.byref(views[i]), C.byref(trk.c), C.byref(rs.c), abi.FRAME_ALL))
        prune_pass(timed_prune)
        if lookahead and not profile and i + 1 < len(views):
            # streaming look-ahead: the request pass of frame i+1 overlaps the raycast of frame i
            trk.set_pose(frames[i + 1].M)
            abi.check(lib.dsb200_recon_prefetch_allocation(recon.h, C.byref(scene.c), C.byref(views[i + 1]), C.byref(trk.c), C.byref(rs.c), None))
            abi.check(lib.dsb200_recon_join_prefetch(recon.h))     # the frame's closing event (and the L2 flush) come after the look-ahead

    def reset_all():
        recon.ResetScene(scene)
        rs.c.no_visible = 0
        abi.check(lib.dsb200_memcpy_h2d(rs.c.visible_type, zeros_vt.ctypes.data, scene.c.total_entries, None))
        abi.check(lib.dsb200_stream_synchronize(None))

    zeros_vt = np.zeros(scene.c.total_entries, np.uint8)
    alloc_stage = np.zeros(5)
    stage_buf = (C.c_float * abi.PROFILE_STAGES)()

    def resident_pass(profile):
        """W warm-up + K timed frames from a fresh scene.  profile=False: the product configuration (expected depths
        overlapped with integration); profile=True: stages serialised and timed one by one with CUDA events."""
        reset_all()
        recon.SetAsync(True)
        abi.check(lib.dsb200_profile_enable(recon.h, 1 if profile else 0))
        stage = np.zeros(abi.PROFILE_STAGES)
        counts = []
        for i in range(Wm):
            frame_async(i, False, profile)
        recon.Sync(scene, rs)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        l0 = lib.dsb200_kernel_launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        torch.cuda.synchronize()
        for k in range(K):
            flush.zero_()                      # L2 flush between timed iterations (outside the event pair)
            evs[k][0].record()
            frame_async(Wm + k, profile, profile)
            evs[k][1].record()
            if profile:
                abi.check(lib.dsb200_profile_read(recon.h, stage_buf))      # synchronises; device-side stage times
                stage += np.array(stage_buf[:])
                abi.check(lib.dsb200_profile_read_allocate(recon.h, stage_buf))
                np.add(alloc_stage, np.array(stage_buf[:])[:5], out=alloc_stage)
            recon.Sync(scene, rs)
            counts.append(rs.c.no_visible)
        torch.cuda.synchronize()
        n_launch = int(lib.dsb200_kernel_launch_count() - l0)
        if world > 1:
            dist.barrier()
        ms = float(sum(a.elapsed_time(b) for a, b in evs))
        abi.check(lib.dsb200_profile_enable(recon.h, 0))
        recon.SetAsync(False)
        return ms, stage, counts, n_launch

    sampler = ClockSampler(local_rank, rank, world)
    sampler.start()
    dev_ms, _, vis_counts, launches = resident_pass(False)
    clocks = sampler.stop()
    blocks_used = CAPS["num_blocks"] - 1 - scene.c.last_free_block
    # what did the timed frames build?  (untimed) digest of hash + free lists + visible list + voxels + the last frame's images,
    # compared with the digest the CPU checker produced for the same frames (tests/golden/bench_digests.json)
    parity = {}
    if rank == 0:
        abi.check(lib.dsb200_recon_abandon_prefetch(recon.h))     # a pending (deep) look-ahead holds frame k+1's entries: undone exactly
        abi.check(lib.dsb200_stream_synchronize(None))
        verdict, key = check_digest(state_digest_of(digest, eng, scene, rs, trk, recon), Wm + K)
        parity = {"resident_pass": verdict, "golden": key}
    _, stage_ms, _, _ = resident_pass(True)

    # ---- leg 2: end to end through the C ABI with pinned host buffers
    # the caller's frames live in pinned host memory (as a camera driver / decoder would deliver them)
    # ... in the camera's own formats: int16 millimetres and packed 8-bit BGR (what DenseSLAM's driver receives as cv::Mat1s /
    # cv::Mat3b, DS/InfiniTamDriver.cpp:280-287); the view is built on the device (dsb200_view_from_raw)
    h_depth = torch.from_numpy(np.stack([f.depth_mm for f in frames]).astype(np.int16)).pin_memory()
    h_rgb = torch.from_numpy(np.ascontiguousarray(np.stack([f.rgb[..., 2::-1] for f in frames]))).pin_memory()
    h_out = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
    raw_d = [eng.DeviceBuffer(npx * 2), eng.DeviceBuffer(npx * 2)]
    raw_c = [eng.DeviceBuffer(npx * 3), eng.DeviceBuffer(npx * 3)]
    mm_to_m = np.float32(1.0 / 1000.0)
    reset_all()

    # Streaming pipeline, as a camera-driven caller would run it: frame k+1 is uploaded (copy stream, second view buffer)
    # while frame k is fused, and frame k's rendering is read back (read-back stream, second image buffer) while frame k+1
    # is fused.  Every byte still crosses PCIe inside the timed region; the host blocks once, after the last read-back.
    view2 = eng.ITMView(W, H, cfg.intr)
    img2 = eng.DeviceBuffer(npx * 4)
    views_e = [view, view2]
    imgs_e = [rs.c.raycast_image, img2.ptr]
    h_outs = [h_out, torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()]
    main = torch.cuda.current_stream()
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()
    ev_ready = [torch.cuda.Event(), torch.cuda.Event()]      # upload of buffer b finished
    ev_free = [torch.cuda.Event(), torch.cuda.Event()]       # fusion that read buffer b finished
    ev_done = [torch.cuda.Event(), torch.cuda.Event()]       # frame that wrote image b finished
    ev_read = [torch.cuda.Event(), torch.cuda.Event()]       # read-back of image b finished

    def upload(i):
        b = i & 1
        s_up.wait_event(ev_free[b])
        abi.check(lib.dsb200_memcpy_h2d(raw_d[b].ptr, h_depth[i].data_ptr(), npx * 2, s_up.cuda_stream))
        abi.check(lib.dsb200_memcpy_h2d(raw_c[b].ptr, h_rgb[i].data_ptr(), npx * 3, s_up.cuda_stream))
        abi.check(lib.dsb200_view_from_raw(raw_d[b].ptr, raw_c[b].ptr, mm_to_m, 0.0, C.byref(views_e[b].c), s_up.cuda_stream))
        ev_ready[b].record(s_up)

    host_s = {"upload": 0.0, "process_frame": 0.0, "prune": 0.0, "lookahead": 0.0, "readback": 0.0}   # host time spent enqueueing

    def frame_e2e(i, last):
        b = i & 1
        t_a = time.perf_counter()
        if i + 1 < last:
            upload(i + 1)                                    # prefetch the next frame
        t_b = time.perf_counter(); host_s["upload"] += t_b - t_a
        main.wait_event(ev_ready[b])
        main.wait_event(ev_read[b])                          # image buffer b was read back two frames ago
        views_e[b].c.timestamp = frames[i].timestamp
        trk.set_pose(frames[i].M)
        rs.c.raycast_image = imgs_e[b]
        abi.check(lib.dsb200_process_frame(recon.h, vis.h, C.byref(scene.c), C.byref(views_e[b].c), C.byref(trk.c), C.byref(rs.c), abi.FRAME_ALL))
        t_c = time.perf_counter(); host_s["process_frame"] += t_c - t_b
        prune_pass()
        t_d = time.perf_counter(); host_s["prune"] += t_d - t_c
        if lookahead and i + 1 < last:
            nb = (i + 1) & 1                                 # request pass of frame i+1 as soon as its upload has landed
            trk.set_pose(frames[i + 1].M)
            abi.check(lib.dsb200_recon_prefetch_allocation(recon.h, C.byref(scene.c), C.byref(views_e[nb].c), C.byref(trk.c), C.byref(rs.c),
                                                           ev_ready[nb].cuda_event))
        t_e = time.perf_counter(); host_s["lookahead"] += t_e - t_d
        ev_free[b].record(main)
        ev_done[b].record(main)
        s_down.wait_event(ev_done[b])
        abi.check(lib.dsb200_memcpy_d2h(h_outs[b].data_ptr(), imgs_e[b], npx * 4, s_down.cuda_stream))
        ev_read[b].record(s_down)
        host_s["readback"] += time.perf_counter() - t_e

    recon.SetAsync(True)
    for b in range(2):
        ev_free[b].record(main); ev_read[b].record(main)
    upload(0)
    for i in range(Wm):
        frame_e2e(i, Wm + K + 1)
    recon.Sync(scene, rs)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for key in host_s:
        host_s[key] = 0.0
    t0 = time.perf_counter()
    e0.record()
    for k in range(K):
        frame_e2e(Wm + k, Wm + K + 1)       # the last timed frame still uploads and looks ahead to a successor
    abi.check(lib.dsb200_recon_join_prefetch(recon.h))
    main.wait_stream(s_down)
    e1.record()
    torch.cuda.synchronize()
    e2e_wall = time.perf_counter() - t0
    recon.Sync(scene, rs)
    abi.check(lib.dsb200_recon_abandon_prefetch(recon.h))         # the successor's look-ahead (outside the timed region: undone)
    recon.SetAsync(False)
    if rank == 0:
        rs.c.raycast_image = imgs_e[(Wm + K - 1) & 1]      # the image buffer the last frame rendered into
        parity["e2e_pass"] = check_digest(state_digest_of(digest, eng, scene, rs, trk, recon), Wm + K)[0]
    rs.c.raycast_image = imgs_e[0]
    e2e_ms = max(float(e0.elapsed_time(e1)), 1000.0 * e2e_wall)
    e2e_detail = {"device_ms_per_step": float(e0.elapsed_time(e1)) / K, "wall_ms_per_step": 1000.0 * e2e_wall / K,
                  "host_enqueue_ms_per_step": {key: 1000.0 * v / K for key, v in host_s.items()}}
    if world > 1:
        dist.barrier()
        t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms = float(t[0]), float(t[1])
    def release():
        recon.close(); vis.close()
        for o in (scene, rs, view, view2, trk, img2) + tuple(raw_d) + tuple(raw_c):
            o.free()

    if rank != 0:
        release()
        return None

    pk, pk_kind = peaks()
    V = float(np.mean(vis_counts))
    A = float(blocks_used) / float(Wm + K)          # new blocks per frame (mean over the run)
    N_tot = float(CAPS["bucket_num"] + CAPS["excess_size"])
    stage_names = ["allocate+visible_list", "integrate", "expected_depths", "raycast", "icp_maps"]
    per_stage = stage_ms / K
    # algorithmic bytes per frame (SURVEY.md section 8d / DESIGN.md), each datum counted once
    # B_alloc = 4 W H (depth) + 3 N_tot (allocation + visibility flags; the kernels scan 1-bit summaries of them, so real
    # traffic is lower) + 20 (V + A) (hash reads) + 36 A (block coordinates + hash writes) + 4 (V + V_prev) (visible lists)
    alg = [4.0 * npx + 3.0 * N_tot + 20.0 * (V + A) + 36.0 * A + 8.0 * V,
           8216.0 * V + 8.0 * npx, 24.0 * V + 16.0 * (W // 8 + 1) * (H // 8 + 1) + 8.0 * npx,
           4116.0 * V + 16.0 * npx, 52.0 * npx]
    dom = int(np.argmax(per_stage))
    achieved = alg[dom] / (per_stage[dom] * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(stage_names[dom])
    stages = {stage_names[i]: {"ms": float(per_stage[i]), "alg_MB": alg[i] / 1e6,
                               "GBps": alg[i] / (per_stage[i] * 1e-3) / 1e9 if per_stage[i] > 0 else None} for i in range(5)}
    out = {"metric": METRIC, "value": world * K / (dev_ms * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
           "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32 math -> i16 sdf / u8 weights", "data": "synthetic",
           "config": {"workload": workload_text(),
                      "timed_region": "per frame: CUDA-event pair around the frame's launches (process_frame, pruning, the look-ahead); the L2 "
                                      "flush and the host's counter read-back between frames are outside the pairs; e2e is one closed wall-clock region",
                      "lookahead": ("allocation of frame k+1 (request pass, ordered allocation as pending entries, visible list) submitted with frame k and joined before frame k's closing event (results bit-identical)"
                                    if lookahead else "off"), "frames": K, "mean_visible_blocks": V,
                      "blocks_allocated": int(blocks_used), "capacity_blocks": CAPS["num_blocks"], "l2": "flushed (256 MiB write) between timed steps"},
           "e2e": {"value": world * K / (e2e_ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": 5 * npx,
                   "d2h_bytes_per_step": 4 * npx, "ms_per_step": e2e_ms / K, "detail": e2e_detail,
                   "input": "int16 mm depth + packed BGR from pinned host memory, view built on the device (dsb200_view_from_raw)",
                   "pipeline": "double-buffered: upload of frame k+1 and read-back of frame k-1 overlap the fusion of frame k; one host sync at the end"
                               + ("; allocation of frame k+1 (request pass, ordered allocation with pending entries, visible list) overlaps the raycast of frame k (dsb200_recon_prefetch_allocation)" if lookahead else "")},
           "gpu_launches": launches, "clocks": clocks,
           "roofline": {"bound": "hbm", "kernel": stage_names[dom], "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                        "frac": achieved / pk["hbm_gbs"], "traffic": traffic, "traffic_captured_at": traffic_src,
                        "algorithmic_bytes_per_launch": alg[dom], "peak_source": pk_kind + " (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)"},
           "stages": stages, "prune_ms_per_frame": (prune_ms[0] / K) if prune else None,
           "decayed_blocks": recon.GetDecayedBlockCount() if prune else None,
           "allocate_kernels_ms": dict(zip(["begin_frame", "alloc_requests", "allocate", "recheck_visibility", "visible_list"],
                                           [float(x) / K for x in alloc_stage]))}
    out["parity_check"] = parity
    # integrate + raycast together against the HBM roofline (the north star's 60 % target is on this pair)
    ir_ms = per_stage[1] + per_stage[3]
    out["roofline"]["integrate_plus_raycast"] = {"algorithmic_bytes": alg[1] + alg[3], "ms": float(ir_ms),
                                                 "achieved": (alg[1] + alg[3]) / (ir_ms * 1e-3) / 1e9,
                                                 "frac": (alg[1] + alg[3]) / (ir_ms * 1e-3) / 1e9 / pk["hbm_gbs"]}
    if world == 1 and full and not args.no_baselines:
        out["gpu_comparator"] = time_reference_cuda(frames, cfg, d_depth, d_rgb, Wm, min(K, REF_MAX_FRAMES), flush)
        if "value" in out["gpu_comparator"]:
            out["gpu_comparator"]["ours_over_reference_cuda"] = out["value"] / out["gpu_comparator"]["value"]
        a2 = argparse.Namespace(gpus=1, steps=min(K, 20), warmup=Wm)
        ref = run_reference(a2, 0, 1, frames=frames, cfg=cfg)
        out["cpu_baseline"] = ref["cpu_baseline"]
    release()
    del d_depth, d_rgb, flush, h_depth, h_rgb
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------ batched scenes
def run_batched(args, rank, world, local_rank):
    """BASELINE configs[4]: S independent sequences per GPU, one scene + one engine pair + one CUDA stream each, frames
    submitted round-robin so the scenes' (latency-bound) kernels overlap.  value = total frames / max time over ranks."""
    import torch
    import torch.distributed as dist
    abi = importlib.import_module(PKG + ".abi")
    eng = importlib.import_module(PKG + ".engine")
    lib = abi.load()
    if lib.dsb200_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device; libdsb200 has no CPU fallback")
    abi.check(lib.dsb200_set_device(local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    K, Wm, S = args.steps, args.warmup, args.scenes
    caps = dict(num_blocks=abi.DEFAULT_LOCAL_BLOCKS, bucket_num=abi.DEFAULT_BUCKETS, excess_size=abi.DEFAULT_EXCESS)   # the reference's capacities
    lanes = []
    for si in range(S):
        cfg, frames = make_frames(K + Wm, seed=si)
        st = torch.cuda.Stream()
        scene = eng.ITMScene(cfg.params(), **caps)
        recon = eng.ITMSceneReconstructionEngine(stream=st.cuda_stream, **caps)
        recon.SetFusionWeightParams(cfg.depth_weighting, 10, 100)
        vis = eng.ITMVisualisationEngine(cfg.W, cfg.H, stream=st.cuda_stream)
        rs = vis.CreateRenderState(scene, (cfg.W, cfg.H))
        trk = eng.ITMTrackingState(cfg.W, cfg.H)
        view = eng.ITMView(cfg.W, cfg.H, cfg.intr)
        d_depth = torch.from_numpy(np.stack([f.depth for f in frames])).to(dev)
        d_rgb = torch.from_numpy(np.stack([f.rgb for f in frames])).to(dev)
        recon.ResetScene(scene)
        recon.SetAsync(True)
        lanes.append(dict(cfg=cfg, frames=frames, st=st, scene=scene, recon=recon, vis=vis, rs=rs, trk=trk, view=view, depth=d_depth, rgb=d_rgb))
    torch.cuda.synchronize()

    def submit(i):
        for L in lanes:
            L["view"].c.depth, L["view"].c.rgb = L["depth"][i].data_ptr(), L["rgb"][i].data_ptr()
            L["view"].c.timestamp = L["frames"][i].timestamp
            L["trk"].set_pose(L["frames"][i].M)
            abi.check(lib.dsb200_process_frame(L["recon"].h, L["vis"].h, C.byref(L["scene"].c), C.byref(L["view"].c), C.byref(L["trk"].c),
                                               C.byref(L["rs"].c), abi.FRAME_ALL))
    for i in range(Wm):
        submit(i)
    for L in lanes:
        L["recon"].Sync(L["scene"], L["rs"])
    sampler = ClockSampler(local_rank, rank, world); sampler.start()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = lib.dsb200_kernel_launch_count()
    main = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for L in lanes:
        L["st"].wait_event(e0)
    for k in range(K):
        submit(Wm + k)
    for L in lanes:
        main.wait_stream(L["st"])
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    launches = int(lib.dsb200_kernel_launch_count() - l0)
    ms = float(e0.elapsed_time(e1))
    vis_mean = []
    for L in lanes:
        L["recon"].Sync(L["scene"], L["rs"])
        vis_mean.append(L["rs"].c.no_visible)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        dist.destroy_process_group()
    if rank != 0:
        return None
    return {"metric": METRIC, "value": world * S * K / (ms * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 math -> i16 sdf / u8 weights", "data": "synthetic", "mode": "batched",
            "config": {"workload": WORKLOADS[WORKLOAD]["text"] + "; %d independent sequences per GPU, one scene + one stream each "
                                   "(a step = one frame of every sequence)" % S, "scenes_per_gpu": S, "frames_per_scene": K,
                       "final_visible_blocks_mean": float(np.mean(vis_mean)), "l2": "not flushed: %d scenes x visible set exceed L2" % S},
            "gpu_launches": launches, "clocks": clocks}


# ------------------------------------------------------------------------------------------------ tracking in the loop
def run_tracked(args, rank, world, local_rank):
    """the whole InfiniTAM frame (ITMMainEngine::ProcessFrame) with NO external poses -- ICP depth
    tracking of frame k against the raycast of frame k-1 (dsb200_tracker_track_camera), then allocate + integrate +
    expected depths + raycast + ICP maps at the tracked pose.  Not the headline metric (DenseSLAM takes its poses from
    ORB-SLAM2); reported as frames/s of the closed loop plus the drift against the synthetic ground truth."""
    import torch
    import torch.distributed as dist
    abi = importlib.import_module(PKG + ".abi")
    eng = importlib.import_module(PKG + ".engine")
    lib = abi.load()
    if lib.dsb200_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device; libdsb200 has no CPU fallback")
    abi.check(lib.dsb200_set_device(local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    K, Wm = args.steps, args.warmup
    cfg, frames = make_frames(K + Wm, seed=0)
    W, H = cfg.W, cfg.H
    scene = eng.ITMScene(cfg.params(), CAPS["num_blocks"], CAPS["bucket_num"], CAPS["excess_size"])
    recon = eng.ITMSceneReconstructionEngine(CAPS["num_blocks"], CAPS["bucket_num"], CAPS["excess_size"])
    recon.SetFusionWeightParams(cfg.depth_weighting, 10, 100)
    vis = eng.ITMVisualisationEngine(W, H)
    rs = vis.CreateRenderState(scene, (W, H))
    trk = eng.ITMTrackingState(W, H)
    view = eng.ITMView(W, H, cfg.intr)
    tracker = eng.ITMDepthTracker((W, H))          # ITMLibSettings defaults: 5 levels, b b r r r, 0.1^2, 1e-3
    d_depth = torch.from_numpy(np.stack([f.depth for f in frames])).to(dev)
    d_rgb = torch.from_numpy(np.stack([f.rgb for f in frames])).to(dev)
    recon.ResetScene(scene)
    recon.SetAsync(True)
    err, evals, track_ms = [], [], []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def frame(i, timed):
        view.c.depth, view.c.rgb, view.c.timestamp = d_depth[i].data_ptr(), d_rgb[i].data_ptr(), frames[i].timestamp
        if i == 0:
            trk.set_pose(frames[0].M)               # the first pose anchors the map, like ITMMainEngine's identity start
        else:
            if timed:
                t0.record()
            tracker.TrackCamera(trk, view)          # starts from the previous frame's estimate
            if timed:
                t1.record(); t1.synchronize()
                track_ms.append(t0.elapsed_time(t1)); evals.append(tracker.report.evaluations)
                err.append(float(np.abs(trk.pose()[0] - frames[i].M).max()))
        recon.ProcessFrame(vis, scene, view, trk, rs, abi.FRAME_ALL)

    for i in range(Wm):
        frame(i, False)
    recon.Sync(scene, rs)
    sampler = ClockSampler(local_rank, rank, world); sampler.start()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = lib.dsb200_kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        frame(Wm + k, True)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    recon.Sync(scene, rs)
    launches = int(lib.dsb200_kernel_launch_count() - l0)
    ms = float(e0.elapsed_time(e1))
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        dist.destroy_process_group()
    if rank != 0:
        return None
    return {"metric": "frames/s track+alloc+integrate+raycast (closed loop, no external poses)", "value": world * K / (ms * 1e-3),
            "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 math (f64 tracker sums) -> i16 sdf / u8 weights", "data": "synthetic",
            "mode": "tracked",
            "config": {"workload": WORKLOADS[WORKLOAD]["text"] + "; poses from the ICP depth tracker (5 levels, ITMLibSettings defaults)",
                       "l2": "not flushed (closed loop: each frame depends on the previous one)"},
            "tracker": {"ms_per_frame": float(np.mean(track_ms)), "evaluations_per_frame": float(np.mean(evals)),
                        "pose_error_max_abs_entry": {"mean": float(np.mean(err)), "last": err[-1], "max": float(np.max(err))},
                        "final_visible_blocks": int(rs.c.no_visible)},
            "gpu_launches": launches, "clocks": clocks}


# ------------------------------------------------------------------------------------------------ sharded scene
C4_CAPS = dict(num_blocks=(1 << 22) + (1 << 20), bucket_num=1 << 23, excess_size=1 << 21)   # ~4 M blocks + margin: 20 GiB of voxels on one GPU
_FRAME_CACHE = {}
C4_TEXT = ("C4 large outdoor scene: the C2 camera (1242x375, 2 cm voxels, depth weighting) driven %d m along a street that repeats every "
           "%d m, fused into ONE scene of ~%.1f M voxel blocks partitioned by block ownership over %d GPU(s)")


def measure_sharded(args, rank, world, local_rank, mode, build_frames=None):
    """BASELINE configs[3]: ONE large scene partitioned by block ownership across the N GPUs (sharded.py).  Rank 0 owns the
    camera stream (the street's 64 distinct frames resident in its HBM); the map is first BUILT untimed (build_frames frames,
    ~2.4 k new blocks each), then K frames are timed: NCCL broadcast of the frame + per-shard allocate / integrate / expected
    depths + raycast (mode `composite`: own blocks, nearest hit per pixel composited slice-wise over peer memory; mode `lookup`:
    whole-scene march with owner-routed probes, bit-identical to one big scene) + ICP maps on the root.
    value = frames/s of the one scene.  Needs an initialised process group (world 1 included)."""
    import torch
    import torch.distributed as dist
    abi = importlib.import_module(PKG + ".abi")
    sharded = importlib.import_module(PKG + ".sharded")
    synth = importlib.import_module(PKG + ".synth")
    lib = abi.load()
    dev = torch.device("cuda", local_rank)
    K, Wm = args.steps, args.warmup
    build = args.shard_frames if build_frames is None else build_frames
    total = build + Wm + K
    cfg = synth.config("C4")
    poses = [synth.pose_to_M(p) for p in synth.trajectory(cfg, total)]
    caps = sharded.shard_caps(C4_CAPS, world, args.brick_log2, halo=(mode == "composite"))
    sc = sharded.DistShardedScene(cfg, caps, brick_log2=args.brick_log2, mode=mode)
    sc.m.recon.SetFusionWeightParams(cfg.depth_weighting, 10, 100)
    packed = None
    if rank == 0:      # only the camera rank renders (64 distinct images, once per process)
        if "C4" not in _FRAME_CACHE:
            _FRAME_CACHE["C4"] = synth.make_sequence("C4", cfg.period)[1]
        packed = [sc.pack_frame(f).to(dev) for f in _FRAME_CACHE["C4"]]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    sc.m.recon.SetAsync(True)

    def frame(i):
        sc.process_frame(packed[i % cfg.period] if rank == 0 else None, M=poses[i], timestamp=0.1 * i)
    for i in range(build + Wm):
        frame(i)
        if i % 64 == 63:
            sc.m.recon.Sync(sc.m.scene, sc.m.rs)       # keeps the snapshot reservations trimmed; raises if a shard ran out of blocks
    sc.m.recon.Sync(sc.m.scene, sc.m.rs)
    if sc.m.scene.c.last_free_block < 0 or sc.m.scene.c.last_free_excess < 0:
        raise RuntimeError("shard %d ran out of voxel blocks / excess entries while building the C4 map (capacity %r)" % (rank, caps))
    sampler = ClockSampler(local_rank, rank, world); sampler.start()      # before the barrier: rank 0 waits for nvidia-smi's first row
    torch.cuda.synchronize(); dist.barrier()
    l0 = lib.dsb200_kernel_launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    vis_counts = []
    for k in range(K):
        flush.zero_()
        evs[k][0].record()
        frame(build + Wm + k)
        evs[k][1].record()
        # The host reads the counters back only every 8th frame (outside the event pairs): with a read-back after every frame the
        # GPUs drain, and the next frame's pair then contains the enqueue latency of ~25 launches + 3 NCCL calls of the SLOWEST of
        # the N host processes (measured at N = 8: median step 0.40 ms, mean 0.64 ms, worst 4.6 ms -- profiles/r02s).
        if k % 8 == 7 or k + 1 == K:
            sc.m.recon.Sync(sc.m.scene, sc.m.rs)
            vis_counts.append(sc.m.rs.c.no_visible)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    launches = int(lib.dsb200_kernel_launch_count() - l0)
    dist.barrier()
    step_ms = np.array([a.elapsed_time(b) for a, b in evs])
    ms = float(step_ms.sum())
    blocks = caps["num_blocks"] - 1 - sc.m.scene.c.last_free_block
    t = torch.tensor([ms, float(blocks), float(np.mean(vis_counts))], dtype=torch.float64, device=dev)
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    per_shard = [torch.zeros(3, dtype=torch.float64, device=dev) for _ in range(world)]
    dist.all_gather(per_shard, t)
    sc.m.recon.SetAsync(False)
    sc.close()
    del flush, packed
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    ms = float(tmax[0])
    npx = cfg.W * cfg.H
    return {"metric": "frames/s alloc+integrate+raycast of ONE sharded scene (1242x375, 2 cm voxels, ~4 M blocks)", "value": K / (ms * 1e-3),
            "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "strong", "mode": "sharded/" + mode,
            "config": {"workload": C4_TEXT % (total, cfg.period, float(tsum[1]) / 1e6, world), "raycast": mode, "brick_blocks": 1 << args.brick_log2,
                       "build_frames": build, "per_shard_capacity": caps, "l2": "flushed (256 MiB write) between timed steps"},
            "blocks_allocated_total": int(tsum[1]), "blocks_allocated_per_shard": [int(x[1]) for x in per_shard],
            "visible_blocks_per_shard": [float(x[2]) for x in per_shard], "ms_per_shard": [float(x[0]) / K for x in per_shard],
            "collectives_per_step": {"broadcast_bytes": 8 * npx + 72, "rendezvous_allreduce_4B": 2,
                                     "peer_reads_bytes_per_gpu": (16 * npx) if mode == "composite" else None},
            "step_ms_rank0": {"p10": float(np.percentile(step_ms, 10)), "p50": float(np.percentile(step_ms, 50)),
                              "p90": float(np.percentile(step_ms, 90)), "max": float(step_ms.max()),
                              "all": [round(float(x), 3) for x in step_ms]},
            "gpu_launches": launches, "clocks": clocks}


def run_sharded(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    abi = importlib.import_module(PKG + ".abi")
    lib = abi.load()
    if lib.dsb200_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device; libdsb200 has no CPU fallback")
    abi.check(lib.dsb200_set_device(local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    init_group(world, dev)
    out = measure_sharded(args, rank, world, local_rank, args.shard_mode)
    dist.destroy_process_group()
    if out is not None:
        out.update({"vs_baseline": None, "dtype": "f32 math -> i16 sdf / u8 weights", "data": "synthetic"})
    return out


def init_group(world, dev):
    import torch.distributed as dist
    if dist.is_initialized():
        return
    if world == 1:     # single process: DistShardedScene still goes through the same code path (count = 1)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29555")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    else:
        dist.init_process_group("nccl", device_id=dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scenes", type=int, default=8, help="--mode batched: independent sequences per GPU")
    ap.add_argument("--mode", default="replicas", choices=["replicas", "sharded", "batched", "tracked"],
                    help="N > 1: independent scenes per GPU (default, weak scaling) or ONE scene sharded by block ownership")
    ap.add_argument("--shard-mode", default="lookup", choices=["lookup", "composite"])
    ap.add_argument("--brick-log2", type=int, default=3, help="sharded scene: ownership bricks of 2^k blocks per axis (halo replication ((B + 1) / B)^3)")
    ap.add_argument("--no-lookahead", action="store_true",
                    help="do not overlap the next frame's allocation request pass with the current frame's raycast")
    ap.add_argument("--no-sharded", action="store_true", help="skip the `sharded` object (the C4 scene split over the N GPUs)")
    ap.add_argument("--shard-frames", type=int, default=1650, help="frames fused untimed to build the C4 map (~2.4 k new blocks each)")
    ap.add_argument("--no-baselines", action="store_true", help="skip the gpu_comparator and cpu_baseline legs (tuning runs)")
    ap.add_argument("--single-workload", action="store_true",
                    help="N = 1: report only --workload (skip the `workloads` sub-objects for the other single-GPU configs)")
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS.keys()),
                    help="BASELINE.json config to run (default C2: the configuration the metric is quoted on)")
    args = ap.parse_args()
    global WORKLOAD, CAPS, METRIC
    WORKLOAD, CAPS = args.workload, WORKLOADS[args.workload]["caps"]
    METRIC = metric_text(WORKLOAD)
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # stdout carries exactly one JSON line: anything libraries print meanwhile (NCCL's version banner ...) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        out = run_reference(args, rank, world)
    elif args.mode == "sharded":
        out = run_sharded(args, rank, world, local_rank)
    elif args.mode == "batched":
        out = run_batched(args, rank, world, local_rank)
    elif args.mode == "tracked":
        out = run_tracked(args, rank, world, local_rank)
    else:
        out = run_ours(args, rank, world, local_rank)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if out is not None:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
