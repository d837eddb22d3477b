"""Times the two auxiliary paths that have no bench.py mode: the sensor-noise normals / uncertainties kernel and the block
swapping calls.  One JSON object on stdout.  Device timing with CUDA events on the stream the library launches on (the legacy
default stream; events come from libcudart through ctypes, so the script needs neither torch nor a minute of imports); L2 is
flushed between timed launches of the normals kernel.

    python tools/time_aux.py > profiles/<round>_aux_timings.json
"""
import ctypes as C
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "denseslam-global-consistency-h_b200"
abi = importlib.import_module(PKG + ".abi")
eng = importlib.import_module(PKG + ".engine")
synth = importlib.import_module(PKG + ".synth")


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 7700.0, "fallback (B200_PROFILING.md nominal HBM3e)"


class Cudart(object):
    def __init__(self):
        for name in ("libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so.12", "libcudart.so"):
            try:
                self.L = C.CDLL(name)
                break
            except OSError:
                continue
        else:
            raise RuntimeError("libcudart not found")
        self.L.cudaEventElapsedTime.argtypes = [C.POINTER(C.c_float), C.c_void_p, C.c_void_p]
        self.L.cudaEventRecord.argtypes = [C.c_void_p, C.c_void_p]
        self.L.cudaMemset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]

    def check(self, rc):
        if rc != 0:
            raise RuntimeError("cudart error %d" % rc)

    def event(self):
        e = C.c_void_p()
        self.check(self.L.cudaEventCreate(C.byref(e)))
        return e

    def record(self, e):
        self.check(self.L.cudaEventRecord(e, None))

    def sync(self):
        self.check(self.L.cudaDeviceSynchronize())

    def elapsed_ms(self, a, b):
        ms = C.c_float()
        self.check(self.L.cudaEventElapsedTime(C.byref(ms), a, b))
        return ms.value


def time_normals(lib, rt, W, H, reps=20):
    rng = np.random.default_rng(1)
    depth = eng.DeviceBuffer(W * H * 4); depth.upload(rng.uniform(0.5, 4.0, (H, W)).astype(np.float32))
    normal = eng.DeviceBuffer(W * H * 16); sigma = eng.DeviceBuffer(W * H * 4); flush = eng.DeviceBuffer(256 << 20)
    rt.check(rt.L.cudaMemset(normal.ptr, 0, W * H * 16)); rt.check(rt.L.cudaMemset(sigma.ptr, 0, W * H * 4))
    k = (C.c_float * 4)(1.0 / 525.0, 1.0 / 525.0, W / 2.0, H / 2.0)
    e0, e1 = rt.event(), rt.event()
    ms = []
    for i in range(reps + 3):
        rt.check(rt.L.cudaMemset(flush.ptr, i & 255, 256 << 20))
        rt.record(e0)
        abi.check(lib.dsb200_compute_normal_and_weights(depth.ptr, normal.ptr, sigma.ptr, W, H, k, None))
        rt.record(e1)
        rt.sync()
        if i >= 3:
            ms.append(rt.elapsed_ms(e0, e1))
    t = float(np.median(ms))
    for b in (depth, normal, sigma, flush):
        b.free()
    alg = 24.0 * W * H                      # 4 B depth read + 16 B normal + 4 B sigma written per pixel (DESIGN.md kernel table)
    peak, src = peaks()
    return {"image": [W, H], "ms": t, "algorithmic_bytes": alg, "GBps": alg / t / 1e6, "frac_of_hbm_peak": alg / t / 1e6 / peak,
            "peak": peak, "peak_source": src, "l2": "flushed (256 MiB memset) between launches"}


def time_swapping(lib, rt):
    """C1-sized scene (640x480, 5 mm voxels): fuse 4 frames, swap the whole map out in transfer-sized calls, let the allocator
    bring the visible part back and combine it.  Wall-clock per call (host store copies and PCIe included: that is the call)."""
    cfg, frames = synth.make_sequence("C1", 4)
    scene = eng.ITMScene(cfg.params(), use_swapping=True)
    recon = eng.ITMSceneReconstructionEngine()
    vis = eng.ITMVisualisationEngine(cfg.W, cfg.H)
    rs = vis.CreateRenderState(scene, (cfg.W, cfg.H))
    view = eng.ITMView(cfg.W, cfg.H, cfg.intr)
    trk = eng.ITMTrackingState(cfg.W, cfg.H)
    swap = eng.ITMSwappingEngine(scene, recon=recon)
    recon.ResetScene(scene)
    for f in frames:
        view.update(f.depth, f.rgb, f.timestamp); trk.set_pose(f.M)
        recon.AllocateSceneFromDepth(scene, view, trk, rs, False, False)
        swap.MarkVisible(rs)
        recon.IntegrateIntoScene(scene, view, trk, rs)
        swap.IntegrateGlobalIntoLocal(scene)
    out_calls = []
    while True:
        rt.sync(); t0 = time.perf_counter()
        n = swap.SaveToGlobalMemory(scene)
        dt = time.perf_counter() - t0
        if n == 0:
            break
        out_calls.append((n, dt))
    f = frames[-1]
    view.update(f.depth, f.rgb, f.timestamp); trk.set_pose(f.M)
    recon.AllocateSceneFromDepth(scene, view, trk, rs, False, False)
    swap.MarkVisible(rs)
    in_calls = []
    while True:
        rt.sync(); t0 = time.perf_counter()
        n = swap.IntegrateGlobalIntoLocal(scene)
        dt = time.perf_counter() - t0
        if n == 0:
            break
        in_calls.append((n, dt))
    def summary(calls):
        blocks = sum(n for n, _ in calls); secs = sum(dt for _, dt in calls)
        return {"calls": len(calls), "blocks": blocks, "ms_per_call": 1e3 * secs / max(len(calls), 1),
                "per_call": [[n, round(1e3 * dt, 4)] for n, dt in calls],
                "blocks_per_s": blocks / secs if secs else 0.0, "GBps": blocks * 4096 / secs / 1e9 if secs else 0.0}
    res = {"workload": "C1 640x480, 5 mm voxels, 4 frames fused, transfer limit 4096 blocks per call",
           "save_to_global_memory_whole_map": summary(out_calls), "integrate_global_into_local_visible_part": summary(in_calls),
           "timing": "host wall clock around each call (selection kernels, id read-back, block kernel moving 4 KB per block over PCIe)"}
    swap.close(); recon.close(); vis.close()
    for o in (scene, rs, view, trk):
        o.free()
    return res


def main():
    lib = abi.load()
    rt = Cudart()
    out = {"normals_and_weights": [time_normals(lib, rt, 640, 480), time_normals(lib, rt, 1242, 375), time_normals(lib, rt, 4096, 4096)],
           "swapping": time_swapping(lib, rt)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
