"""Deterministic synthetic inputs for the fusion hot path.

Everything the engines consume is produced here from fixed seeds: analytic scenes ray-traced
to depth, sensor-like noise / holes, int16-mm quantisation followed by the reference's
convertDepthAffineToFloat (Engine/DeviceAgnostic/ITMViewBuilder.h:22-28: x * (1/1000) + 0,
<=0 or >32000 -> -1) so that oracle and GPU consume bit-identical float depth, a procedural
rgb texture, and world->camera poses (column-major float32, ORUtils/Matrix.h:7-11).
"""
import numpy as np

from . import abi


C4_PERIOD = 64     # metres (= frames at 1 m per frame) after which the C4 street, its trajectory and its sensor noise repeat


# ---------------------------------------------------------------- configs (BASELINE.json)
class Config(object):
    def __init__(self, name, W, H, intr, voxel_size, mu, max_w, vf_min, vf_max, scene, depth_weighting=False,
                 bf=None):
        self.name, self.W, self.H, self.intr = name, W, H, np.asarray(intr, np.float32)
        self.voxel_size, self.mu, self.max_w, self.vf_min, self.vf_max = voxel_size, mu, max_w, vf_min, vf_max
        self.scene, self.depth_weighting, self.bf = scene, depth_weighting, bf
        self.period = None      # periodic street (C4): only `period` distinct images exist, frame i shows image i % period

    def params(self):
        return abi.SceneParams(self.voxel_size, self.mu, self.max_w, self.vf_min, self.vf_max, 0)


def config(name):
    if name == "C1":   # ICL-NUIM-shaped, 5 mm voxels
        return Config("C1", 640, 480, (481.2, 480.0, 319.5, 239.5), 0.005, 0.02, 100, 0.2, 3.0, "room")
    if name == "C2":   # KITTI-stereo-shaped, 2 cm voxels (intrinsics: ORB-SLAM2 KITTI00-02.yaml)
        return Config("C2", 1242, 375, (718.856, 718.856, 607.1928, 185.2157), 0.02, 0.08, 100, 0.5, 30.0, "street",
                      depth_weighting=True, bf=386.1448)
    if name == "C3":   # TUM-shaped (TUM1.yaml), decay + slide window
        return Config("C3", 640, 480, (517.306, 516.469, 318.643, 255.314), 0.005, 0.02, 100, 0.2, 3.0, "room")
    if name == "C4":   # large outdoor scene (BASELINE configs[3]): the C2 camera on a street that repeats every PERIOD metres
        c = Config("C4", 1242, 375, (718.856, 718.856, 607.1928, 185.2157), 0.02, 0.08, 100, 0.5, 30.0, "street",
                   depth_weighting=True, bf=386.1448)
        c.period = C4_PERIOD
        return c
    if name == "tiny_loop":  # the same construction at CPU-test size
        c = Config("tiny_loop", 312, 96, (180.0, 180.0, 152.3, 47.6), 0.08, 0.32, 100, 0.5, 30.0, "street", depth_weighting=True, bf=96.7)
        c.period = 16
        return c
    if name == "tiny":  # small CPU-test configuration (coarse voxels so the oracle is fast)
        return Config("tiny", 160, 120, (120.3, 120.0, 79.5, 59.5), 0.02, 0.08, 100, 0.2, 3.0, "room")
    if name == "tiny_street":
        return Config("tiny_street", 312, 96, (180.0, 180.0, 152.3, 47.6), 0.08, 0.32, 100, 0.5, 30.0, "street",
                      depth_weighting=True, bf=96.7)
    raise KeyError(name)


# ---------------------------------------------------------------- analytic scenes
class Scene(object):
    """Room: camera inside an axis-aligned box, plus spheres and solid boxes."""

    def __init__(self, room_min, room_max, spheres, boxes):
        self.room_min = np.asarray(room_min, np.float64)
        self.room_max = np.asarray(room_max, np.float64)
        self.spheres = [(np.asarray(c, np.float64), float(r)) for c, r in spheres]
        self.boxes = [(np.asarray(a, np.float64), np.asarray(b, np.float64)) for a, b in boxes]


def room_scene(seed=0):
    rng = np.random.default_rng(20240501 + seed)
    spheres = []
    for _ in range(6):
        c = rng.uniform([-1.6, -0.2, -1.2], [1.6, 1.0, 1.2])
        spheres.append((c, rng.uniform(0.15, 0.35)))
    boxes = []
    for _ in range(2):
        c = rng.uniform([-1.5, 0.6, -1.0], [1.5, 1.0, 1.0])
        s = rng.uniform(0.2, 0.4, size=3)
        boxes.append((c - s, c + s))
    return Scene([-2.5, -1.25, -2.0], [2.5, 1.25, 2.0], spheres, boxes)


def street_scene(seed=0, length=1100.0):
    rng = np.random.default_rng(20240501 + 7 + seed)
    boxes = []
    z = 8.0
    while z < length:
        side = -1.0 if rng.random() < 0.5 else 1.0
        cx = side * rng.uniform(2.5, 5.5)
        c = np.array([cx, 1.65 - 0.75, z])
        s = np.array([0.9, 0.75, 2.1])
        boxes.append((c - s, c + s))
        z += rng.uniform(6.0, 18.0)
    return Scene([-8.0, -12.0, -60.0], [8.0, 1.65, length + 60.0], [], boxes)


def periodic_street_scene(period, seed=0):
    """a street whose parked boxes repeat every `period` metres: boxes of one period, replicated over periods -1 .. 2 so that a
    camera anywhere in period 0 sees the same world as a camera `period` metres further on"""
    rng = np.random.default_rng(20240501 + 11 + seed)
    base = []
    z = 2.0
    while z < period - 3.0:
        side = -1.0 if rng.random() < 0.5 else 1.0
        cx = side * rng.uniform(2.5, 5.5)
        c = np.array([cx, 1.65 - 0.75, z])
        s_ = np.array([0.9, 0.75, 2.1])
        base.append((c - s_, c + s_))
        z += rng.uniform(6.0, 14.0)
    boxes = []
    for k in (-1, 0, 1, 2):
        off = np.array([0.0, 0.0, k * float(period)])
        boxes += [(a + off, b + off) for a, b in base]
    return Scene([-8.0, -12.0, -3.0 * period], [8.0, 1.65, 4.0 * period], [], boxes)


def _min3(a):     # np.min(a, axis=1) for (N, 3): column-wise is several times faster than a reduction along a length-3 axis
    return np.minimum(np.minimum(a[:, 0], a[:, 1]), a[:, 2])


def _max3(a):
    return np.maximum(np.maximum(a[:, 0], a[:, 1]), a[:, 2])


def _raytrace(scene, origin, dirs, near_boxes=None):
    """Nearest positive hit distance t along dirs (N,3), origin (3,). dirs are scaled so that camera z == t."""
    o = origin[None, :]
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / dirs
        # room (inside): exit distance
        t1 = (scene.room_min[None, :] - o) * inv
        t2 = (scene.room_max[None, :] - o) * inv
        t = _min3(np.maximum(t1, t2))
        for c, r in scene.spheres:
            oc = origin - c
            a = np.sum(dirs * dirs, axis=1)
            b = 2.0 * dirs @ oc
            cc = oc @ oc - r * r
            disc = b * b - 4 * a * cc
            ts = (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a)
            ok = (disc > 0) & (ts > 1e-6)
            t = np.where(ok & (ts < t), ts, t)
        boxes = scene.boxes if near_boxes is None else near_boxes
        for a_, b_ in boxes:
            ta = (a_[None, :] - o) * inv
            tb = (b_[None, :] - o) * inv
            tn = _max3(np.minimum(ta, tb))
            tf = _min3(np.maximum(ta, tb))
            ok = (tn < tf) & (tn > 1e-6)
            t = np.where(ok & (tn < t), tn, t)
    return t


def look_at(eye, target, up=(0.0, -1.0, 0.0)):
    """camera->world 4x4 (x right, y down, z forward)."""
    eye, target, up = (np.asarray(v, np.float64) for v in (eye, target, up))
    z = target - eye
    z /= np.linalg.norm(z)
    x = np.cross(-up, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    T = np.eye(4)
    T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = x, y, z, eye
    return T


def trajectory(cfg, n_frames, seed=0):
    """List of camera->world float64 4x4."""
    poses = []
    if cfg.scene == "room":
        rng = np.random.default_rng(99 + seed)
        jitter = cfg.name == "C3"
        for i in range(n_frames):
            ang = 2 * np.pi * (i / 100.0) + 0.37 * seed
            eye = np.array([1.2 * np.cos(ang), -0.1 + 0.15 * np.sin(2 * ang), 1.2 * np.sin(ang) * 0.8])
            tgt = np.array([0.0, 0.2, 0.0]) - eye * 0.8
            if jitter:
                eye = eye + rng.normal(0, 0.01, 3)
                tgt = tgt + rng.normal(0, 0.03, 3)
            poses.append(look_at(eye, tgt))
    elif cfg.period:
        for i in range(n_frames):      # sway and yaw repeat with the street
            yaw = np.deg2rad(2.0) * np.sin(2 * np.pi * i / float(cfg.period))
            eye = np.array([0.3 * np.sin(2 * np.pi * i / float(cfg.period)), 0.0, 1.0 * i])
            tgt = eye + np.array([np.sin(yaw), 0.0, np.cos(yaw)])
            poses.append(look_at(eye, tgt))
    else:
        for i in range(n_frames):
            yaw = np.deg2rad(2.0) * np.sin(2 * np.pi * i / 60.0)
            eye = np.array([0.3 * np.sin(i / 25.0), 0.0, 1.0 * i])
            tgt = eye + np.array([np.sin(yaw), 0.0, np.cos(yaw)])
            poses.append(look_at(eye, tgt))
    return poses


def pose_to_M(cam_to_world):
    """world->camera, column-major float32[16] (what pose_d->GetM() holds)."""
    M = np.linalg.inv(cam_to_world)
    return np.ascontiguousarray(M.T.reshape(-1), dtype=np.float32)   # column-major == transpose flattened


class Frame(object):
    __slots__ = ("depth", "rgb", "M", "timestamp", "depth_mm")


def render_frame(cfg, scene, cam_to_world, index=0, seed=0, noise=True):
    W, H = cfg.W, cfg.H
    fx, fy, cx, cy = [float(v) for v in cfg.intr]
    xs, ys = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    dc = np.stack([(xs - cx) / fx, (ys - cy) / fy, np.ones_like(xs)], axis=-1).reshape(-1, 3)
    R, o = cam_to_world[:3, :3], cam_to_world[:3, 3]
    dw = dc @ R.T
    near = None
    if cfg.scene == "street":
        near = [b for b in scene.boxes if -5.0 < b[0][2] - o[2] < 45.0]
    t_hit = _raytrace(scene, o, dw, near)
    z = t_hit.reshape(H, W)
    rng = np.random.default_rng(1000003 * (seed + 1) + index)
    if cfg.scene == "street":
        if noise:
            disp = cfg.bf / np.maximum(z, 1e-3)
            disp = np.round(disp * 16.0 + rng.normal(0, 0.5, z.shape)) / 16.0   # 1/16 px quantisation + matching noise
            z = cfg.bf / np.maximum(disp, 1e-3)
        z = np.where((z < 0.5) | (z > 30.0), 0.0, z)
    elif noise:
        z = z + rng.normal(0, 1.0, z.shape) * (0.0012 + 0.0019 * (z - 0.4) ** 2)
    if noise:
        z = np.where(rng.random(z.shape) < 0.02, 0.0, z)   # 2 % holes
    mm = np.clip(np.round(z * 1000.0), 0, 32767).astype(np.int16)
    depth = np.where((mm <= 0) | (mm > 32000), np.float32(-1.0),
                     mm.astype(np.float32) * np.float32(1.0 / 1000.0) + np.float32(0.0)).astype(np.float32)
    # procedural texture from the world hit point (checker + hash noise)
    pw = o[None, :] + dw * t_hit[:, None]
    chk = (np.floor(pw[:, 0] * 4) + np.floor(pw[:, 1] * 4) + np.floor(pw[:, 2] * 4)).astype(np.int64) & 1
    base = np.stack([90 + 120 * chk, 60 + 60 * ((np.floor(pw[:, 1] * 2).astype(np.int64)) & 1) + 70 * chk,
                     200 - 110 * chk], axis=-1)
    tex = np.clip(base + rng.integers(-12, 13, size=base.shape), 0, 255).astype(np.uint8)
    rgb = np.concatenate([tex, np.full((tex.shape[0], 1), 255, np.uint8)], axis=-1).reshape(H, W, 4)
    f = Frame()
    f.depth, f.rgb, f.M, f.timestamp, f.depth_mm = np.ascontiguousarray(depth), np.ascontiguousarray(rgb), \
        pose_to_M(cam_to_world), float(index) * 0.1, mm
    return f


def _render_job(job):
    name, index, seed, noise, periodic = job
    cfg = config(name)
    if periodic:
        return render_frame(cfg, periodic_street_scene(cfg.period, seed), trajectory(cfg, cfg.period, seed)[index], index, seed, noise)
    scene = room_scene(seed) if cfg.scene == "room" else street_scene(seed)
    return render_frame(cfg, scene, trajectory(cfg, index + 1, seed)[index], index, seed, noise)


def _render_many(jobs):
    """frames are independent: render them on the host's cores, in worker PROCESSES started with subprocess (no fork of a process
    that may hold a CUDA context, no multiprocessing re-import of the caller's __main__).  Same code, same seeds, same results as
    the serial loop; any failure of a worker falls back to it."""
    import os
    import pickle
    import subprocess
    import sys
    import tempfile
    workers = min(16, os.cpu_count() or 1, len(jobs) // 4)
    if workers < 2 or os.environ.get("DSB200_SYNTH_SERIAL"):
        return [_render_job(j) for j in jobs]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys, pickle, importlib; sys.path.insert(0, %r); m = importlib.import_module(%r); "
            "jobs = pickle.load(open(sys.argv[1], 'rb')); pickle.dump([m._render_job(j) for j in jobs], open(sys.argv[2], 'wb'), protocol=4)"
            % (root, __name__))
    try:
        with tempfile.TemporaryDirectory() as d:
            procs = []
            for w in range(workers):
                part = jobs[w::workers]
                fin, fout = os.path.join(d, "in%d" % w), os.path.join(d, "out%d" % w)
                pickle.dump(part, open(fin, "wb"))
                env = dict(os.environ, DSB200_SYNTH_SERIAL="1", OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
                procs.append((subprocess.Popen([sys.executable, "-c", code, fin, fout], env=env, stdout=subprocess.DEVNULL), fout, w))
            out = [None] * len(jobs)
            for p, fout, w in procs:
                if p.wait() != 0:
                    raise RuntimeError("render worker failed")
                for k, f in enumerate(pickle.load(open(fout, "rb"))):
                    out[w + k * workers] = f
            return out
    except Exception:
        return [_render_job(j) for j in jobs]


def make_sequence(cfg_name, n_frames, seed=0, noise=True, start=0):
    cfg = config(cfg_name) if isinstance(cfg_name, str) else cfg_name
    poses = trajectory(cfg, start + n_frames, seed)
    named = isinstance(cfg_name, str)
    if cfg.period:
        # periodic street: render the `period` distinct images once (camera inside period 0), then every frame i is image
        # i % period seen from pose i.  The frames SHARE their image arrays.
        scene = periodic_street_scene(cfg.period, seed)
        base_poses = trajectory(cfg, cfg.period, seed)
        need = sorted({i % cfg.period for i in range(start, start + n_frames)})
        if named and len(need) >= 8:
            images = dict(zip(need, _render_many([(cfg_name, k, seed, noise, True) for k in need])))
        else:
            images = {k: render_frame(cfg, scene, base_poses[k], k, seed, noise) for k in need}
        out = []
        for i in range(start, start + n_frames):
            k = i % cfg.period
            b = images[k]
            f = Frame()
            f.depth, f.rgb, f.depth_mm, f.M, f.timestamp = b.depth, b.rgb, b.depth_mm, pose_to_M(poses[i]), float(i) * 0.1
            out.append(f)
        return cfg, out
    if named and n_frames >= 8 and cfg.W * cfg.H >= 200000:
        return cfg, _render_many([(cfg_name, start + i, seed, noise, False) for i in range(n_frames)])
    scene = room_scene(seed) if cfg.scene == "room" else street_scene(seed)
    return cfg, [render_frame(cfg, scene, poses[start + i], start + i, seed, noise) for i in range(n_frames)]
