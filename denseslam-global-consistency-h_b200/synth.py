"""Deterministic synthetic inputs for the fusion hot path.

Everything the engines consume is produced here from fixed seeds: analytic scenes ray-traced
to depth, sensor-like noise / holes, int16-mm quantisation followed by the reference's
convertDepthAffineToFloat (Engine/DeviceAgnostic/ITMViewBuilder.h:22-28: x * (1/1000) + 0,
<=0 or >32000 -> -1) so that oracle and GPU consume bit-identical float depth, a procedural
rgb texture, and world->camera poses (column-major float32, ORUtils/Matrix.h:7-11).
"""
import numpy as np

from . import abi


# ---------------------------------------------------------------- configs (BASELINE.json)
class Config(object):
    def __init__(self, name, W, H, intr, voxel_size, mu, max_w, vf_min, vf_max, scene, depth_weighting=False,
                 bf=None):
        self.name, self.W, self.H, self.intr = name, W, H, np.asarray(intr, np.float32)
        self.voxel_size, self.mu, self.max_w, self.vf_min, self.vf_max = voxel_size, mu, max_w, vf_min, vf_max
        self.scene, self.depth_weighting, self.bf = scene, depth_weighting, bf

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


def _raytrace(scene, origin, dirs, near_boxes=None):
    """Nearest positive hit distance t along dirs (N,3), origin (3,). dirs are scaled so that camera z == t."""
    o = origin[None, :]
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / dirs
        # room (inside): exit distance
        t1 = (scene.room_min[None, :] - o) * inv
        t2 = (scene.room_max[None, :] - o) * inv
        t = np.min(np.maximum(t1, t2), axis=1)
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
            tn = np.max(np.minimum(ta, tb), axis=1)
            tf = np.min(np.maximum(ta, tb), axis=1)
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
    z = _raytrace(scene, o, dw, near).reshape(H, W)
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
    pw = o[None, :] + dw * _raytrace(scene, o, dw, near)[:, None]
    chk = (np.floor(pw[:, 0] * 4) + np.floor(pw[:, 1] * 4) + np.floor(pw[:, 2] * 4)).astype(np.int64) & 1
    base = np.stack([90 + 120 * chk, 60 + 60 * ((np.floor(pw[:, 1] * 2).astype(np.int64)) & 1) + 70 * chk,
                     200 - 110 * chk], axis=-1)
    tex = np.clip(base + rng.integers(-12, 13, size=base.shape), 0, 255).astype(np.uint8)
    rgb = np.concatenate([tex, np.full((tex.shape[0], 1), 255, np.uint8)], axis=-1).reshape(H, W, 4)
    f = Frame()
    f.depth, f.rgb, f.M, f.timestamp, f.depth_mm = np.ascontiguousarray(depth), np.ascontiguousarray(rgb), \
        pose_to_M(cam_to_world), float(index) * 0.1, mm
    return f


def make_sequence(cfg_name, n_frames, seed=0, noise=True, start=0):
    cfg = config(cfg_name) if isinstance(cfg_name, str) else cfg_name
    scene = room_scene(seed) if cfg.scene == "room" else street_scene(seed)
    poses = trajectory(cfg, start + n_frames, seed)
    return cfg, [render_frame(cfg, scene, poses[start + i], start + i, seed, noise) for i in range(n_frames)]
