"""Golden vectors generated from the UNMODIFIED reference CPU engines (tests/golden/make_golden.py).

CPU leg: the oracle (oracle/restate.cpp) must reproduce every digest and every stored array -> the oracle
is pinned to reference outputs even where /root/reference (and oracle/_ref) do not exist.
GPU leg (-m gpu): libdsb200 through the C ABI must reproduce the same reference outputs directly.
The fixtures carry their own inputs, so neither leg depends on the synthetic generator staying stable;
test_fixture_inputs_match_generator only reports generator drift."""
import os
import sys

import numpy as np
import pytest

import harness as hz
from harness import abi, synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLDEN)
import make_golden as mg  # noqa: E402

FIXTURES = ["tiny_room_ref.npz", "tiny_room_swapping_ref.npz", "tiny_street_ref.npz"]


def load(name):
    z = np.load(os.path.join(GOLDEN, name))
    cfg = synth.Config("fixture:" + name, int(z["W"]), int(z["H"]), z["intr"], float(z["voxel_size"]), float(z["mu"]), int(z["max_w"]),
                       float(z["vf_min"]), float(z["vf_max"]), "fixture")
    frames = []
    for i in range(int(z["n_frames"])):
        f = synth.Frame()
        mm = z["depth_mm"][i]
        # convertDepthAffineToFloat, Engine/DeviceAgnostic/ITMViewBuilder.h:22-28
        f.depth = np.where((mm <= 0) | (mm > 32000), np.float32(-1.0), mm.astype(np.float32) * np.float32(1.0 / 1000.0) + np.float32(0.0)).astype(np.float32)
        f.rgb = np.ascontiguousarray(np.concatenate([z["rgb"][i], np.full(mm.shape + (1,), 255, np.uint8)], -1))
        f.M, f.timestamp, f.depth_mm = z["M"][i].astype(np.float32), float(z["timestamps"][i]), mm
        frames.append(f)
    return z, cfg, frames


def check_backend_against_fixture(back, z, cfg, frames):
    skeys, ikeys = list(z["state_keys"]), list(z["image_keys"])
    for i, f in enumerate(frames):
        back.set_frame(f)
        assert back.process_frame() == 0
        s = back.state(copy=False) if back.name != "gpu" else back.state()      # oracle: views of the live arrays, digested at once
        d = mg.state_digests(s["hash"], s["voxels"], s["visible"], s["visible_type"], s["last_free"])
        for k, want in zip(skeys, z["state_digests"][i]):
            assert d[k] == want, "frame %d: %s differs from the reference" % (i, k)
        im = back.images()
        d = mg.image_digests(im["range"], im["raycast"], im["points"], im["normals"], im["rendering"], cfg.W, cfg.H)
        for k, want in zip(ikeys, z["image_digests"][i]):
            assert d[k] == want, "frame %d: image %s differs from the reference" % (i, k)
        assert len(s["visible"]) == int(z["no_visible"][i])
    # last frame, element by element (so a digest mismatch above can be localised)
    live = z["hash_live"]
    assert np.array_equal(np.nonzero(s["hash"]["ptr"] != -2)[0], live)
    assert np.array_equal(s["hash"]["pos"][live], z["hash_pos"]) and np.array_equal(s["hash"]["offset"][live], z["hash_offset"])
    assert np.array_equal(s["hash"]["ptr"][live], z["hash_ptr"])
    assert np.array_equal(s["visible"], z["visible"])
    assert np.array_equal(np.nonzero(s["visible_type"])[0], z["vt_idx"]) and np.array_equal(s["visible_type"][z["vt_idx"]], z["vt_val"])
    assert tuple(s["last_free"]) == tuple(z["last_free"])
    H8, W8 = cfg.H // 8 + 1, cfg.W // 8 + 1
    hz.assert_images_equal(dict(range=z["range"], raycast=z["raycast"], points=z["points"], normals=z["normals"], rendering=z["rendering"]),
                           dict(range=im["range"][:H8, :W8], raycast=im["raycast"], points=im["points"], normals=im["normals"],
                                rendering=im["rendering"][..., 0]), "last frame")
    # free-view render
    back.free_view(z["free_M"], cfg.intr, abi.RENDER_SHADED_GREYSCALE)
    fi = back.free_images()
    assert np.array_equal(fi["visible"], z["free_visible"])
    assert mg.sha(fi["raycast"]) == str(z["free_raycast_sha"])
    assert np.array_equal(fi["image"][..., 0], z["free_image"])


@pytest.mark.parametrize("name", FIXTURES)
def test_oracle_reproduces_reference_golden(name):
    z, cfg, frames = load(name)
    ora = hz.OracleBackend(cfg, use_swapping=bool(int(z["use_swapping"])))
    check_backend_against_fixture(ora, z, cfg, frames)
    ora.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", FIXTURES)
def test_gpu_reproduces_reference_golden(name):
    z, cfg, frames = load(name)
    gpu = hz.GpuBackend(cfg, use_swapping=bool(int(z["use_swapping"])))
    check_backend_against_fixture(gpu, z, cfg, frames)
    gpu.close()


def test_fixture_inputs_match_generator():
    z, cfg, frames = load("tiny_room_ref.npz")
    c2, f2 = synth.make_sequence("tiny", len(frames))
    assert np.array_equal(np.asarray(c2.intr), z["intr"])
    for a, b in zip(frames, f2):
        assert np.array_equal(a.depth, b.depth) and np.array_equal(a.rgb, b.rgb) and np.array_equal(a.M, b.M)
