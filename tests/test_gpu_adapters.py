"""The drop-in EXECUTED on the GPU: adapters/run_check.cu constructs the reference's own ITMScene / ITMView /
ITMTrackingState / ITMRenderState_VH with MEMORYDEVICE_CUDA, drives them through the *_B200 adapter subclasses in
ITMDenseMapper::ProcessFrame + ITMTrackingController::Prepare order and compares every frame, in the same process, with the
reference's ITMSceneReconstructionEngine_CPU / ITMVisualisationEngine_CPU (hash, free-list mirrors, visible list, voxels,
raycast, ICP maps: bit for bit); plus the free-view path, the view builder, pruning through the adapter, and the
exhaustion -> std::runtime_error contract.  The binary is built where /root/reference exists (`make -C adapters run_check`)
and travels to the GPU box in oracle/_ref/."""
import os
import struct
import subprocess

import numpy as np
import pytest

import harness as hz
from harness import synth

BIN = os.path.join(hz.ROOT, "oracle", "_ref", "adapter_run_check")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/adapter_run_check not built")]


def write_frames(path, cfg, frames):
    with open(path, "wb") as f:
        f.write(struct.pack("<3i", cfg.W, cfg.H, len(frames)))
        f.write(np.asarray(cfg.intr, np.float32).tobytes())
        f.write(struct.pack("<2fi2f", cfg.voxel_size, cfg.mu, cfg.max_w, cfg.vf_min, cfg.vf_max))
        for fr in frames:
            f.write(np.ascontiguousarray(fr.depth, np.float32).tobytes())
            f.write(np.ascontiguousarray(fr.rgb, np.uint8).tobytes())
            f.write(np.ascontiguousarray(fr.M, np.float32).tobytes())
            f.write(struct.pack("<d", fr.timestamp))


@pytest.mark.parametrize("name,n", [("tiny", 6), ("tiny_street", 4)])
def test_adapters_run_on_gpu_and_match_reference_cpu_engines(tmp_path, name, n):
    cfg, frames = synth.make_sequence(name, n)
    path = str(tmp_path / "frames.bin")
    write_frames(path, cfg, frames)
    out = subprocess.run([BIN, path], capture_output=True, text=True, timeout=900)
    print(out.stdout[-4000:])
    assert out.returncode == 0, out.stdout[-4000:] + out.stderr[-2000:]
    assert "adapters: OK" in out.stdout
