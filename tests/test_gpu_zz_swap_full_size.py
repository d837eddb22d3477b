"""Full-size leg of the swapping engine: a C1 map (640 x 480, 5 mm voxels, the reference's default capacities) swapped out whole
in transfer-sized calls and swapped back in where visible, held to the size-independent properties of
harness.check_swap_round_trip (which tests/test_swap_properties.py validates on the oracle).  Written after the GPU budget of
round 1 was spent: the calls themselves ran on the B200 (tools/time_aux.py), this file's assertions have not; it sorts last so
that nothing else depends on it."""
import pytest

import harness as hz
from harness import synth

pytestmark = pytest.mark.gpu


def test_c1_whole_map_swap_out_and_back():
    cfg, frames = synth.make_sequence("C1", 4)
    gpu = hz.GpuBackend(cfg, use_swapping=True)
    sw = hz.SwapOnGpu(gpu)
    moved, visible = hz.check_swap_round_trip(gpu, sw, frames, min_blocks=4000)
    assert moved > 0x1000 and visible > 1000          # more than one transfer-limited call each way
    sw.close(); gpu.close()
