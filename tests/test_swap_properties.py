"""Size-independent properties of the swapping engine (harness.check_swap_round_trip) on the oracle, which is pinned to
ITMSwappingEngine_CPU; tests/test_gpu_zz_swap_full_size.py asks the same of the GPU path on a full-size C1 map."""
import harness as hz
from harness import synth


def test_oracle_whole_map_swap_out_and_back():
    for name, n, caps, transfer, min_blocks in (("tiny", 4, dict(num_blocks=8192, bucket_num=16384, excess_size=4096), 0x1000, 400),
                                                ("tiny_street", 3, dict(num_blocks=8192, bucket_num=16384, excess_size=4096), 128, 400)):
        cfg, frames = synth.make_sequence(name, n)
        ora = hz.OracleBackend(cfg, caps=caps, use_swapping=True)
        sw = hz.SwapOnOracle(ora, transfer)
        live, visible = hz.check_swap_round_trip(ora, sw, frames, min_blocks)
        assert live >= visible
        ora.close()
