"""csrc/exactdiv.cuh: the FMA correction sequences that replace `/` in the integrate kernel are bit-identical to IEEE
division.  CPU leg: the same sequences evaluated with fmaf on the host over every integer divisor the kernel tabulates
(quick subset by default, full 2^23-significand sweep with DSB200_EXHAUSTIVE=1: ~15 s on 8 cores, result 0 mismatches).
GPU leg (-m gpu): the device helpers themselves against `/` on 2^28 operand tuples, and the plain-division kernel
variants against the fast ones through full fusion sequences."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import harness as hz
from harness import abi, synth


def test_fma_division_sequences_match_ieee_on_host(tmp_path):
    exe = str(tmp_path / "exactdiv_check")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fopenmp", os.path.join(hz.ROOT, "oracle", "exactdiv_check.c"), "-o", exe, "-lm"])
    mode = "0" if os.environ.get("DSB200_EXHAUSTIVE") else "1"
    out = subprocess.run([exe, mode], capture_output=True, text=True)
    bad1, bad2, n = [int(x) for x in out.stdout.split()]
    assert out.returncode == 0 and bad1 == 0 and bad2 == 0 and n > 1e7, out.stdout
    # numerators below the run-time-divisor path's window: the quotient may be inexact, `quotient + principal point` never is
    out = subprocess.run([exe, "2"], capture_output=True, text=True)
    bad, inexact, n = [int(x) for x in out.stdout.split()]
    assert out.returncode == 0 and bad == 0 and n == 20000000, out.stdout


@pytest.mark.gpu
def test_device_division_helpers_match_ieee():
    lib = abi.load()
    bad = C.c_ulonglong(123)
    abi.check(lib.dsb200_selftest_division(1 << 28, 7, C.byref(bad)))
    assert bad.value == 0


@pytest.mark.gpu
@pytest.mark.parametrize("cfg_name,weights", [("tiny", None), ("tiny_street", (True, 10, 100))])
def test_plain_division_variant_equals_fast_variant_and_oracle(cfg_name, weights):
    cfg, frames = synth.make_sequence(cfg_name, 5)
    fast, plain, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    abi.check(abi.load().dsb200_debug_force_plain_division(plain.recon.h, 1))
    for b in (fast, plain, ora):
        if weights:
            b.set_weights(*weights)
    for i, f in enumerate(frames):
        for b in (fast, plain, ora):
            b.set_frame(f); b.process_frame()
        hz.assert_state_equal(ora.state(), fast.state(), "fast frame %d" % i)
        hz.assert_state_equal(ora.state(), plain.state(), "plain frame %d" % i)
    # de-integration goes through the same helpers
    for b in (fast, plain, ora):
        b.set_frame(frames[1]); b.deintegrate()
    hz.assert_state_equal(ora.state(), fast.state(), "fast de-integration")
    hz.assert_state_equal(ora.state(), plain.state(), "plain de-integration")
    fast.close(); plain.close(); ora.close()


@pytest.mark.gpu
def test_pose_outside_fast_division_window_falls_back():
    # a pose entry of 1e-35 fails the host-side operand-window test -> plain-division kernels, same results
    cfg, frames = synth.make_sequence("tiny", 2)
    gpu, ora = hz.GpuBackend(cfg, hz.SMALL_CAPS), hz.OracleBackend(cfg, hz.SMALL_CAPS)
    for f in frames:
        M = f.M.copy(); M[1] = np.float32(1e-35)
        gpu.set_frame(f, M=M); ora.set_frame(f, M=M)
        gpu.process_frame(); ora.process_frame()
    hz.assert_state_equal(ora.state(), gpu.state(), "tiny pose entry")
    gpu.close(); ora.close()
