"""The C-ABI library loads and exports every symbol include/dsb200.h declares (no GPU needed)."""
import ctypes as C
import os
import re

import pytest

import harness as hz
from harness import abi

HEADER = os.path.join(hz.ROOT, "include", "dsb200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dsb200_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(abi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return abi.load()


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 40
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, "declared in dsb200.h but not exported: %s" % missing
    # and the ctypes mirror declares a prototype for each of them
    assert sorted(lib._declared) == syms


def test_struct_layouts_match_reference_sizes():
    assert abi.HASH_DTYPE.itemsize == 20      # sizeof(ITMHashEntry), SURVEY 2.4
    assert abi.VOXEL_DTYPE.itemsize == 8      # sizeof(ITMVoxel_s_rgb)
    assert C.sizeof(abi.SceneParams) == 24
    assert C.sizeof(abi.WeightParams) == 12
    assert abi.DEFAULT_BUCKETS + abi.DEFAULT_EXCESS == 1441792   # ITMVoxelBlockHash::noTotalEntries


def test_product_fails_loudly_without_gpu(lib):
    if lib.dsb200_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.dsb200_recon_create(C.byref(h), 1024, 1024, 256, None)
    assert rc == abi.ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.dsb200_last_error()
    with pytest.raises(abi.Dsb200Error):
        abi.check(rc)


def test_host_matrix_helpers_match_oracle(lib):
    import numpy as np
    o = hz.po.Oracle()
    rng = np.random.default_rng(5)
    for _ in range(50):
        M = rng.normal(size=16).astype(np.float32)
        out = (C.c_float * 16)()
        assert lib.dsb200_mat4_inverse(abi.f16(M), out)
        assert np.array_equal(np.array(out[:], np.float32).view(np.uint32), o.mat4_inverse(M).view(np.uint32))
        B = rng.normal(size=16).astype(np.float32)
        lib.dsb200_mat4_multiply(abi.f16(M), abi.f16(B), out)
        assert np.array_equal(np.array(out[:], np.float32).view(np.uint32), o.mat4_multiply(M, B).view(np.uint32))


def test_host_pose_helpers_match_reference_fixture_and_oracle(lib):
    """dsb200_pose_from_params / dsb200_pose_coerce are host code (csrc/tracker_host.cpp): no GPU needed.  Bit-exact with the
    reference's ITMPose known answers in tests/golden/tiny_tracker_ref.npz and with the oracle over every angle branch."""
    import importlib
    import numpy as np
    eng = importlib.import_module(hz.PKG + ".engine")
    t = np.load(os.path.join(hz.ROOT, "tests", "golden", "tiny_tracker_ref.npz"))
    bits = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32)
    for p, M, cM, cp in zip(t["pose_params"], t["pose_M"], t["pose_coerced_M"], t["pose_coerced_params"]):
        assert np.array_equal(bits(eng.pose_from_params(p)), bits(M))
        m2, p2 = eng.pose_coerce(M)
        assert np.array_equal(bits(m2), bits(cM)) and np.array_equal(bits(p2), bits(cp))
    o = hz.po.Oracle()
    rng = np.random.default_rng(13)
    for k in range(1200):
        p = np.concatenate([rng.normal(size=3), rng.normal(size=3) * [1e-5, 5e-4, 0.05, 1.0, 1.8, 3.1][k % 6]]).astype(np.float32)
        M = eng.pose_from_params(p)
        assert np.array_equal(bits(M), bits(o.pose_from_params(p)))
        noisy = M + rng.normal(size=16).astype(np.float32) * np.float32(1e-4) * (np.arange(16) % 4 != 3)
        a, b = eng.pose_coerce(noisy), o.pose_coerce(noisy)
        assert np.array_equal(bits(a[0]), bits(b[0])) and np.array_equal(bits(a[1]), bits(b[1]))


def test_tracker_entry_points_fail_loudly_without_gpu(lib):
    if lib.dsb200_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.dsb200_tracker_create(C.byref(h), 64, 48, 3, None)
    assert rc < 0 and not h.value
    assert lib.dsb200_tracker_create(C.byref(h), 64, 48, 0, None) == abi.ERR_INVALID      # level count checked before any CUDA call
    assert lib.dsb200_recon_join_prefetch(None) == abi.ERR_INVALID


def test_swap_and_view_entry_points_fail_loudly_without_gpu(lib):
    """argument checks come before any CUDA call; with no device the constructors report it instead of falling back"""
    assert lib.dsb200_swap_mark_visible(None, None) == abi.ERR_INVALID
    assert lib.dsb200_swap_integrate_global_into_local(None, None, None) == abi.ERR_INVALID
    assert lib.dsb200_swap_save_to_global_memory(None, None, None, None) == abi.ERR_INVALID
    assert lib.dsb200_swap_save_file(None, b"/tmp/x") == abi.ERR_INVALID and lib.dsb200_swap_load_file(None, None) == abi.ERR_INVALID
    assert not lib.dsb200_swap_stored_block(None, 0)
    assert lib.dsb200_swap_destroy(None) == abi.OK
    assert lib.dsb200_recon_notify_scene_modified(None) == abi.ERR_INVALID
    assert lib.dsb200_compute_normal_and_weights(None, None, None, 64, 48, None, None) == abi.ERR_INVALID
    h = C.c_void_p()
    assert lib.dsb200_swap_create(C.byref(h), 0, 16, None) == abi.ERR_INVALID
    if lib.dsb200_device_count() > 0:
        pytest.skip("a GPU is present")
    assert lib.dsb200_swap_create(C.byref(h), 1024, 16, None) == abi.ERR_NO_DEVICE and not h.value
    assert b"no CPU fallback" in lib.dsb200_last_error()
    x = (C.c_float * 4)(0.1, 0.2, 0.3, 0.4)
    assert lib.dsb200_selftest_acosf(x, x, 4) == abi.ERR_NO_DEVICE
