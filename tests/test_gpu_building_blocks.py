"""GPU checks of the two bit-exactness building blocks: the device atan2f (must equal glibc's atan2f of this
image bit for bit) and the device replay of libstdc++ std::sort (must give std::sort's permutation)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from lego_loam_b200 import capi

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _libm_atan2f(y, x):
    libm = ctypes.CDLL("libm.so.6")
    libm.atan2f.restype = ctypes.c_float
    libm.atan2f.argtypes = [ctypes.c_float, ctypes.c_float]
    return np.array([libm.atan2f(float(a), float(b)) for a, b in zip(y, x)], np.float32)


def test_device_atan2f_equals_glibc():
    rng = np.random.default_rng(5)
    n = 400000
    y = rng.uniform(-100, 100, n).astype(np.float32)
    x = rng.uniform(-100, 100, n).astype(np.float32)
    y[::7] *= np.float32(1e-6); x[::11] *= np.float32(1e-5)
    # segmentation form d2*sin(a), d1 - d2*cos(a)
    d1 = rng.uniform(1, 100, n // 4).astype(np.float32); d2 = (d1 * rng.uniform(0.5, 1.0, n // 4)).astype(np.float32)
    sa, ca = np.float32(np.sin(np.float32(0.0034906585))), np.float32(np.cos(np.float32(0.0034906585)))
    y[: n // 4] = d2 * sa; x[: n // 4] = d1 - d2 * ca
    sp = np.array([[0, 1], [-0.0, 1], [0, -1], [-0.0, -1], [1, 0], [-1, 0], [1, -0.0], [np.inf, 1], [1, np.inf], [1, -np.inf],
                   [-np.inf, -np.inf], [1e30, 1e-30], [1e-30, -1e30], [3, 1]], np.float32)
    y = np.concatenate([y, sp[:, 0]]); x = np.concatenate([x, sp[:, 1]])
    got = capi.selftest_atan2f(y, x)
    want = _libm_atan2f(y, x)
    bad = np.flatnonzero(got.view(np.uint32) != want.view(np.uint32))
    assert bad.size == 0, "%d mismatches, first: y=%r x=%r got=%r want=%r" % (bad.size, y[bad[0]], x[bad[0]], got[bad[0]], want[bad[0]])


def test_device_std_sort_replay_equals_libstdcxx(tmp_path):
    exe = str(tmp_path / "std_sort_cli")
    subprocess.check_call(["g++", "-O2", os.path.join(ROOT, "tests", "native", "std_sort_cli.cpp"), "-o", exe])
    rng = np.random.default_rng(11)
    segs = []
    for c in range(3000):
        n = int(rng.integers(1, 420)) if c % 40 else int(rng.integers(1, 2100))
        mode = c % 6
        if mode == 0: v = rng.integers(0, 5, n).astype(np.float32)
        elif mode == 1: v = (rng.integers(-1000, 1001, n).astype(np.float32) * np.float32(1e-3)) ** 2
        elif mode == 2: v = np.arange(n, dtype=np.float32)
        elif mode == 3: v = np.arange(n, 0, -1).astype(np.float32)
        elif mode == 4:
            k = n // 2
            v = np.concatenate([np.where(np.arange(k) % 2 == 0, np.arange(k) + 1, k + np.arange(k)), (np.arange(n - k) + 1) * 2]).astype(np.float32)
        else: v = rng.random(n).astype(np.float32)
        segs.append(v)
    offsets = np.concatenate([[0], np.cumsum([len(s) for s in segs])]).astype(np.int32)
    values = np.concatenate(segs)
    blob = np.int32(len(segs)).tobytes() + offsets.tobytes() + values.tobytes()
    want = np.frombuffer(subprocess.run([exe], input=blob, stdout=subprocess.PIPE, check=True).stdout, np.int32)
    got = capi.selftest_std_sort(values, offsets)
    assert np.array_equal(got, want)
    # the warp-parallel statement fa_sort_sectors_kernel uses (segments of at most 416 values)
    keep = [i for i, s in enumerate(segs) if len(s) <= 416]
    sub_off = np.concatenate([[0], np.cumsum([len(segs[i]) for i in keep])]).astype(np.int32)
    sub_val = np.concatenate([segs[i] for i in keep])
    sub_want = np.concatenate([want[offsets[i]:offsets[i + 1]] for i in keep])
    assert np.array_equal(capi.selftest_warp_sort(sub_val, sub_off), sub_want)


def _vg_cases():
    rng = np.random.default_rng(3)
    clouds = []
    # wall-like slabs, dense duplicates, single point, empty, wide cloud, > 1 tile, > 2^16 voxels
    for n, span, leafy in ((0, 1, 0), (1, 1, 0), (37, 2, 0), (5000, 30, 0), (4096, 5, 0), (4097, 80, 0), (20000, 3, 0), (70000, 120, 0)):
        p = np.zeros((n, 4), np.float32)
        p[:, 0] = rng.uniform(-span, span, n); p[:, 1] = rng.uniform(-2, 4, n); p[:, 2] = rng.uniform(-span, span * 0.7, n)
        p[:, 3] = rng.integers(0, 16, n) + rng.random(n).astype(np.float32) * 0.1
        clouds.append(p)
    return clouds


@pytest.mark.parametrize("leaf", [0.2, 0.4, 1.0])
def test_device_voxel_grid_equals_oracle(leaf):
    """Batched device VoxelGrid (hand-written segmented radix sort + run reduction) vs the oracle's restatement of
    pcl::VoxelGrid: same voxel membership, same output order, and -- members being summed in input order on both
    sides with the same float32 operations -- the same bits."""
    import refpy
    clouds = _vg_cases()
    got = capi.selftest_voxel_grid(clouds, leaf)
    for c, g in zip(clouds, got):
        want = refpy.voxel_grid(c, leaf)
        assert g.shape == want.shape, (len(c), g.shape, want.shape)
        assert np.array_equal(g.view(np.uint32), want.view(np.uint32)), "n=%d leaf=%g" % (len(c), leaf)


def test_device_sincosf_is_glibc_sincosf():
    """pm::sinf_glibc / cosf_glibc (csrc/parity_math.cuh) against the host C library, bit for bit: glibc evaluates both in
    double (range reduction + polynomial, FMA build on x86-64), which the device repeats operation by operation."""
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library("m"))
    libm.sinf.restype = ctypes.c_float; libm.sinf.argtypes = [ctypes.c_float]
    libm.cosf.restype = ctypes.c_float; libm.cosf.argtypes = [ctypes.c_float]
    rng = np.random.default_rng(17)
    x = np.concatenate([rng.uniform(-4, 4, 150000), rng.uniform(-119, 119, 100000), rng.uniform(-0.01, 0.01, 50000),
                        rng.uniform(-0.8, 0.8, 100000), [0.0, -0.0, 1e-30, 0.78539816, 0.78539822, 1.5707964, 3.1415927, -3.1415927]]).astype(np.float32)
    s, c = capi.selftest_sincosf(x)
    ws = np.array([libm.sinf(float(v)) for v in x], np.float32)
    wc = np.array([libm.cosf(float(v)) for v in x], np.float32)
    bad_s = int(np.count_nonzero(s.view(np.uint32) != ws.view(np.uint32)))
    bad_c = int(np.count_nonzero(c.view(np.uint32) != wc.view(np.uint32)))
    # the host's libm may be the non-FMA build on a CPU without FMA: that build differs in ~2e-7 of the inputs
    assert bad_s <= 1 and bad_c <= 1, "sinf mismatches %d, cosf mismatches %d of %d" % (bad_s, bad_c, x.size)
