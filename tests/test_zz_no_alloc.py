"""GPU test of lego_voxel_grid_noalloc (include/lego_b200.h): the allocation-free VoxelGrid entry point, whose sort scratch
is a static __device__ array, against the existing allocating entry point (lego_selftest_voxel_grid) on the same clouds,
two calls in a row (the second finds the first call's contents in the shared scratch).  Named test_zz_* so that pytest
collects it last: a failure here must not hide the parity suite."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from lego_loam_b200.capi import load, selftest_voxel_grid, voxel_grid_noalloc

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 2e-4     # the coordinate tolerance of the parity suite (tests/test_gpu_map_optimization.py::_cloud_close)


def _cloud(rng, n, extent):
    c = np.empty((n, 4), np.float32)
    c[:, :3] = rng.uniform(-extent, extent, (n, 3)) * np.array([1.0, 1.0, 0.1])
    c[:, 3] = rng.integers(0, 16, n) + rng.uniform(0, 0.1, n)
    return c


def _run_noalloc(cloud, leaf):
    import torch
    d_in = torch.from_numpy(cloud).cuda()
    d_out = torch.full((max(len(cloud), 1), 4), -7.0, dtype=torch.float32, device="cuda")
    d_n = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()                      # the entry point enqueues on the default stream; torch's work is done
    rc = voxel_grid_noalloc(d_in.data_ptr(), len(cloud), leaf, d_out.data_ptr(), d_out.shape[0], d_n.data_ptr())
    assert rc == 0, load().lego_last_error().decode()
    torch.cuda.synchronize()
    m = int(d_n.item())
    return d_out[:m].cpu().numpy()


def test_noalloc_voxel_grid_matches_allocating_entry_point_twice():
    rng = np.random.default_rng(5)
    first = _cloud(rng, 60000, 40.0)              # 15 tiles of the radix sort
    second = _cloud(rng, 9000, 12.0)              # smaller: stale keys / histograms of the first call lie beyond it
    for cloud, leaf in ((first, 0.4), (second, 0.2), (first, 0.4)):
        want = selftest_voxel_grid([cloud], leaf)[0]
        got = _run_noalloc(cloud, leaf)
        assert 0 < len(want) < len(cloud)
        assert got.shape == want.shape, "%s vs %s voxels" % (got.shape, want.shape)
        assert np.allclose(got, want, rtol=RTOL, atol=ATOL), "max abs diff %g" % np.abs(got - want).max()


def test_noalloc_voxel_grid_limits():
    import torch
    lib = load()
    assert lib.lego_voxel_grid_noalloc_max_points() == 262144
    buf = torch.zeros((16, 4), dtype=torch.float32, device="cuda")
    n_out = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = voxel_grid_noalloc(buf.data_ptr(), 262145, 0.4, buf.data_ptr(), 16, n_out.data_ptr())
    assert rc == -4 and "262144" in lib.lego_last_error().decode()          # LEGO_ERR_CAPACITY, names the limit
    assert voxel_grid_noalloc(buf.data_ptr(), 16, 0.0, buf.data_ptr(), 16, n_out.data_ptr()) == -1      # leaf <= 0
    assert voxel_grid_noalloc(None, 16, 0.4, buf.data_ptr(), 16, n_out.data_ptr()) == -1
    # an empty cloud is a valid call and yields an empty cloud
    n_out.fill_(5)
    torch.cuda.synchronize()
    assert voxel_grid_noalloc(buf.data_ptr(), 0, 0.4, buf.data_ptr(), 16, n_out.data_ptr()) == 0
    torch.cuda.synchronize()
    assert int(n_out.item()) == 0
