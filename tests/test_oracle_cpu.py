"""CPU tests of the oracle itself: the third-party restatements in oracle/thirdparty (exposed by oracle/port.cpp)
against independent numpy / scipy formulations.  These libraries (PCL, FLANN, OpenCV) are not vendored in the
reference, so this is what pins the restatements ("parity unpinned by the reference's own tests")."""
import numpy as np
import pytest

import refpy


def _numpy_voxel_grid(p, leaf):
    """pcl::VoxelGrid as published (voxel_grid.hpp::applyFilter): float32 index arithmetic, ascending voxel index,
    float32 mean of the members in input order."""
    if len(p) == 0:
        return p.reshape(0, 4)
    inv = np.float32(1.0) / np.float32(leaf)
    mn, mx = p[:, :3].min(0), p[:, :3].max(0)
    minb = np.floor(mn * inv).astype(np.int64); maxb = np.floor(mx * inv).astype(np.int64)
    div = maxb - minb + 1
    ijk = (np.floor(p[:, :3] * inv) - minb.astype(np.float32)).astype(np.int64)
    idx = ijk[:, 0] + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    order = np.argsort(idx, kind="stable")
    out = []
    start = 0
    sidx = idx[order]
    while start < len(order):
        end = start
        while end < len(order) and sidx[end] == sidx[start]:
            end += 1
        s = np.zeros(4, np.float32)
        for k in order[start:end]:
            s = (s + p[k]).astype(np.float32)
        out.append(s / np.float32(end - start))
        start = end
    return np.array(out, np.float32)


@pytest.mark.parametrize("leaf", [0.2, 0.4, 1.0])
def test_oracle_voxel_grid_matches_numpy(leaf):
    rng = np.random.default_rng(0)
    for n, span in ((0, 1), (1, 1), (50, 2), (3000, 25)):
        p = np.zeros((n, 4), np.float32)
        p[:, :3] = rng.uniform(-span, span, (n, 3)); p[:, 3] = rng.integers(0, 16, n)
        got = refpy.voxel_grid(p, leaf)
        want = _numpy_voxel_grid(p, leaf)
        assert got.shape == want.shape
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_oracle_knn_matches_ckdtree():
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(1)
    pts = np.zeros((5000, 4), np.float32); pts[:, :3] = rng.uniform(-20, 20, (5000, 3))
    q = np.zeros((300, 4), np.float32); q[:, :3] = rng.uniform(-22, 22, (300, 3))
    idx, d2 = refpy.knn(pts, q, 5)
    dd, ii = cKDTree(pts[:, :3].astype(np.float64)).query(q[:, :3].astype(np.float64), k=5)
    assert np.array_equal(idx, ii.astype(np.int32))
    assert np.allclose(d2, dd ** 2, rtol=1e-5)


def test_oracle_small_solvers_match_numpy_and_cv2():
    rng = np.random.default_rng(2)
    lib = refpy.port()
    for _ in range(200):
        A = rng.normal(size=(5, 3)).astype(np.float32) * 10
        b = -np.ones(5, np.float32)
        x = np.zeros(3, np.float32)
        lib.oracle_solve_qr(A.ctypes.data, b.ctypes.data, x.ctypes.data, 5, 3)
        want = np.linalg.lstsq(A.astype(np.float64), b.astype(np.float64), rcond=None)[0]
        assert np.allclose(x, want, rtol=2e-4, atol=2e-6)
        M = rng.normal(size=(6, 6)).astype(np.float32); S = (M @ M.T).astype(np.float32)
        ev = np.zeros(6, np.float32); V = np.zeros((6, 6), np.float32)
        lib.oracle_eigen_sym(S.ctypes.data, ev.ctypes.data, V.ctypes.data, 6)
        w = np.linalg.eigvalsh(S.astype(np.float64))[::-1]
        assert np.allclose(ev, w, rtol=1e-4, atol=1e-4)
        assert np.allclose(V @ S @ V.T, np.diag(ev), atol=2e-3 * max(1.0, float(w[0])))   # rows are eigenvectors
    try:
        import cv2
    except ImportError:
        return
    A = rng.normal(size=(5, 3)).astype(np.float32); b = -np.ones((5, 1), np.float32)
    ok, xc = cv2.solve(A, b, flags=cv2.DECOMP_QR)
    x = np.zeros(3, np.float32)
    lib.oracle_solve_qr(A.ctypes.data, b.ctypes.data, x.ctypes.data, 5, 3)
    assert np.allclose(x, xc[:, 0], rtol=1e-4, atol=1e-6)


def test_oracle_transform_fusion_is_identity_before_the_first_mapping_message():
    """transformFusion.cpp compiled into the oracle: until TransformFusion has seen an /aft_mapped_to_init message its
    transformBefMapped / transformAftMapped are zero, so the fused pose must equal the odometry pose (through the
    quaternion round trip of the message); afterwards it must differ from it by the mapping correction only."""
    import lego_loam_b200  # noqa: F401
    from lego_loam_b200.synth import SynthSequence
    sensor = "vlp16"
    if not refpy.available(sensor):
        pytest.skip("oracle/_ref not built")
    seq = SynthSequence(sensor, seed=3, scene="street", speed=4.0)
    ref = refpy.RefPipeline(sensor)
    seen_mapping = False
    checked = 0
    for k in range(8):
        xyzi, ring = seq.scan(k)
        f = ref.feed(xyzi, ring, seq.stamp(k), stages=15)
        if f & 16 and not seen_mapping:
            # transformFusion.cpp:98-179 with zero Bef/Aft reduces to transformMapped = transformSum up to float rounding
            assert np.max(np.abs(ref.get("transformMapped") - ref.get("transformSum"))) < 2e-6
            checked += 1
        if f & 8:
            seen_mapping = True
        if f & 16:
            q = ref.get("integratedToInit")[:4]
            assert abs(float(np.dot(q, q)) - 1.0) < 1e-5   # a unit quaternion is published
    assert checked >= 1 and seen_mapping
