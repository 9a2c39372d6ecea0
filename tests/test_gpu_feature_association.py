"""GPU parity: FeatureAssociation kernels (de-skew, curvature, occlusion, sorted feature selection, per-ring
VoxelGrid, scan-to-scan LM odometry) vs the reference's own code (oracle/_ref) on identical synthetic scans.
Bar (BASELINE.json north_star): selected feature indices / labels bit-exact, smoothness and coordinates within
1e-5 relative, poses within 1e-4 m / 1e-4 rad."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu

POSE_TOL = 1e-4      # metres / radians (north_star)
COORD_RTOL = 1e-5    # relative, coordinates and smoothness (north_star)


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _feed(pl, ref, seq, k, stages):
    xyzi, ring = seq.scan(k)
    ref.feed(xyzi, ring, seq.stamp(k), stages=stages)
    x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
    r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
    pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=stages)


def _close(a, b, rtol=COORD_RTOL, atol=1e-6):
    return a.shape == b.shape and np.allclose(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("sensor,scene,kw,nscans", [
    ("vlp16", "street", {}, 12),
    ("vlp16", "room", {"dropout": 0.03}, 8),
    ("vlp16_noring", "street", {}, 4),
    ("hdl32", "street", {}, 6),
    ("hdl64", "street", {}, 5),
    ("vls128", "street", {"dropout": 0.02}, 3),
    ("os1_16", "street", {}, 8),
    ("os1_64", "room", {}, 5),
])
def test_features_bit_exact_and_odometry(sensor, scene, kw, nscans):
    import refpy
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=5, scene=scene, speed=4.0, **kw)
    ref = refpy.RefPipeline(sensor)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    for k in range(nscans):
        _feed(pl, ref, seq, k, 3)
        tag = "%s scan %d: " % (sensor, k)
        # de-skewed cloud: xyz are copies (bit-exact), intensity = ring + scanPeriod * relTime (float ops mirrored)
        adj_g, adj_r = pl.download("segmentedCloudAdj"), ref.cloud("segmentedCloudAdj")
        assert adj_g.shape == adj_r.shape, tag + "segmented size"
        assert np.array_equal(_bits(adj_g), _bits(adj_r)), tag + "adjustDistortion"
        S = len(adj_r)
        cg, cr = pl.download("cloudCurvature"), ref.get("cloudCurvature")
        assert np.array_equal(_bits(cg[5:S - 5]), _bits(cr[5:S - 5])), tag + "cloudCurvature"
        assert np.array_equal(pl.download("cloudNeighborPicked"), ref.get("cloudNeighborPicked")), tag + "cloudNeighborPicked"
        assert np.array_equal(pl.download("cloudLabel"), ref.get("cloudLabel")), tag + "cloudLabel"
        for name in ("cornerPointsSharp", "cornerPointsLessSharp", "surfPointsFlat"):
            g, r = pl.download(name), ref.cloud(name)
            assert g.shape == r.shape and np.array_equal(_bits(g), _bits(r)), tag + name
        g, r = pl.download("surfPointsLessFlat"), ref.cloud("surfPointsLessFlat")
        assert g.shape == r.shape, tag + "surfPointsLessFlat count %s vs %s" % (g.shape, r.shape)
        assert _close(g, r), tag + "surfPointsLessFlat values"
        # odometry
        tc_g, tc_r = pl.download("transformCur"), ref.get("transformCur")
        ts_g, ts_r = pl.download("transformSum"), ref.get("transformSum")
        assert np.max(np.abs(tc_g - tc_r)) < POSE_TOL, tag + "transformCur %s vs %s" % (tc_g, tc_r)
        assert np.max(np.abs(ts_g - ts_r)) < POSE_TOL, tag + "transformSum %s vs %s" % (ts_g, ts_r)
        g, r = pl.download("laserCloudCornerLast"), ref.cloud("laserCloudCornerLast")
        # TransformToEnd of the less-sharp points with the scan's transformCur: bit-identical since parity_math.cuh states glibc's
        # sinf / cosf exactly (north star: coordinates within 1e-5 relative)
        assert g.shape == r.shape and np.array_equal(_bits(g), _bits(r)), tag + "laserCloudCornerLast"
        g, r = pl.download("laserCloudSurfLast"), ref.cloud("laserCloudSurfLast")
        assert g.shape == r.shape and np.allclose(g, r, rtol=1e-5, atol=1e-6), tag + "laserCloudSurfLast"
    assert np.abs(ref.get("transformSum")).max() > 0.05, "trajectory did not move: the odometry check would be vacuous"
    pl.close(); ref.close()
