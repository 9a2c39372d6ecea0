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


def test_interleaved_ring_labels_at_high_yaw_rate():
    """featureAssociation.cpp:522-523 stores ring + 0.1 * relTime; on a tight circle (0.5 rad/s yaw) relTime is often slightly
    negative, so int(intensity) -- the "ring" of the adjacent-ring scans, featureAssociation.cpp:1062-1110, 1173-1226 -- labels
    a ring-r point r - 1 and the labels interleave inside a ring's stretch.  The odometry must follow the reference's linear
    scans there (break at the first label beyond +-2.5) and must be deterministic: a version that assumed sorted labels had a
    write race in its offset table and gave two different transformCur from identical inputs (scan 326 of this sequence)."""
    import refpy
    sensor = "vlp16"
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=1, scene="room", speed=4.0, n_scans=340, circle_radius=8.0)
    ref = refpy.RefPipeline(sensor)
    pls = [Pipeline(make_params(sensor), lanes=1, max_points=seq.cap) for _ in range(3)]
    interleaved = 0
    worst = 0.0
    for k in range(334):
        xyzi, ring = seq.scan(k)
        ref.feed(xyzi, ring, seq.stamp(k), stages=3)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        for pl in pls:
            pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=3)
        tc = [pl.download("transformCur") for pl in pls]
        assert all(np.array_equal(_bits(t), _bits(tc[0])) for t in tc), "scan %d: pipelines fed the same scans disagree: %s" % (k, tc)
        labels = ref.cloud("laserCloudSurfLast")[:, 3].astype(np.int32)
        interleaved += int((np.diff(labels) < 0).any())
        if k % 4 == 0 or k > 300:
            for name in ("transformCur", "transformSum"):
                d = float(np.max(np.abs(pls[0].download(name) - ref.get(name))))
                worst = max(worst, d)
                assert d < POSE_TOL, "scan %d %s: %g" % (k, name, d)
    assert interleaved >= 1, "the sequence did not produce interleaved ring labels (scan 325 does on the reference)"
    print("interleaved-label scans checked:", interleaved, "worst pose difference", worst)
    for pl in pls:
        pl.close()
    ref.close()


def test_odometry_cluster_sizes_change_nothing(monkeypatch):
    """The scan-to-scan LM kernel runs a lane on a thread-block cluster of 1, 2, 4 or 8 CTAs (csrc/odometry.cu:
    fa_odometry_cluster_kernel) that share the correspondence searches; transformCur, the iteration counts and the
    correspondences' effect on every later scan must not depend on the cluster size.  Two lanes with different feature counts."""
    sensor = "hdl32"
    seqs = [SynthSequence(sensor, seed=23, scene="street", speed=7.0), SynthSequence(sensor, seed=24, scene="street", speed=3.0, dropout=0.15)]
    cap = seqs[0].cap
    pls = []
    for cl in (1, 2, 4, 8, 16):
        monkeypatch.setenv("LEGO_ODO_CLUSTER", str(cl))
        pls.append(Pipeline(make_params(sensor), lanes=2, max_points=cap))
    monkeypatch.delenv("LEGO_ODO_CLUSTER")
    pls.append(Pipeline(make_params(sensor), lanes=2, max_points=cap))       # the default for two lanes (16)
    moved = 0.0
    for k in range(16):
        x = np.zeros((2, cap, 4), np.float32); r = np.zeros((2, cap), np.uint16); n = []
        for l, q in enumerate(seqs):
            xyzi, ring = q.scan(k)
            x[l, :len(xyzi)] = xyzi; r[l, :len(ring)] = ring; n.append(len(xyzi))
        outs = []
        for pl in pls:
            pl.step(x, r, n, [k * 0.1] * 2, stage_mask=3)
            outs.append([(pl.download("transformCur", l).copy(), pl.download("lmState", l).copy(), pl.download("transformSum", l).copy()) for l in range(2)])
        for o in outs[1:]:
            for l in range(2):
                assert np.array_equal(o[l][0].view(np.uint32), outs[0][l][0].view(np.uint32)), "scan %d lane %d transformCur" % (k, l)
                assert np.array_equal(o[l][1], outs[0][l][1]), "scan %d lane %d lmState %s vs %s" % (k, l, o[l][1], outs[0][l][1])
                assert np.array_equal(o[l][2].view(np.uint32), outs[0][l][2].view(np.uint32))
        moved = max(moved, float(np.abs(outs[0][0][0]).max()))
    assert moved > 0.05
    for pl in pls:
        pl.close()
