"""GPU parity of the optional IMU path against the reference's own code (oracle/_ref; the IMU code
is all inside featureAssociation.cpp / mapOptmization.cpp, so this parity is PINNED by the verbatim build):
imuHandler + AccumulateIMUShiftAndRotation (featureAssociation.cpp:392-459), the per-point IMU interpolation and
TransformToStartIMU de-skew of adjustDistortion (525-614), updateInitialGuess (1639-1664), the IMU terms of TransformToEnd
(928-952) and integrateTransformation / PluginIMURotation (955-1013, 1697-1725), checkSystemInitialization (1632-1633) and
the roll / pitch blend of mapOptimization::transformUpdate (mapOptmization.cpp:463-490).
An ideal IMU riding on the synthetic lidar publishes at 100 Hz; both sides get every message with a stamp up to the end
of the sweep before that sweep's scan, as the reference's subscribers would have."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu

POSE_TOL = 1e-4


def _close(g, r, what, rtol=1e-5, atol=2e-4):
    assert g.shape == r.shape, "%s: %s vs %s points" % (what, g.shape, r.shape)
    assert np.allclose(g, r, rtol=rtol, atol=atol), "%s: max abs diff %g" % (what, np.abs(g - r).max())


@pytest.mark.parametrize("sensor,speed,nscans", [("vlp16", 3.0, 42), ("hdl32", 5.0, 14)])
def test_imu_path_matches_reference(sensor, speed, nscans):
    import refpy
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=5, scene="street", speed=speed, yaw_amp=0.25, sway_amp=1.5)
    ref = refpy.RefPipeline(sensor)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    j = 0
    mapped = 0
    prior_used = False
    for k in range(nscans):
        msgs, j = seq.imu_until(j, seq.stamp(k) + 0.1)
        for (t, q, w, a) in msgs:
            ref.imu(t, q, w, a); pl.imu(0, t, q, w, a)
        xyzi, ring = seq.scan(k)
        rf = ref.feed(xyzi, ring, seq.stamp(k), stages=7)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=7)
        poses, flags = pl.poses()
        tag = "%s scan %d: " % (sensor, k)
        # ImageProjection does not see the IMU: labels stay bit-exact
        assert np.array_equal(pl.download("labelMat"), ref.get("labelMat")), tag + "labelMat"
        # de-skewed cloud: same points, coordinates to 1e-5 relative (device sinf/cosf differ from glibc in the last ulp)
        _close(pl.download("segmentedCloudAdj"), ref.cloud("segmentedCloudAdj"), tag + "segmentedCloudAdj")
        # curvature / picks come from the ranges, which the IMU does not touch: still bit-exact
        assert np.array_equal(pl.download("cloudLabel"), ref.get("cloudLabel")), tag + "cloudLabel"
        for name in ("cornerPointsSharp", "cornerPointsLessSharp", "surfPointsFlat", "surfPointsLessFlat"):
            _close(pl.download(name), ref.cloud(name), tag + name)
        st = pl.download("imuState")
        ws, wl = ref.get("imuStart"), ref.get("imuLast")
        assert np.allclose(st[0:6], ws, rtol=1e-5, atol=1e-6), tag + "imu start state %s vs %s" % (st[0:6], ws)
        if k > 0:
            got_last = np.concatenate([st[27:30], st[30:33], st[18:21]])
            assert np.allclose(got_last, wl, rtol=1e-4, atol=1e-5), tag + "imu last / velocity / angular prior %s vs %s" % (got_last, wl)
            prior_used |= bool(np.any(np.abs(wl[6:9]) > 1e-4))
        assert np.max(np.abs(pl.download("transformCur") - ref.get("transformCur"))) < POSE_TOL, tag + "transformCur"
        assert np.max(np.abs(poses[0, :6] - ref.get("transformSum"))) < POSE_TOL, tag + "transformSum %s vs %s" % (poses[0, :6], ref.get("transformSum"))
        assert bool(flags[0] & 8) == bool(rf & 8), tag + "mapping gate"
        if rf & 8:
            mapped += 1
            _close(pl.download("laserCloudSurfLastDS"), ref.cloud("laserCloudSurfLastDS"), tag + "laserCloudSurfLastDS")
        for name, sl in (("transformAftMapped", slice(6, 12)), ("transformBefMapped", slice(12, 18))):
            assert np.max(np.abs(poses[0, sl] - ref.get(name))) < POSE_TOL, tag + "%s %s vs %s" % (name, poses[0, sl], ref.get(name))
    assert mapped >= 3 and prior_used, "mapping ran %d times, rotation prior used: %s" % (mapped, prior_used)
    pl.close(); ref.close()


def test_imu_changes_the_result():
    """Sanity: the same scans with and without IMU messages give different odometry (the path is not a no-op),
    and a lane without IMU next to a lane with IMU keeps its no-IMU result bit for bit."""
    sensor = "vlp16"
    seq = SynthSequence(sensor, seed=5, scene="street", speed=3.0, yaw_amp=0.25, sway_amp=1.5)
    p = make_params(sensor)
    pl2 = Pipeline(p, lanes=2, max_points=seq.cap)      # lane 0 with IMU, lane 1 without
    pl1 = Pipeline(p, lanes=1, max_points=seq.cap)      # without
    j = 0
    for k in range(8):
        msgs, j = seq.imu_until(j, seq.stamp(k) + 0.1)
        for (t, q, w, a) in msgs:
            pl2.imu(0, t, q, w, a)
        xyzi, ring = seq.scan(k)
        x = np.zeros((2, seq.cap, 4), np.float32); x[:, :len(xyzi)] = xyzi
        r = np.zeros((2, seq.cap), np.uint16); r[:, :len(ring)] = ring
        pl2.step(x, r, [len(xyzi)] * 2, [seq.stamp(k)] * 2, stage_mask=7)
        pl1.step(x[:1], r[:1], [len(xyzi)], [seq.stamp(k)], stage_mask=7)
    p2, _ = pl2.poses(); p1, _ = pl1.poses()
    assert np.array_equal(p2[1], p1[0]), "the lane without IMU changed: %s vs %s" % (p2[1], p1[0])
    assert np.max(np.abs(p2[0, :6] - p2[1, :6])) > 1e-6, "IMU messages had no effect at all"
    pl2.close(); pl1.close()
