"""GPU parity: mapOptimization scan-to-map (local-map assembly, VoxelGrid filters, voxel-hash 5-NN, line / plane
residuals, 6x6 LM, key-frame bookkeeping) vs the reference's own code (oracle/_ref) on identical synthetic scans.
Bar (BASELINE.json north_star): per-scan 6-DoF poses within 1e-4 m / 1e-4 rad after LM convergence; the map clouds
are compared point for point (same voxel membership, same order) with coordinates within 1e-5 relative."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu

POSE_TOL = 1e-4      # metres / radians (north_star)


def _feed(pl, ref, seq, k, stages=7):
    xyzi, ring = seq.scan(k)
    rf = ref.feed(xyzi, ring, seq.stamp(k), stages=stages)
    x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
    r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
    pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=stages)
    poses, flags = pl.poses()
    return rf, poses[0], int(flags[0])


def _cloud_close(g, r, what):
    assert g.shape == r.shape, "%s: %s vs %s points" % (what, g.shape, r.shape)
    assert np.allclose(g, r, rtol=1e-5, atol=2e-4), "%s: max abs diff %g" % (what, np.abs(g - r).max())


@pytest.mark.parametrize("sensor,scene,kw,nscans", [
    ("vlp16", "street", {}, 42),
    ("vlp16", "room", {"dropout": 0.02}, 26),
    ("hdl32", "street", {}, 18),
    ("hdl64", "street", {}, 14),
    ("vls128", "street", {}, 10),
    ("os1_16", "street", {}, 22),
    ("os1_64", "street", {}, 14),
])
def test_scan_to_map_matches_reference(sensor, scene, kw, nscans):
    import refpy
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=9, scene=scene, speed=4.0, **kw)
    ref = refpy.RefPipeline(sensor)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    mapped = 0
    for k in range(nscans):
        rf, poses, flags = _feed(pl, ref, seq, k)
        tag = "%s scan %d: " % (sensor, k)
        assert bool(flags & 8) == bool(rf & 8), tag + "mapping gate: device %d reference %d" % (flags, rf)
        ts = ref.get("transformSum")
        assert np.max(np.abs(poses[:6] - ts)) < POSE_TOL, tag + "transformSum"
        if rf & 8:
            mapped += 1
            assert np.max(np.abs(pl.download("moTransformSum") - ref.get("moTransformSum"))) < POSE_TOL, tag + "odometry message round trip"
            for name in ("laserCloudCornerLastDS", "laserCloudSurfLastDS", "laserCloudOutlierLastDS", "laserCloudSurfTotalLastDS",
                         "laserCloudCornerFromMapDS", "laserCloudSurfFromMapDS"):
                _cloud_close(pl.download(name), ref.cloud(name), tag + name)
            assert int(pl.download("moState")[1]) == int(ref.get("numKeyFrames")[0]), tag + "number of key frames"
            assert int(pl.download("moState")[0]) == int(ref.get("moDegenerate")[0]), tag + "isDegenerate"
        for name, sl in (("transformAftMapped", slice(6, 12)), ("transformBefMapped", slice(12, 18))):
            g, r = poses[sl], ref.get(name)
            assert np.max(np.abs(g - r)) < POSE_TOL, tag + "%s %s vs %s" % (name, g, r)
        g, r = pl.download("transformTobeMapped"), ref.get("transformTobeMapped")
        assert np.max(np.abs(g - r)) < POSE_TOL, tag + "transformTobeMapped %s vs %s" % (g, r)
    assert mapped >= 3, "mapping ran only %d times" % mapped
    assert np.abs(ref.get("transformAftMapped")).max() > 0.05, "trajectory did not move: the mapping check would be vacuous"
    pl.close(); ref.close()


def test_mapping_lanes_independent():
    """3 lanes with different sequences in lock step give the poses of 3 single-lane runs."""
    sensor = "vlp16"
    seqs = [SynthSequence(sensor, seed=20 + s, scene="street", speed=3.0 + s) for s in range(3)]
    cap = seqs[0].cap
    pl3 = Pipeline(make_params(sensor), lanes=3, max_points=cap)
    singles = [Pipeline(make_params(sensor), lanes=1, max_points=cap) for _ in range(3)]
    for k in range(14):
        x = np.zeros((3, cap, 4), np.float32); r = np.zeros((3, cap), np.uint16); n = np.zeros(3, np.int32)
        for i, s in enumerate(seqs):
            a, b = s.scan(k)
            x[i, :len(a)] = a; r[i, :len(b)] = b; n[i] = len(a)
        pl3.step(x, r, n, np.full(3, k * 0.1), stage_mask=7)
        p3, f3 = pl3.poses()
        for i in range(3):
            singles[i].step(x[i:i + 1], r[i:i + 1], n[i:i + 1], [k * 0.1], stage_mask=7)
            p1, f1 = singles[i].poses()
            assert np.array_equal(p3[i], p1[0]) and f3[i] == f1[0], (k, i, p3[i], p1[0])
    assert np.abs(p3[:, 6:12]).max() > 0.05
    pl3.close()
    for s in singles:
        s.close()


def test_long_free_running_trajectory_stays_within_bar():
    """300 scans, both sides integrating their own state (nothing teacher-forced): ImageProjection outputs and feature
    clouds bit-exact throughout, odometry and mapping poses within the north_star bar.  (With last-ulp differences in the
    pose trigonometry this comparison used to leave the bar after ~230 scans -- one point moving to the neighbouring voxel
    is enough to send the two LM runs down different paths; since every sin / cos / asin / atan2 repeats glibc's
    evaluation, csrc/parity_math.cuh, the two trajectories coincide.)"""
    import os
    import sys
    require_ref("vlp16")
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import long_parity
    worst, first_bit_diff = long_parity.run("vlp16", 300, 11, verbose=False)
    assert first_bit_diff is None, "bit-exact output differs first at %s" % (first_bit_diff,)
    assert worst["sum"] < POSE_TOL and worst["aft"] < POSE_TOL, worst


def test_neighbour_cache_changes_nothing(monkeypatch):
    """The scan-to-map 5-NN of LM iterations >= 1 is answered from a per-query candidate cache when that is provably exact
    (csrc/voxel_hash.cuh).  With LEGO_KNN_CACHE=0 every iteration searches the voxel hash instead: poses, iteration
    counts and row counts of the two runs must be bit-identical."""
    sensor = "hdl32"
    seq = SynthSequence(sensor, seed=17, scene="street", speed=6.0, dropout=0.01)
    monkeypatch.setenv("LEGO_KNN_CACHE", "0")
    plain = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    monkeypatch.setenv("LEGO_KNN_CACHE", "1")
    cached = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    mapped = iters = 0
    for k in range(34):
        xyzi, ring = seq.scan(k)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        out = []
        for pl in (plain, cached):
            pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=7)
            poses, flags = pl.poses()
            out.append((poses[0].copy(), int(flags[0]), pl.download("moState").copy(), pl.download("transformTobeMapped").copy()))
        assert out[0][1] == out[1][1]
        assert np.array_equal(out[0][0].view(np.uint32), out[1][0].view(np.uint32)), "scan %d: poses differ" % k
        assert np.array_equal(out[0][2], out[1][2]), "scan %d: moState %s vs %s" % (k, out[0][2], out[1][2])
        assert np.array_equal(out[0][3].view(np.uint32), out[1][3].view(np.uint32))
        if out[0][1] & 8:
            mapped += 1; iters += int(out[0][2][3])
    assert mapped >= 6 and iters >= 2 * mapped, "too few LM iterations (%d in %d calls) for the cache to matter" % (iters, mapped)
    plain.close(); cached.close()


def test_persistent_lm_kernel_changes_nothing(monkeypatch):
    """scan2MapOptimization runs as one persistent block per lane (csrc/map_optimization.cu: mo_lm_persistent_kernel); with
    LEGO_LM_PERSISTENT=0 the same device functions run as ten launches of {re-rank, search, fit, solve} over all lanes.
    Poses, iteration counts, row counts and the neighbour results must be bit-identical, with lanes of different scan
    sizes (dropout) and different iteration counts in one step."""
    sensor = "hdl32"
    seqs = [SynthSequence(sensor, seed=17, scene="street", speed=6.0, dropout=0.01), SynthSequence(sensor, seed=18, scene="street", speed=2.0),
            SynthSequence(sensor, seed=19, scene="street", speed=9.0, dropout=0.2)]
    cap = seqs[0].cap
    monkeypatch.setenv("LEGO_LM_PERSISTENT", "0")
    loop = Pipeline(make_params(sensor), lanes=3, max_points=cap)
    monkeypatch.setenv("LEGO_LM_PERSISTENT", "1")
    pers = Pipeline(make_params(sensor), lanes=3, max_points=cap)
    mapped = 0; iters = set()
    for k in range(30):
        x = np.zeros((3, cap, 4), np.float32); r = np.zeros((3, cap), np.uint16); n = []
        for l, q in enumerate(seqs):
            xyzi, ring = q.scan(k)
            x[l, :len(xyzi)] = xyzi; r[l, :len(ring)] = ring; n.append(len(xyzi))
        out = []
        for pl in (loop, pers):
            pl.step(x, r, n, [seqs[0].stamp(k)] * 3, stage_mask=7)
            poses, flags = pl.poses()
            out.append((poses.copy(), flags.copy(), [pl.download("moState", l).copy() for l in range(3)],
                        [pl.download("transformTobeMapped", l).copy() for l in range(3)]))
        assert np.array_equal(out[0][1], out[1][1])
        assert np.array_equal(out[0][0].view(np.uint32), out[1][0].view(np.uint32)), "scan %d: poses differ" % k
        for l in range(3):
            assert np.array_equal(out[0][2][l], out[1][2][l]), "scan %d lane %d: moState %s vs %s" % (k, l, out[0][2][l], out[1][2][l])
            assert np.array_equal(out[0][3][l].view(np.uint32), out[1][3][l].view(np.uint32))
            if out[0][1][l] & 8:
                mapped += 1; iters.add(int(out[0][2][l][3]))
    assert mapped >= 15 and len(iters) >= 2, "want mapping calls with different iteration counts, got %d calls, iterations %s" % (mapped, sorted(iters))
    loop.close(); pers.close()


def test_incremental_local_map_changes_nothing(monkeypatch):
    """The local map of extractSurroundingKeyFrames is kept sorted on the device and updated per call (key frames that left
    are deleted, the new ones merged in: csrc/incremental_map.cu).  With LEGO_INCR_MAP=0 every call assembles and filters
    the whole map like the reference (mapOptmization.cpp:1041-1063).  Both paths must give the same down-sampled maps
    and poses bit for bit, long enough for key frames to LEAVE the 50 m neighbourhood (10 m/s: from scan ~50 on)."""
    sensor = "vlp16"
    seqs = [SynthSequence(sensor, seed=61, scene="street", speed=10.0), SynthSequence(sensor, seed=62, scene="street", speed=7.0, dropout=0.01)]
    cap = seqs[0].cap
    monkeypatch.setenv("LEGO_INCR_MAP", "0")
    full = Pipeline(make_params(sensor), lanes=2, max_points=cap)
    monkeypatch.setenv("LEGO_INCR_MAP", "1")
    incr = Pipeline(make_params(sensor), lanes=2, max_points=cap)
    mapped = 0
    max_kf, departed_seen = 0, False
    for k in range(200):
        x = np.zeros((2, cap, 4), np.float32); r = np.zeros((2, cap), np.uint16); n = np.zeros(2, np.int32)
        for l, q in enumerate(seqs):
            xyzi, ring = q.scan(k)
            n[l] = len(xyzi); x[l, :n[l]] = xyzi; r[l, :n[l]] = ring
        res = []
        for pl in (full, incr):
            pl.step(x, r, n, [k * 0.1, k * 0.1], stage_mask=7)
            res.append(pl.poses())
        assert np.array_equal(res[0][1], res[1][1])
        assert np.array_equal(res[0][0].view(np.uint32), res[1][0].view(np.uint32)), "scan %d: poses differ" % k
        if res[0][1][0] & 8:
            mapped += 1
            for lane in range(2):
                for name in ("laserCloudCornerFromMapDS", "laserCloudSurfFromMapDS"):
                    a, b = full.download(name, lane), incr.download(name, lane)
                    assert a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32)), "scan %d lane %d %s" % (k, lane, name)
                assert np.array_equal(full.download("moState", lane), incr.download("moState", lane))
            in_map = len(incr.download("surroundingExistingKeyPosesID", 0))
            saved = int(incr.download("moState", 0)[1])
            max_kf = max(max_kf, in_map)
            departed_seen |= saved > in_map + 2
    assert mapped >= 45 and departed_seen and max_kf >= 8, (mapped, departed_seen, max_kf)
    assert int(incr.download("moState", 0)[5]) == 0
    full.close(); incr.close()
