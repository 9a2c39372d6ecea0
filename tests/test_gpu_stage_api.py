"""GPU: the three per-stage C-ABI handles (lego_ip_* / lego_fa_* / lego_mo_*), chained through HOST buffers the way the
reference chains its nodes through ROS messages, against the reference's own code and against the fused pipeline."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200 import capi
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu


def test_stage_handles_match_reference_and_pipeline():
    import refpy
    sensor = "vlp16"
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=4, scene="street", speed=4.0)
    p = make_params(sensor)
    ip, fa, mo = capi.ImageProjection(p), capi.FeatureAssociation(p), capi.MapOptimization(p)
    pl = capi.Pipeline(p, lanes=1, max_points=seq.cap)
    ref = refpy.RefPipeline(sensor)
    mapped = 0
    for k in range(14):
        xyzi, ring = seq.scan(k)
        t = seq.stamp(k)
        rf = ref.feed(xyzi, ring, t, stages=7)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [t], stage_mask=7)
        poses, _ = pl.poses()
        # --- stage by stage through host memory
        info, seg, outl = ip.process(xyzi, ring, t)
        assert np.array_equal(seg.numpy().view(np.uint32), ref.cloud("segmentedCloud").view(np.uint32)), "segmentedCloud scan %d" % k
        assert np.array_equal(outl.numpy().view(np.uint32), ref.cloud("outlierCloud").view(np.uint32)), "outlierCloud scan %d" % k
        assert info.count == len(ref.cloud("segmentedCloud"))
        seg.dev = None; outl.dev = None   # force the host-memory path of lego_fa_process
        ts, out = fa.process(info, seg, outl, t)
        assert np.array_equal(out.cornerPointsSharp.numpy().view(np.uint32), ref.cloud("cornerPointsSharp").view(np.uint32))
        assert np.array_equal(out.surfPointsFlat.numpy().view(np.uint32), ref.cloud("surfPointsFlat").view(np.uint32))
        assert np.max(np.abs(ts - ref.get("transformSum"))) < 1e-4
        assert np.array_equal(ts, poses[0, :6]), "stage handle and pipeline disagree on transformSum"
        # the init scan forwards corner + surf only (FA:1605-1637); `published` means all three clouds + odometry went out
        assert bool(out.published) == bool((rf & 4) and (rf & 2))
        if out.published:
            for v in (out.laserCloudCornerLast, out.laserCloudSurfLast, out.outlierCloudLast):
                v.dev = None
            aft, bef, done = mo.process(out.laserCloudCornerLast, out.laserCloudSurfLast, out.outlierCloudLast, ts, t)
            assert done == bool(rf & 8)
            mapped += int(done)
            assert np.max(np.abs(aft - ref.get("transformAftMapped"))) < 1e-4, (k, aft, ref.get("transformAftMapped"))
            assert np.max(np.abs(bef - ref.get("transformBefMapped"))) < 1e-4
            assert np.array_equal(aft, poses[0, 6:12]), "stage handle and pipeline disagree on transformAftMapped"
    assert mapped >= 3
    # lego_mo_global_map: the node's map products (mapOptmization.cpp:724-800) from the stage handle
    ref.global_map()
    for which, name in ((capi.MAP_GLOBAL, "globalMapKeyFramesDS"), (capi.MAP_CORNER, "pcd:cornerMap"), (capi.MAP_SURFACE, "pcd:surfaceMap"),
                        (capi.MAP_TRAJECTORY, "pcd:trajectory")):
        g, w = mo.global_map(which), ref.cloud(name)
        assert g.shape == w.shape and len(w) > 0, "%s: %s vs %s points" % (name, g.shape, w.shape)
        assert np.allclose(g, w, rtol=1e-5, atol=2e-4), "%s: max abs diff %g" % (name, np.abs(g - w).max())
        assert np.array_equal(g.view(np.uint32), pl.global_map(which).view(np.uint32)), "stage handle and pipeline disagree on " + name
    for h in (ip, fa, mo, pl):
        h.close()


def test_image_projection_error_codes():
    p = make_params("vlp16")
    ip = capi.ImageProjection(p)
    xyzi = np.zeros((10, 4), np.float32); xyzi[:, 0] = 5.0
    ring = np.zeros(10, np.uint16)
    with pytest.raises(capi.LegoError, match=r"\(-3\)"):       # LEGO_ERR_NOT_DENSE, imageProjection.cpp:172-177
        ip.process(xyzi, ring, 0.0, is_dense=False)
    big = np.zeros((p.N_SCAN * p.Horizon_SCAN * 2, 4), np.float32)
    with pytest.raises(capi.LegoError, match=r"\(-4\)"):       # LEGO_ERR_CAPACITY
        ip.process(big, np.zeros(len(big), np.uint16), 0.0)
    info, seg, outl = ip.process(np.zeros((0, 4), np.float32), np.zeros(0, np.uint16), 0.0)   # empty scan
    assert seg.count == 0 and outl.count == 0 and info.count == 0
    ip.close()
    # without the ring channel a non-dense cloud is filtered (removeNaNFromPointCloud, imageProjection.cpp:168)
    pn = make_params("vlp16_noring")
    ipn = capi.ImageProjection(pn)
    seq = SynthSequence("vlp16_noring", seed=2, scene="room")
    xyzi, _ = seq.scan(0)
    a = ipn.process(xyzi, None, 0.0)[1].numpy()
    holes = np.insert(xyzi, [100, 5000, 5000], np.nan, axis=0)
    b = ipn.process(holes, None, 0.1, is_dense=False)[1].numpy()
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    ipn.close()


def test_streaming_pose_readback_equals_blocking():
    """lego_pipeline_poses_async / _wait (scan k+1 submitted before the result of scan k is collected) returns what
    lego_pipeline_get_poses returns for the same scans."""
    sensor = "vlp16"
    p = make_params(sensor)
    seqs = [SynthSequence(sensor, seed=s, scene="street", speed=3.0) for s in (2, 9)]
    cap = max(s.cap for s in seqs)
    a = capi.Pipeline(p, lanes=2, max_points=cap)
    b = capi.Pipeline(p, lanes=2, max_points=cap)
    n = 12

    def scan(k):
        x = np.zeros((2, cap, 4), np.float32); r = np.zeros((2, cap), np.uint16); npts = []
        for l, s in enumerate(seqs):
            xyzi, ring = s.scan(k)
            x[l, :len(xyzi)] = xyzi; r[l, :len(ring)] = ring; npts.append(len(xyzi))
        return x, r, npts, [s.stamp(k) for s in seqs]

    want = []
    for k in range(n):
        x, r, npts, st = scan(k)
        a.step(x, r, npts, st, stage_mask=7)
        want.append(a.poses())
    got = []
    held = []   # host arrays stay alive until their upload has been consumed
    for k in range(n):
        x, r, npts, st = scan(k)
        held.append((x, r))
        b.step(x, r, npts, st, stage_mask=7)
        b.poses_async(k & 1)
        if k > 0:
            got.append(b.poses_wait((k - 1) & 1))
    got.append(b.poses_wait((n - 1) & 1))
    for k in range(n):
        assert np.array_equal(got[k][0], want[k][0]), "poses differ at scan %d" % k
        assert np.array_equal(got[k][1], want[k][1]), "flags differ at scan %d" % k
    a.close(); b.close()


def test_transform_fusion_matches_reference():
    """Stage bit 3 of the pipeline and the lego_tf_* handle against the reference's transformFusion.cpp (compiled verbatim
    into the oracle): fused pose transformMapped and the published /integrated_to_init pose within 1e-4 m / rad of the
    reference, and pipeline == stage handle exactly (same kernels, poses passed through host memory)."""
    import refpy
    sensor = "vlp16"
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=12, scene="street", speed=5.0)
    p = make_params(sensor)
    pl = capi.Pipeline(p, lanes=1, max_points=seq.cap)
    tf = capi.TransformFusion(p)
    ref = refpy.RefPipeline(sensor)
    fused = mapped_updates = 0
    for k in range(30):
        xyzi, ring = seq.scan(k)
        t = seq.stamp(k)
        rf = ref.feed(xyzi, ring, t, stages=15)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [t], stage_mask=15)
        poses, flags = pl.poses()
        assert bool(pl.download("tfPublished")[0]) == bool(rf & 16), "publish flag differs at scan %d" % k
        if rf & 16:
            fused += 1
            want_m, want_i = ref.get("transformMapped"), ref.get("integratedToInit")
            got_m, got_i = pl.download("transformMapped"), pl.download("integratedToInit")
            assert np.max(np.abs(got_m - want_m)) < 1e-4, "transformMapped scan %d: %s vs %s" % (k, got_m, want_m)
            assert np.max(np.abs(got_i - want_i)) < 1e-4, "integratedToInit scan %d: %s vs %s" % (k, got_i, want_i)
            m2, i2 = tf.laser_odometry(poses[0, :6])
            assert np.array_equal(m2, got_m) and np.array_equal(i2, got_i), "stage handle differs from the pipeline at scan %d" % k
        if flags[0] & 8:   # the mapping body ran: /aft_mapped_to_init was published
            mapped_updates += 1
            tf.aft_mapped(poses[0, 6:12], poses[0, 12:18])
    assert fused >= 25 and mapped_updates >= 5
    pl.close(); tf.close()
