"""GPU parity: ImageProjection kernels vs the reference's own code (oracle/_ref) on identical synthetic scans.
Bar (BASELINE.json north_star): ground masks, labels, cloud_info indices bit-exact; range / coordinates bit-exact
too, since the kernels reproduce the reference's float operations one for one."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _compare_ip(pl, ref, lane=0):
    errs = []

    def eq(name, a, b):
        if a.shape != b.shape or not np.array_equal(a, b):
            n = int(np.sum(a != b)) if a.shape == b.shape else -1
            errs.append("%s: shape %s vs %s, %d mismatches" % (name, a.shape, b.shape, n))

    eq("rangeMat", _bits(pl.download("rangeMat", lane)), _bits(ref.get("rangeMat")))
    eq("groundMat", pl.download("groundMat", lane).astype(np.int32), ref.get("groundMat"))
    eq("labelMat", pl.download("labelMat", lane), ref.get("labelMat"))
    eq("startRingIndex", pl.download("startRingIndex", lane), ref.get("startRingIndex"))
    eq("endRingIndex", pl.download("endRingIndex", lane), ref.get("endRingIndex"))
    eq("segColInd", pl.download("segColInd", lane), ref.get("segColInd"))
    eq("segGroundFlag", pl.download("segGroundFlag", lane).astype(np.int32), ref.get("segGroundFlag"))
    eq("segRange", _bits(pl.download("segRange", lane)), _bits(ref.get("segRange")))
    eq("segmentedCloud", _bits(pl.download("segmentedCloud", lane)), _bits(ref.cloud("segmentedCloud")))
    eq("outlierCloud", _bits(pl.download("outlierCloud", lane)), _bits(ref.cloud("outlierCloud")))
    eq("orientation", _bits(pl.download("orientation", lane)), _bits(ref.get("orientation")))
    fc_g, fc_r = pl.download("fullCloud", lane), ref.cloud("fullCloud")
    valid = fc_r[:, 3] != -1
    eq("fullCloud.valid", fc_g[:, 3] != -1, valid)
    eq("fullCloud", _bits(fc_g[valid]), _bits(fc_r[valid]))
    return errs


@pytest.mark.parametrize("sensor,scene,kw", [
    ("vlp16", "street", {}),
    ("vlp16", "room", {}),
    ("vlp16", "street", {"dropout": 0.03, "dup_cols": 0.01}),
    ("vlp16_noring", "street", {}),
    ("hdl32", "street", {}),
    ("hdl64", "street", {}),
    ("vls128", "street", {"dropout": 0.02}),
    ("os1_16", "street", {}),                       # utility.h:88-94: 1024-column images
    ("os1_64", "street", {"dropout": 0.02}),        # utility.h:96-102
])
def test_image_projection_bit_exact(sensor, scene, kw):
    import refpy
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=3, scene=scene, speed=3.0, **kw)
    ref = refpy.RefPipeline(sensor)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    for k in (0, 1, 7):
        xyzi, ring = seq.scan(k)
        ref.feed(xyzi, ring, seq.stamp(k), stages=1)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=1)
        errs = _compare_ip(pl, ref)
        assert not errs, "scan %d: %s" % (k, "; ".join(errs))
    pl.close(); ref.close()


def test_image_projection_lanes_independent():
    """4 lanes fed 4 different scans in one launch give the same bytes as 4 single-lane runs."""
    sensor = "vlp16"
    seqs = [SynthSequence(sensor, seed=s, scene="street") for s in range(4)]
    cap = seqs[0].cap
    x = np.zeros((4, cap, 4), np.float32); r = np.zeros((4, cap), np.uint16); n = np.zeros(4, np.int32)
    for i, s in enumerate(seqs):
        a, b = s.scan(2)
        x[i, :len(a)] = a; r[i, :len(b)] = b; n[i] = len(a)
    pl4 = Pipeline(make_params(sensor), lanes=4, max_points=cap)
    pl4.step(x, r, n, np.zeros(4), stage_mask=1)
    for i in range(4):
        pl1 = Pipeline(make_params(sensor), lanes=1, max_points=cap)
        pl1.step(x[i:i + 1], r[i:i + 1], n[i:i + 1], [0.0], stage_mask=1)
        for name in ("labelMat", "segmentedCloud", "segColInd", "startRingIndex", "outlierCloud"):
            assert np.array_equal(pl4.download(name, i).view(np.uint8), pl1.download(name, 0).view(np.uint8)), (i, name)
        pl1.close()
    pl4.close()


@pytest.mark.parametrize("sensor", ["vlp16", "hdl64", "vls128"])
def test_tiled_connected_components_change_nothing(sensor, monkeypatch):
    """labelComponents as shared-memory tiles (csrc/image_projection.cu: ip_ccl_tile / border / finish) with the tan-ratio
    shortcut of the edge test, against the three-pass form on the global forest with atan2f in every edge test
    (LEGO_CCL_TILE=0, LEGO_SEG_EDGE_EXACT=1): roots, labels, segmented and outlier clouds, ring indices bit-identical, on
    scans with dropouts (ragged components) and for tile widths 256 / 64 / 32."""
    from lego_loam_b200.capi import Pipeline
    from lego_loam_b200.params import make_params
    from lego_loam_b200.synth import SynthSequence
    seq = SynthSequence(sensor, seed=5, scene="street", speed=5.0, dropout=0.15)
    monkeypatch.setenv("LEGO_CCL_TILE", "0"); monkeypatch.setenv("LEGO_SEG_EDGE_EXACT", "1")
    old = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    monkeypatch.setenv("LEGO_CCL_TILE", "1"); monkeypatch.setenv("LEGO_SEG_EDGE_EXACT", "0")
    new = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    names = ("labelMat", "segmentedCloud", "outlierCloud", "startRingIndex", "endRingIndex", "segGroundFlag",
             "segColInd", "segRange", "ipCounts")
    segments = 0
    for k in range(4):
        xyzi, ring = seq.scan(k)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        for pl in (old, new):
            pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=1)
        for nm in names:
            a, b = old.download(nm), new.download(nm)
            assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8)), "scan %d: %s differs" % (k, nm)
        lab = old.download("labelMat")
        segments += int(lab[(lab > 0) & (lab < 999999)].max(initial=0))
    assert segments > 20
    old.close(); new.close()
