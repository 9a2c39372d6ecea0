"""GPU parity: the map products outside the per-scan path -- publishGlobalMap's cloud and the three
clouds visualizeGlobalMapThread saves at shutdown (mapOptmization.cpp:724-800) -- against the reference's own code
(oracle/_ref) after both sides ran the same scans.  Key-pose selection (radius search + VoxelGrid over the poses) is
host code like the reference's; key-frame assembly and the VoxelGrid filters run on the device."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200 import capi
from lego_loam_b200.capi import MAP_CORNER, MAP_GLOBAL, MAP_SURFACE, MAP_TRAJECTORY, Pipeline
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu

NAMES = ((MAP_GLOBAL, "pcd:finalCloud"), (MAP_CORNER, "pcd:cornerMap"), (MAP_SURFACE, "pcd:surfaceMap"), (MAP_TRAJECTORY, "pcd:trajectory"))


def _run(pl, ref, seq, nscans, lane=0, lanes=1):
    for k in range(nscans):
        xyzi, ring = seq.scan(k)
        ref.feed(xyzi, ring, seq.stamp(k), stages=7)
        x = np.zeros((lanes, seq.cap, 4), np.float32); x[lane, :len(xyzi)] = xyzi
        r = np.zeros((lanes, seq.cap), np.uint16); r[lane, :len(ring)] = ring
        n = np.zeros(lanes, np.int32); n[lane] = len(xyzi)
        pl.step(x, r, n, np.full(lanes, seq.stamp(k)), stage_mask=7)


def _same_cloud(g, w, what):
    assert g.shape == w.shape, "%s: %s vs %s points" % (what, g.shape, w.shape)
    assert len(w) > 0, what + ": empty cloud, the check would be vacuous"
    assert np.allclose(g, w, rtol=1e-5, atol=2e-4), "%s: max abs diff %g" % (what, np.abs(g - w).max())


@pytest.mark.parametrize("sensor,scene,nscans", [("vlp16", "street", 62), ("hdl32", "room", 30)])
def test_global_map_and_saved_clouds_match_reference(sensor, scene, nscans, tmp_path):
    import refpy
    require_ref(sensor)
    seq = SynthSequence(sensor, seed=13, scene=scene, speed=5.0)
    ref = refpy.RefPipeline(sensor)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    _run(pl, ref, seq, nscans)
    assert ref.global_map() == 4, "the reference saves four PCD files"
    assert len(ref.cloud("pcd:trajectory")) >= 5, "too few key frames"
    _same_cloud(pl.global_map(MAP_GLOBAL), ref.cloud("globalMapKeyFramesDS"), "globalMapKeyFramesDS")
    for which, name in NAMES:
        g = pl.global_map(which)
        _same_cloud(g, ref.cloud(name), name)
        # the file the product writes for ITS cloud against the text the reference side wrote for the reference's
        path = tmp_path / (name[4:] + ".pcd")
        capi.write_pcd_ascii(str(path), g)
        text = path.read_bytes()
        assert text.split(b"DATA ascii\n")[0] == ref.pcd_text(name[4:]).split(b"DATA ascii\n")[0], name + ": PCD header"
        if np.array_equal(g.view(np.uint32), ref.cloud(name).view(np.uint32)):
            assert text == ref.pcd_text(name[4:]), name + ": PCD bytes"
    # the trajectory is host state: exact
    assert np.array_equal(pl.global_map(MAP_TRAJECTORY), ref.cloud("pcd:trajectory"))
    pl.close(); ref.close()


def test_global_map_empty_before_first_key_frame_and_per_lane():
    """No key frame yet -> empty clouds (publishGlobalMap returns early, mapOptmization.cpp:762-763); with two lanes only
    the lane that was fed has map points, and its map equals the single-lane run's."""
    sensor = "vlp16"
    seq = SynthSequence(sensor, seed=5, scene="street", speed=4.0)
    pl2 = Pipeline(make_params(sensor), lanes=2, max_points=seq.cap)
    for which in (MAP_GLOBAL, MAP_CORNER, MAP_SURFACE, MAP_TRAJECTORY):
        assert pl2.global_map(which, lane=1).shape == (0, 4)
    pl1 = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)

    class _NoRef:
        def feed(self, *a, **k):
            return 0
    _run(pl2, _NoRef(), seq, 22, lane=1, lanes=2)
    _run(pl1, _NoRef(), seq, 22)
    for which in (MAP_GLOBAL, MAP_CORNER, MAP_SURFACE, MAP_TRAJECTORY):
        a, b = pl2.global_map(which, lane=1), pl1.global_map(which, lane=0)
        assert len(b) > 0 and np.array_equal(a.view(np.uint32), b.view(np.uint32)), which
        if which != MAP_TRAJECTORY:   # lane 0 only saw empty scans: it may hold key poses, but no points
            assert pl2.global_map(which, lane=0).shape == (0, 4)
    with pytest.raises(capi.LegoError):
        pl1.global_map(7)
    with pytest.raises(capi.LegoError):
        pl1.global_map(MAP_GLOBAL, lane=3)
    pl1.close(); pl2.close()
