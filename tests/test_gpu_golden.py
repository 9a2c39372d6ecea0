"""GPU: the CUDA pipeline against the committed golden fixtures (tests/golden/*.json, written from the reference's own
code by tools/make_golden.py) -- no oracle needed at run time.  Labels, segmented cloud and feature clouds bit-exact
(CRC32 of the bytes), poses within the north-star tolerance 1e-4 m / rad."""
import glob
import json
import os
import zlib

import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.json")))
POSE_TOL = 1e-4


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_pipeline_matches_golden(path):
    g = json.load(open(path))
    seq = SynthSequence(g["sensor"], seed=g["seed"], scene=g["scene"], speed=g["speed"], circle_radius=g.get("circle_radius", 0.0))
    pl = Pipeline(make_params(g["sensor"], loop_closure=bool(g.get("loop"))), lanes=1, max_points=seq.cap)
    j = 0
    loops = 0
    for k, e in enumerate(g["scans"]):
        if g.get("imu"):
            msgs, j = seq.imu_until(j, seq.stamp(k) + 0.1)
            for (t, q, w, a) in msgs:
                pl.imu(0, t, q, w, a)
            assert crc(np.concatenate([np.concatenate([[t], q, w, a]) for (t, q, w, a) in msgs]) if msgs else np.zeros(0)) == e["imu_crc"], "IMU generator drifted at scan %d" % k
        xyzi, ring = seq.scan(k)
        assert crc(xyzi) == e["input_crc"], "synthetic generator drifted at scan %d" % k
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=g["stages"])
        poses, flags = pl.poses()
        if "labelMat_crc" in e:
            assert crc(pl.download("labelMat")) == e["labelMat_crc"], "labelMat scan %d" % k
            assert crc(pl.download("segmentedCloud").view(np.uint32)) == e["segmentedCloud_crc"], "segmentedCloud scan %d" % k
            if not g.get("imu"):   # with IMU the feature coordinates carry the de-skew: compared through segmentedCloudAdj below
                assert crc(pl.download("cornerPointsSharp").view(np.uint32)) == e["cornerPointsSharp_crc"], "cornerPointsSharp scan %d" % k
                assert crc(pl.download("cornerPointsLessSharp").view(np.uint32)) == e["cornerPointsLessSharp_crc"], "cornerPointsLessSharp scan %d" % k
                assert crc(pl.download("surfPointsFlat").view(np.uint32)) == e["surfPointsFlat_crc"], "surfPointsFlat scan %d" % k
            assert len(pl.download("surfPointsLessFlat")) == e["surfPointsLessFlat_n"]
        if "segmentedCloudAdj_crc" in e:   # IMU de-skew: bit-exact because sin / cos repeat glibc's evaluation
            assert crc(pl.download("segmentedCloudAdj").view(np.uint32)) == e["segmentedCloudAdj_crc"], "segmentedCloudAdj scan %d" % k
            assert crc(pl.download("cornerPointsSharp").view(np.uint32)) == e["cornerPointsSharp_crc"], "cornerPointsSharp scan %d" % k
        assert np.max(np.abs(poses[0, :6] - np.array(e["transformSum"], np.float32))) < POSE_TOL, "transformSum scan %d" % k
        assert bool(flags[0] & 8) == ("transformAftMapped" in e)
        if "transformAftMapped" in e:
            assert np.max(np.abs(poses[0, 6:12] - np.array(e["transformAftMapped"], np.float32))) < POSE_TOL, "transformAftMapped scan %d" % k
        if "transformMapped" in e:
            assert np.max(np.abs(pl.download("transformMapped") - np.array(e["transformMapped"], np.float32))) < POSE_TOL, "transformMapped scan %d" % k
        if "loopStats" in e:
            assert [int(v) for v in pl.download("loopStats")] == e["loopStats"], "performLoopClosure cadence scan %d" % k
            assert len(pl.download("cloudKeyPoses6D")) // 6 == e["numKeyFrames"], "key frames scan %d" % k
        if "loopClosure" in e:
            loops += e["loopClosure"][2]
            assert [int(v) for v in pl.download("loopClosure")] == e["loopClosure"], "loop record scan %d" % k
            assert np.max(np.abs(pl.download("icpTransform") - np.array(e["icpTransform"], np.float32))) < 1e-5, "ICP scan %d" % k
            kp = pl.download("cloudKeyPoses6D")
            assert np.max(np.abs(kp[-6:] - np.array(e["lastKeyPose"], np.float32))) < POSE_TOL, "latest key pose scan %d" % k
            assert np.max(np.abs(kp[:60] - np.array(e["firstKeyPoses"], np.float32))) < POSE_TOL, "corrected key poses scan %d" % k
    if g.get("loop"):
        assert loops >= 2, "the fixture closes no loop"
    pl.close()
