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
    seq = SynthSequence(g["sensor"], seed=g["seed"], scene=g["scene"], speed=g["speed"])
    pl = Pipeline(make_params(g["sensor"]), lanes=1, max_points=seq.cap)
    for k, e in enumerate(g["scans"]):
        xyzi, ring = seq.scan(k)
        assert crc(xyzi) == e["input_crc"], "synthetic generator drifted at scan %d" % k
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=g["stages"])
        poses, flags = pl.poses()
        assert crc(pl.download("labelMat")) == e["labelMat_crc"], "labelMat scan %d" % k
        assert crc(pl.download("segmentedCloud").view(np.uint32)) == e["segmentedCloud_crc"], "segmentedCloud scan %d" % k
        assert crc(pl.download("cornerPointsSharp").view(np.uint32)) == e["cornerPointsSharp_crc"], "cornerPointsSharp scan %d" % k
        assert crc(pl.download("cornerPointsLessSharp").view(np.uint32)) == e["cornerPointsLessSharp_crc"], "cornerPointsLessSharp scan %d" % k
        assert crc(pl.download("surfPointsFlat").view(np.uint32)) == e["surfPointsFlat_crc"], "surfPointsFlat scan %d" % k
        assert len(pl.download("surfPointsLessFlat")) == e["surfPointsLessFlat_n"]
        assert np.max(np.abs(poses[0, :6] - np.array(e["transformSum"], np.float32))) < POSE_TOL, "transformSum scan %d" % k
        assert bool(flags[0] & 8) == ("transformAftMapped" in e)
        if "transformAftMapped" in e:
            assert np.max(np.abs(poses[0, 6:12] - np.array(e["transformAftMapped"], np.float32))) < POSE_TOL, "transformAftMapped scan %d" % k
        if "transformMapped" in e:
            assert np.max(np.abs(pl.download("transformMapped") - np.array(e["transformMapped"], np.float32))) < POSE_TOL, "transformMapped scan %d" % k
    pl.close()
