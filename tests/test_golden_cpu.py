"""CPU: the oracle (reference sources compiled verbatim, oracle/_ref) must reproduce the committed golden fixtures
tests/golden/*.json bit for bit (CRCs) and float for float (poses).  The fixtures were written by tools/make_golden.py
from the same sources; a drift here means the oracle build changed (compiler, libm, shim), not the product."""
import glob
import json
import os
import zlib

import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
import refpy
from lego_loam_b200.synth import SynthSequence

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.json")))


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def test_golden_fixtures_exist():
    assert len(GOLDEN) >= 2


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_reproduces_golden(path):
    g = json.load(open(path))
    cfg = g["sensor"] + ("_loop" if g.get("loop") else "")
    if not refpy.available(cfg):
        pytest.skip("oracle/_ref for %s not built" % cfg)
    seq = SynthSequence(g["sensor"], seed=g["seed"], scene=g["scene"], speed=g["speed"], circle_radius=g.get("circle_radius", 0.0))
    ref = refpy.RefPipeline(cfg)
    j = 0
    for k, e in enumerate(g["scans"]):
        if g.get("imu"):
            msgs, j = seq.imu_until(j, seq.stamp(k) + 0.1)
            for (t, q, w, a) in msgs:
                ref.imu(t, q, w, a)
            assert crc(np.concatenate([np.concatenate([[t], q, w, a]) for (t, q, w, a) in msgs]) if msgs else np.zeros(0)) == e["imu_crc"], "IMU generator drifted at scan %d" % k
        xyzi, ring = seq.scan(k)
        assert crc(xyzi) == e["input_crc"], "synthetic generator drifted at scan %d" % k
        f = ref.feed(xyzi, ring, seq.stamp(k), stages=g["stages"])
        assert f == e["flags"]
        if "labelMat_crc" in e:
            assert crc(ref.get("labelMat")) == e["labelMat_crc"]
            assert crc(ref.cloud("segmentedCloud").view(np.uint32)) == e["segmentedCloud_crc"]
            assert crc(ref.cloud("cornerPointsSharp").view(np.uint32)) == e["cornerPointsSharp_crc"]
            assert crc(ref.cloud("surfPointsFlat").view(np.uint32)) == e["surfPointsFlat_crc"]
        if "segmentedCloudAdj_crc" in e:
            assert crc(ref.cloud("segmentedCloudAdj").view(np.uint32)) == e["segmentedCloudAdj_crc"]
        assert np.array_equal(ref.get("transformSum"), np.array(e["transformSum"], np.float32))
        if "transformAftMapped" in e:
            assert np.array_equal(ref.get("transformAftMapped"), np.array(e["transformAftMapped"], np.float32))
        if "transformMapped" in e:
            assert np.array_equal(ref.get("transformMapped"), np.array(e["transformMapped"], np.float32))
        if "loopStats" in e:
            assert [int(v) for v in ref.get("loopStats")] == e["loopStats"]
        if "loopClosure" in e:
            assert [int(v) for v in ref.get("loopClosure")] == e["loopClosure"]
            assert np.array_equal(ref.get("icpTransform"), np.array(e["icpTransform"], np.float32))
            assert np.array_equal(ref.get("cloudKeyPoses6D")[-6:], np.array(e["lastKeyPose"], np.float32))
