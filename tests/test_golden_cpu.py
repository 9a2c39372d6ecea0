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
    if not refpy.available(g["sensor"]):
        pytest.skip("oracle/_ref for %s not built" % g["sensor"])
    seq = SynthSequence(g["sensor"], seed=g["seed"], scene=g["scene"], speed=g["speed"])
    ref = refpy.RefPipeline(g["sensor"])
    for k, e in enumerate(g["scans"]):
        xyzi, ring = seq.scan(k)
        assert crc(xyzi) == e["input_crc"], "synthetic generator drifted at scan %d" % k
        f = ref.feed(xyzi, ring, seq.stamp(k), stages=g["stages"])
        assert f == e["flags"]
        assert crc(ref.get("labelMat")) == e["labelMat_crc"]
        assert crc(ref.cloud("segmentedCloud").view(np.uint32)) == e["segmentedCloud_crc"]
        assert crc(ref.cloud("cornerPointsSharp").view(np.uint32)) == e["cornerPointsSharp_crc"]
        assert crc(ref.cloud("surfPointsFlat").view(np.uint32)) == e["surfPointsFlat_crc"]
        assert np.array_equal(ref.get("transformSum"), np.array(e["transformSum"], np.float32))
        if "transformAftMapped" in e:
            assert np.array_equal(ref.get("transformAftMapped"), np.array(e["transformAftMapped"], np.float32))
        if "transformMapped" in e:
            assert np.array_equal(ref.get("transformMapped"), np.array(e["transformMapped"], np.float32))
