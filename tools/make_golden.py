#!/usr/bin/env python
"""Writes tests/golden/<sensor>_<scene>_seed<k>.json from the reference's own code (oracle/_ref: the reference sources
compiled verbatim): per scan the poses and CRC32s of the bit-exact outputs.  The fixture pins the oracle build itself
(a different compiler / libm on another machine would show up here) and gives the GPU tests a target that does not
need the oracle at run time.   python tools/make_golden.py [sensor] [scene] [seed] [n_scans]"""
import json
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lego_loam_b200  # noqa: E402,F401
import refpy  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def make(sensor="vlp16", scene="street", seed=7, n=24, speed=4.0):
    seq = SynthSequence(sensor, seed=seed, scene=scene, speed=speed)
    ref = refpy.RefPipeline(sensor)
    scans = []
    for k in range(n):
        xyzi, ring = seq.scan(k)
        f = ref.feed(xyzi, ring, seq.stamp(k), stages=15)
        e = {"flags": int(f), "input_crc": crc(xyzi), "labelMat_crc": crc(ref.get("labelMat")),
             "segmentedCloud_crc": crc(ref.cloud("segmentedCloud").view(np.uint32)),
             "cornerPointsSharp_crc": crc(ref.cloud("cornerPointsSharp").view(np.uint32)),
             "cornerPointsLessSharp_crc": crc(ref.cloud("cornerPointsLessSharp").view(np.uint32)),
             "surfPointsFlat_crc": crc(ref.cloud("surfPointsFlat").view(np.uint32)),
             "surfPointsLessFlat_n": int(len(ref.cloud("surfPointsLessFlat"))),
             "transformSum": [float(v) for v in ref.get("transformSum")]}
        if f & 8:
            e["transformAftMapped"] = [float(v) for v in ref.get("transformAftMapped")]
        if f & 16:
            e["transformMapped"] = [float(v) for v in ref.get("transformMapped")]
        scans.append(e)
    return {"sensor": sensor, "scene": scene, "seed": seed, "speed": speed, "stages": 15,
            "generator": "tools/make_golden.py (oracle/_ref = reference sources compiled verbatim against oracle/shim)", "scans": scans}


if __name__ == "__main__":
    sensor = sys.argv[1] if len(sys.argv) > 1 else "vlp16"
    scene = sys.argv[2] if len(sys.argv) > 2 else "street"
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 7
    n = int(sys.argv[4]) if len(sys.argv) > 4 else 24
    out = os.path.join(ROOT, "tests", "golden", "%s_%s_seed%d.json" % (sensor, scene, seed))
    with open(out, "w") as f:
        json.dump(make(sensor, scene, seed, n), f, indent=1)
    print(out)
