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


def make(sensor="vlp16", scene="street", seed=7, n=24, speed=4.0, imu=False, loop=False, circle_radius=0.0):
    """imu: an ideal 100 Hz IMU rides along (optional IMU path); loop: the reference is the build with utility.h:104
    loopClosureEnableFlag = true and the trajectory a closed circle (entries are then reduced to poses + loop records)."""
    seq = SynthSequence(sensor, seed=seed, scene=scene, speed=speed, circle_radius=circle_radius)
    ref = refpy.RefPipeline(sensor + ("_loop" if loop else ""))
    stages = 7 if loop else 15
    scans = []
    j = 0
    for k in range(n):
        e = {}
        if imu:
            msgs, j = seq.imu_until(j, seq.stamp(k) + 0.1)
            for (t, q, w, a) in msgs:
                ref.imu(t, q, w, a)
            e["imu_crc"] = crc(np.concatenate([np.concatenate([[t], q, w, a]) for (t, q, w, a) in msgs]) if msgs else np.zeros(0))
        xyzi, ring = seq.scan(k)
        f = ref.feed(xyzi, ring, seq.stamp(k), stages=stages)
        e.update({"flags": int(f), "input_crc": crc(xyzi), "transformSum": [float(v) for v in ref.get("transformSum")]})
        if not loop:
            e.update({"labelMat_crc": crc(ref.get("labelMat")),
                      "segmentedCloud_crc": crc(ref.cloud("segmentedCloud").view(np.uint32)),
                      "cornerPointsSharp_crc": crc(ref.cloud("cornerPointsSharp").view(np.uint32)),
                      "cornerPointsLessSharp_crc": crc(ref.cloud("cornerPointsLessSharp").view(np.uint32)),
                      "surfPointsFlat_crc": crc(ref.cloud("surfPointsFlat").view(np.uint32)),
                      "surfPointsLessFlat_n": int(len(ref.cloud("surfPointsLessFlat")))})
        if imu:
            e["segmentedCloudAdj_crc"] = crc(ref.cloud("segmentedCloudAdj").view(np.uint32))
        if f & 8:
            e["transformAftMapped"] = [float(v) for v in ref.get("transformAftMapped")]
            if loop:
                e["numKeyFrames"] = int(ref.get("numKeyFrames")[0])
                e["loopStats"] = [int(v) for v in ref.get("loopStats")]
                try:
                    lc = ref.get("loopClosure")
                    if lc[1]:
                        e["loopClosure"] = [int(v) for v in lc]
                        e["icpTransform"] = [float(v) for v in ref.get("icpTransform")]
                        e["lastKeyPose"] = [float(v) for v in ref.get("cloudKeyPoses6D")[-6:]]
                        e["firstKeyPoses"] = [float(v) for v in ref.get("cloudKeyPoses6D")[:60]]
                except KeyError:
                    pass
        if f & 16:
            e["transformMapped"] = [float(v) for v in ref.get("transformMapped")]
        scans.append(e)
    return {"sensor": sensor, "scene": scene, "seed": seed, "speed": speed, "stages": stages, "imu": bool(imu), "loop": bool(loop),
            "circle_radius": circle_radius,
            "generator": "tools/make_golden.py (oracle/_ref = reference sources compiled verbatim against oracle/shim)", "scans": scans}


if __name__ == "__main__":
    sensor = sys.argv[1] if len(sys.argv) > 1 else "vlp16"
    scene = sys.argv[2] if len(sys.argv) > 2 else "street"
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 7
    n = int(sys.argv[4]) if len(sys.argv) > 4 else 24
    mode = sys.argv[5] if len(sys.argv) > 5 else ""
    kw = {}
    if mode == "imu":
        kw = dict(imu=True, speed=3.0)
    elif mode == "loop":
        kw = dict(loop=True, speed=0.8, circle_radius=4.0)
    out = os.path.join(ROOT, "tests", "golden", "%s_%s_seed%d%s.json" % (sensor, scene, seed, "_" + mode if mode else ""))
    with open(out, "w") as f:
        json.dump(make(sensor, scene, seed, n, **kw), f, indent=None if mode == "loop" else 1)
    print(out)
