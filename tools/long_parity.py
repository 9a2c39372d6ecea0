#!/usr/bin/env python
"""Free-running parity over a long trajectory: CUDA pipeline vs the reference's own code (oracle/_ref), both fed the
same scans, neither corrected by the other.  Prints the worst pose differences and the first scan where a bit-exact
quantity (labels, feature clouds) differs.   python tools/long_parity.py [sensor] [n_scans] [seed]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lego_loam_b200  # noqa: E402,F401
import refpy  # noqa: E402
from lego_loam_b200.capi import Pipeline  # noqa: E402
from lego_loam_b200.params import make_params  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402


def run(sensor="vlp16", n=300, seed=11, speed=4.0, verbose=True):
    seq = SynthSequence(sensor, seed=seed, scene="street", speed=speed, n_scans=n)
    ref = refpy.RefPipeline(sensor)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
    worst = dict(sum=0.0, aft=0.0)
    first_bit_diff = None
    first_map_diff = None
    reported = False
    for k in range(n):
        xyzi, ring = seq.scan(k)
        ref.feed(xyzi, ring, seq.stamp(k), stages=7)
        x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
        r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
        pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=7)
        poses, flags = pl.poses()
        ds = float(np.max(np.abs(poses[0, :6] - ref.get("transformSum"))))
        da = float(np.max(np.abs(poses[0, 6:12] - ref.get("transformAftMapped"))))
        worst["sum"] = max(worst["sum"], ds); worst["aft"] = max(worst["aft"], da)
        if first_bit_diff is None:
            for name in ("cornerPointsSharp", "surfPointsFlat", "cornerPointsLessSharp"):
                g, w = pl.download(name), ref.cloud(name)
                if g.shape != w.shape or not np.array_equal(g.view(np.uint32), w.view(np.uint32)):
                    first_bit_diff = (k, name); break
            if first_bit_diff is None and not np.array_equal(pl.download("labelMat"), ref.get("labelMat")):
                first_bit_diff = (k, "labelMat")
        if flags[0] & 8:
            ids_g, ids_r = pl.download("surroundingExistingKeyPosesID"), ref.get("surroundingExistingKeyPosesID")
            sizes_g = [len(pl.download(nm)) for nm in ("laserCloudCornerFromMapDS", "laserCloudSurfFromMapDS", "laserCloudSurfTotalLastDS")]
            sizes_r = [len(ref.cloud(nm)) for nm in ("laserCloudCornerFromMapDS", "laserCloudSurfFromMapDS", "laserCloudSurfTotalLastDS")]
            if first_map_diff is None and (not np.array_equal(ids_g, ids_r) or sizes_g != sizes_r):
                first_map_diff = (k, "ids equal %s" % np.array_equal(ids_g, ids_r), sizes_g, sizes_r, int(pl.download("moState")[3]))
                print("first map difference:", first_map_diff, flush=True)
            it_g = int(pl.download("moState")[3])
            if verbose and da > 1e-4 and not reported:
                reported = True
                print("pose difference above 1e-4 first at scan %d: %.3e, device LM iterations %d, ids equal %s, sizes %s vs %s"
                      % (k, da, it_g, np.array_equal(ids_g, ids_r), sizes_g, sizes_r), flush=True)
        if verbose and (k % 50 == 49 or k == n - 1):
            print("scan %4d  |d transformSum| %.2e  |d transformAftMapped| %.2e  worst so far %.2e / %.2e  keyframes %d vs %d  first bit diff %s"
                  % (k, ds, da, worst["sum"], worst["aft"], int(pl.download("moState")[1]), int(ref.get("numKeyFrames")[0]), first_bit_diff), flush=True)
    pl.close(); ref.close()
    return worst, first_bit_diff


if __name__ == "__main__":
    a = sys.argv[1:]
    run(a[0] if a else "vlp16", int(a[1]) if len(a) > 1 else 300, int(a[2]) if len(a) > 2 else 11)
