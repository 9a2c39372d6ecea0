#!/usr/bin/env python
"""Four pipelines fed the same scans in lock step (stages IP + FA): reports, per scan, the first named array in which two of
them differ.  The product must be deterministic; this hunts for a timing-dependent outcome."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lego_loam_b200  # noqa: E402,F401
from lego_loam_b200.capi import Pipeline  # noqa: E402
from lego_loam_b200.params import make_params  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402

n_scans = int(sys.argv[1]) if len(sys.argv) > 1 else 340
first_check = int(sys.argv[2]) if len(sys.argv) > 2 else 300
P = 4
L = 2
seqs = [SynthSequence("vlp16", seed=s, scene="room", speed=4.0, n_scans=n_scans, circle_radius=8.0) for s in range(L)]
cap = seqs[0].cap
pls = [Pipeline(make_params("vlp16"), lanes=L, max_points=cap) for _ in range(P)]
NAMES = ["labelMat", "segmentedCloud", "segmentedCloudAdj", "cloudCurvature", "cloudNeighborPicked", "cloudLabel", "cornerPointsSharp",
         "cornerPointsLessSharp", "surfPointsFlat", "surfPointsLessFlat", "transformCur", "transformSum", "laserCloudCornerLast", "laserCloudSurfLast", "lmState"]
reported = 0
for k in range(n_scans):
    x = np.zeros((L, cap, 4), np.float32); r = np.zeros((L, cap), np.uint16); n = np.zeros(L, np.int32)
    for l, q in enumerate(seqs):
        xyzi, ring = q.scan(k)
        n[l] = len(xyzi); x[l, :n[l]] = xyzi; r[l, :n[l]] = ring
    for pl in pls:
        pl.step(x, r, n, [0.1 * k] * L, stage_mask=3)
    poses = [pl.poses()[0] for pl in pls]
    same = all(np.array_equal(p.view(np.uint32), poses[0].view(np.uint32)) for p in poses)
    if k >= first_check or not same:
        for lane in range(L):
            for name in NAMES:
                try:
                    arrs = [np.ascontiguousarray(pl.download(name, lane)) for pl in pls]
                except Exception as ex:  # noqa: BLE001
                    continue
                ok = all(a.shape == arrs[0].shape and np.array_equal(a.view(np.uint8), arrs[0].view(np.uint8)) for a in arrs)
                if not ok:
                    which = [i for i, a in enumerate(arrs) if a.shape != arrs[0].shape or not np.array_equal(a.view(np.uint8), arrs[0].view(np.uint8))]
                    a, b = arrs[0], arrs[which[0]]
                    if a.shape == b.shape:
                        d = np.nonzero(a.reshape(-1) != b.reshape(-1))[0]
                        print("scan %d lane %d: %s differs in pipelines %s: %d elements, first at %d: %r vs %r" % (k, lane, name, which, len(d), d[0], a.reshape(-1)[d[0]], b.reshape(-1)[d[0]]))
                    else:
                        print("scan %d lane %d: %s shapes %s vs %s" % (k, lane, name, a.shape, b.shape))
                    reported += 1
                    for i in (0, which[0]):
                        print("     pipeline %d: lmState %s matP %s featCounts %s lastCounts %s" % (i, pls[i].download("lmState", lane), np.round(pls[i].download("matP", lane), 4),
                              pls[i].download("featCounts", lane), pls[i].download("lastCounts", lane)))
                        print("                 transformCur %s" % pls[i].download("transformCur", lane))
                    break
    if reported >= 6:
        break
print("done, scans run:", k + 1, "differences reported:", reported)
