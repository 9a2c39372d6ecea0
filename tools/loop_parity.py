"""Free-running comparison of the loop-closure pipeline with the reference (oracle/_ref vlp16_loop) on the closed-circle
trajectory: prints, per mapping call, the pose differences, map sizes and loop records of both sides.  Diagnostic /
report tool (the asserted version is tests/test_gpu_loop_closure.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lego_loam_b200  # noqa: E402,F401
from lego_loam_b200.capi import Pipeline  # noqa: E402
from lego_loam_b200.params import make_params  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402
import refpy  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 346
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 3
seq = SynthSequence("vlp16", seed=seed, scene="room", speed=0.8, circle_radius=4.0)
ref = refpy.RefPipeline("vlp16_loop")
pl = Pipeline(make_params("vlp16", loop_closure=True), lanes=1, max_points=seq.cap)
worst = 0.0
for k in range(n):
    xyzi, ring = seq.scan(k)
    rf = ref.feed(xyzi, ring, seq.stamp(k), stages=7)
    x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
    r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
    pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=7)
    poses, flags = pl.poses()
    if not (rf & 8):
        continue
    d_aft = np.max(np.abs(poses[0, 6:12] - ref.get("transformAftMapped")))
    d_sum = np.max(np.abs(poses[0, :6] - ref.get("transformSum")))
    gk, wk = pl.download("cloudKeyPoses6D").reshape(-1, 6), ref.get("cloudKeyPoses6D").reshape(-1, 6)
    d_kp = np.max(np.abs(gk - wk)) if gk.shape == wk.shape else -1
    sizes = [len(pl.download(nm)) == len(ref.cloud(nm)) for nm in ("laserCloudCornerFromMapDS", "laserCloudSurfFromMapDS", "laserCloudSurfTotalLastDS")]
    worst = max(worst, d_aft)
    line = "scan %3d  d_sum %.2e  d_aft %.2e  d_keyposes %.2e  kf %d/%d  mapsizes_equal %s iters %d" % (
        k, d_sum, d_aft, d_kp, len(gk), len(wk), sizes, int(pl.download("moState")[3]))
    try:
        wl = ref.get("loopClosure"); gl = pl.download("loopClosure")
        if wl[1] or gl[1]:
            line += "  LOOP ref %s dev %s  dT %.2e fit %.5f/%.5f" % (list(wl), list(gl), np.max(np.abs(pl.download("icpTransform") - ref.get("icpTransform"))),
                                                                   float(pl.download("icpFitness")[0]), float(ref.get("icpFitness")[0]))
    except KeyError:
        pass
    print(line)
print("worst d_aft %.3e" % worst, "loopStats", list(pl.download("loopStats")), list(ref.get("loopStats")))
