"""Free-running comparison of the IMU path with the reference: per scan, the pose differences (diagnostic tool; the
asserted version is tests/test_gpu_imu.py)."""
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

sensor = sys.argv[1] if len(sys.argv) > 1 else "vlp16"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 60
use_imu = (sys.argv[3] if len(sys.argv) > 3 else "1") == "1"
seed = int(sys.argv[4]) if len(sys.argv) > 4 else 5
speed = float(sys.argv[5]) if len(sys.argv) > 5 else 3.0
yaw_amp = float(sys.argv[6]) if len(sys.argv) > 6 else 0.25
sway = float(sys.argv[7]) if len(sys.argv) > 7 else 1.5
seq = SynthSequence(sensor, seed=seed, scene="street", speed=speed, yaw_amp=yaw_amp, sway_amp=sway)
ref = refpy.RefPipeline(sensor)
pl = Pipeline(make_params(sensor), lanes=1, max_points=seq.cap)
j = 0
for k in range(n):
    if use_imu:
        msgs, j = seq.imu_until(j, seq.stamp(k) + 0.1)
        for (t, q, w, a) in msgs:
            ref.imu(t, q, w, a); pl.imu(0, t, q, w, a)
    xyzi, ring = seq.scan(k)
    rf = ref.feed(xyzi, ring, seq.stamp(k), stages=7)
    x = np.zeros((1, seq.cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
    r = np.zeros((1, seq.cap), np.uint16); r[0, :len(ring)] = ring
    pl.step(x, r, [len(xyzi)], [seq.stamp(k)], stage_mask=7)
    poses, flags = pl.poses()
    line = "scan %3d d_cur %.2e d_sum %.2e" % (k, np.max(np.abs(pl.download("transformCur") - ref.get("transformCur"))), np.max(np.abs(poses[0, :6] - ref.get("transformSum"))))
    def dmax(a, b):
        return float(np.abs(a - b).max()) if a.shape == b.shape and a.size else -1.0
    line += " d_segAdj %.2e d_cornerLast %.2e d_surfLast %.2e" % (dmax(pl.download("segmentedCloudAdj"), ref.cloud("segmentedCloudAdj")),
        dmax(pl.download("laserCloudCornerLast"), ref.cloud("laserCloudCornerLast")), dmax(pl.download("laserCloudSurfLast"), ref.cloud("laserCloudSurfLast")))
    if rf & 8:
        line += "  MAP d_tobe %.2e d_aft %.2e d_init %.2e iters %d sizes %s" % (
            np.max(np.abs(pl.download("transformTobeMapped") - ref.get("transformTobeMapped"))), np.max(np.abs(poses[0, 6:12] - ref.get("transformAftMapped"))),
            -1, int(pl.download("moState")[3]),
            [len(pl.download(nm)) - len(ref.cloud(nm)) for nm in ("laserCloudCornerFromMapDS", "laserCloudSurfFromMapDS", "laserCloudSurfTotalLastDS")])
    print(line)
