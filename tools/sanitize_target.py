#!/usr/bin/env python
"""Small full-pipeline run for compute-sanitizer (memcheck / racecheck, one tool per gpurun call):
2 lanes x 14 VLP-16 scans through ImageProjection, FeatureAssociation, mapOptimization (incremental local map: 3 mapping calls)
and TransformFusion, plus one lego_voxel_grid_noalloc call.  Exercises the lock-free union-find, the atomicCAS hash insert,
the per-warp work queues of the odometry kernel and the miss-list compaction."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lego_loam_b200  # noqa: E402,F401
from lego_loam_b200.capi import Pipeline, selftest_voxel_grid  # noqa: E402
from lego_loam_b200.params import make_params  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402

n_scans = int(sys.argv[1]) if len(sys.argv) > 1 else 14
seqs = [SynthSequence("vlp16", seed=71 + l, scene="street", speed=6.0, dropout=0.01) for l in range(2)]
cap = seqs[0].cap
pl = Pipeline(make_params("vlp16"), lanes=2, max_points=cap)
for k in range(n_scans):
    x = np.zeros((2, cap, 4), np.float32); r = np.zeros((2, cap), np.uint16); n = np.zeros(2, np.int32)
    for l, q in enumerate(seqs):
        xyzi, ring = q.scan(k)
        n[l] = len(xyzi); x[l, :n[l]] = xyzi; r[l, :n[l]] = ring
    pl.step(x, r, n, [0.1 * k] * 2, stage_mask=15)
poses, flags = pl.poses()
assert np.isfinite(poses).all() and np.abs(poses[:, 6:12]).max() > 0.05
print("labels", int(pl.download("labelMat").max()), "key frames", int(pl.download("moState")[1]), "poses", poses[0, 6:12])
pl.close()
out = selftest_voxel_grid([np.random.default_rng(0).uniform(-5, 5, (5000, 4)).astype(np.float32)], 0.4)
print("voxel grid", len(out[0]))
print("sanitize target ok")
