#!/usr/bin/env python
"""Runs the same loop-closure sequences through fresh pipelines several times (host input, host input again, device input)
and reports the first scan at which the poses of two runs differ, with the loop-closure bookkeeping around it."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lego_loam_b200  # noqa: E402,F401
from lego_loam_b200.capi import Pipeline  # noqa: E402
from lego_loam_b200.params import make_params  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402

LOOP = int(os.environ.get("LOOP", "1"))
STAGES = int(os.environ.get("STAGES", "7"))
n_scans = int(sys.argv[1]) if len(sys.argv) > 1 else 400
L = int(sys.argv[2]) if len(sys.argv) > 2 else 2
seqs = [SynthSequence("vlp16", seed=s, scene="room", speed=4.0, n_scans=n_scans, circle_radius=8.0) for s in range(L)]
cap = seqs[0].cap
X = np.zeros((n_scans, L, cap, 4), np.float32); R = np.zeros((n_scans, L, cap), np.uint16); N = np.zeros((n_scans, L), np.int32)
for l, q in enumerate(seqs):
    x, r, n = q.batch(0, n_scans)
    X[:, l] = x; R[:, l] = r; N[:, l] = n


def run(mode):
    import torch
    pl = Pipeline(make_params("vlp16", loop_closure=bool(LOOP), max_keyframes=1024), lanes=L, max_points=cap)
    out, loops = [], []
    if mode == "device":
        dX = torch.from_numpy(X).cuda(); dR = torch.from_numpy(R.view(np.int16)).cuda(); torch.cuda.synchronize()
    for k in range(n_scans):
        st = [0.1 * k] * L
        if mode == "device":
            pl.step(dX[k].data_ptr(), dR[k].data_ptr(), N[k], st, on_device=True, stride_points=cap, stage_mask=STAGES)
        elif mode == "async":
            pl.step(X[k], R[k], N[k], st, stage_mask=STAGES); pl.poses_async(k % 8)
        else:
            pl.step(X[k], R[k], N[k], st, stage_mask=STAGES)
        p = pl.poses_wait(k % 8)[0] if mode == "async" else pl.poses()[0]
        out.append(p.copy())
        loops.append([tuple(pl.download("loopStats", l)) for l in range(L)] if LOOP else [])
    pl.close()
    return np.array(out), loops


order = sys.argv[3].split(",") if len(sys.argv) > 3 else ["host", "host", "async", "device"]
base, bl = run(order[0])
for mode in order[1:]:
    o, lo = run(mode)
    diff = np.abs(o - base).reshape(n_scans, -1).max(1)
    bad = np.nonzero(diff > 0)[0]
    if len(bad) == 0:
        print(mode, "identical to the first", order[0], "run | final loop stats", lo[-1])
        continue
    k = bad[0]
    comp = np.abs(o[k] - base[k])            # [L, 18]: transformSum | transformAftMapped | transformBefMapped
    print(mode, "vs first %s run: first difference at scan %d: max |d transformSum| %.2e, |d aft| %.2e, |d bef| %.2e; max over the run %.3e; loop stats there %s vs %s; final %s"
          % (order[0], k, comp[:, :6].max(), comp[:, 6:12].max(), comp[:, 12:].max(), diff.max(), lo[k], bl[k], lo[-1]))
    print("   scans with a loop call / closure up to there:", [i for i in range(1, k + 1) if lo[i] != lo[i - 1]][-6:])
