"""Loop-closure workload (BASELINE.json config 5 shape, reduced): `lanes` VLP-16 sequences drive a closed circle in the
room scene with loopClosureEnableFlag on; every lane replays the same seeded scans (device-resident, stride 0) so all
lanes detect and close their loops in the same mapping calls -- the worst case for the 16 shared ICP slots.
Prints one JSON line: scans/s over the whole run (device time on the pipeline's stream), loop statistics, the per-kernel
CUDA-event profile of the loop-closure kernels from a second pass, and the reference (oracle/_ref vlp16_loop) on one
host core over the same scans.  Report tool; bench.py stays on BASELINE.json's configs[1]."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lego_loam_b200  # noqa: E402,F401
from lego_loam_b200.capi import Pipeline  # noqa: E402
from lego_loam_b200.params import make_params  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--lanes", type=int, default=64)
ap.add_argument("--scans", type=int, default=346)
ap.add_argument("--no-cpu", action="store_true")
args = ap.parse_args()

import torch  # noqa: E402

seq = SynthSequence("vlp16", seed=3, scene="room", speed=0.8, circle_radius=4.0)
xyzi, ring, npts = seq.batch(0, args.scans)
d_x = torch.from_numpy(xyzi).cuda(); d_r = torch.from_numpy(ring.view(np.int16)).cuda()
torch.cuda.synchronize()
cap = seq.cap
L = args.lanes


def run(profile):
    pl = Pipeline(make_params("vlp16", loop_closure=True), lanes=L, max_points=cap)
    if profile:
        pl.profile(1)
    pl.timer_record(0)
    t0 = time.time()
    for k in range(args.scans):
        pl.step(d_x.data_ptr() + k * cap * 16, d_r.data_ptr() + k * cap * 2, np.full(L, npts[k], np.int32), np.full(L, k * 0.1),
                stage_mask=7, on_device=True, stride_points=0)
    pl.timer_record(1)
    pl.sync()
    wall = time.time() - t0
    ms = pl.timer_elapsed_ms(0, 1)
    out = {"device_ms": ms, "wall_s": wall, "loopStats": [int(v) for v in pl.download("loopStats")], "launches": pl.launch_count(),
           "keyframes": int(len(pl.download("cloudKeyPoses6D")) // 6)}
    if profile:
        prof = pl.profile_read()
        out["kernels_ms"] = {k: [round(v[0], 3), v[1]] for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:14]}
        out["loop_kernels_ms"] = {k: [round(v[0], 3), v[1]] for k, v in prof.items() if k.startswith("lc_") or "history" in k}
    pl.close()
    return out


run(False)   # warm-up (allocations, module load)
timed = run(False)
prof = run(True)
res = {"workload": "VLP-16 closed circle (room), %d scans, loop closure on, %d lanes (same scans in every lane)" % (args.scans, L),
       "scans_per_sec": L * args.scans / (timed["device_ms"] * 1e-3), "device_ms": timed["device_ms"], "wall_s": timed["wall_s"],
       "loopStats_per_lane": timed["loopStats"], "keyframes_per_lane": timed["keyframes"], "gpu_launches": timed["launches"],
       "profile_pass": prof}
if not args.no_cpu:
    import refpy
    if refpy.available("vlp16_loop"):
        ref = refpy.RefPipeline("vlp16_loop", snapshots=False)
        t0 = time.time()
        for k in range(args.scans):
            ref.feed(xyzi[k, :npts[k]], ring[k, :npts[k]], k * 0.1, stages=7)
        dt = time.time() - t0
        res["cpu_reference"] = {"scans_per_sec": args.scans / dt, "cores": 1, "kind": "reference (oracle/_ref vlp16_loop)", "seconds": dt}
print(json.dumps(res))
