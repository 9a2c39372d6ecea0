import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch
import bench, lego_loam_b200
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params
L, H, T = 592, 4, 24
MASK = int(sys.argv[1]) if len(sys.argv) > 1 else 7
DEPTH = int(sys.argv[2]) if len(sys.argv) > 2 else 3
seeds = bench.lane_seeds(0, L)
xyzi, ring, npts, stamps, cap = bench.gen_inputs("vlp16", seeds, T, 4.0, pinned=True)[:5]
params = make_params("vlp16")
bounds = [(h * L) // H for h in range(H + 1)]
group = [Pipeline(params, lanes=bounds[h+1]-bounds[h], max_points=cap) for h in range(H)]
st = lambda t: np.full(L, stamps[t])
sub, col = [], []
def submit(t):
    t0 = time.perf_counter()
    for h, pl in enumerate(group):
        l0, l1 = bounds[h], bounds[h+1]
        pl.step(xyzi[t, l0:l1], ring[t, l0:l1], npts[t, l0:l1], st(t)[l0:l1], stage_mask=MASK)
        pl.poses_async(t % 8)
    sub.append(time.perf_counter() - t0)
def collect(t):
    t0 = time.perf_counter()
    r = [pl.poses_wait(t % 8)[0] for pl in group]
    col.append(time.perf_counter() - t0)
for t in range(4):
    submit(t); collect(t)
sub.clear(); col.clear()
t0 = time.perf_counter()
for t in range(4, T):
    submit(t)
    if t - (DEPTH - 1) >= 4: collect(t - (DEPTH - 1))
for t in range(T - (DEPTH - 1), T): collect(t)
tot = time.perf_counter() - t0
print("mask", MASK, "depth", DEPTH, "per step ms", 1e3 * tot / (T - 4))
print("submit ms", [round(1e3 * x, 2) for x in sub])
print("collect ms", [round(1e3 * x, 2) for x in col])
