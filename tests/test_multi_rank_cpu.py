"""CPU, world_size 2 over gloo: the host logic of the N>1 path.  The hot path shards across independent sequences
only ("replicas"), so what has to hold is: ranks own disjoint sequences, the step time is the max over
ranks, only rank 0 reports, and the reference arm runs on rank 0 alone."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        seeds = bench.rank_seeds(rank, world, 4)
        gathered = [None] * world
        dist.all_gather_object(gathered, seeds)
        slow = bench.max_over_ranks(10.0 + rank, world)         # rank 1 is the slow one
        r, w, l = bench.dist_env()
        # same synthetic generator, different seeds -> different scans on every rank
        inp = bench.Inputs("vlp16", seeds[:2], 3, "drive", 3.0, 3, pinned=False)
        assert inp.blob.shape == (3, 2, inp.cap * 14) and inp.npts.shape == (3, 2)
        digest = float(inp.scan_xyz(0, 0).sum())
        # one slot per (scan, sequence): the scan's points and, directly behind them, its ring channel
        from lego_loam_b200.synth import SynthSequence
        xyzi, ring = SynthSequence("vlp16", seed=seeds[1], scene="street", speed=3.0, n_scans=64).scan(2)
        n = int(inp.npts[2, 1])
        assert n == len(xyzi) and np.array_equal(inp.scan_xyz(2, 1), xyzi[:, :3])
        assert np.array_equal(inp.blob[2, 1, n * 12:n * 14].view(np.uint16), ring)
        # lane l replays sequence l % D: the per-lane address table points at the slots of scan 0
        xt, seq = inp.tables(inp.blob.ctypes.data, 5)
        assert list(seq) == [0, 1, 0, 1, 0] and int(xt[1] - xt[0]) == inp.cap * 14 and inp.scan_bytes == 2 * inp.cap * 14
        digests = [None] * world
        dist.all_gather_object(digests, digest)
        class A: sensor = "vlp16"; steps = 1; warmup = 0; speed = 3.0; stages = 3; gpus = world; workload = "drive"; scans_per_step = 0
        ref_line = None
        if rank != 0:
            ref_line = bench.run_reference_arm(A)                # must be None without doing work
        q.put((rank, gathered, slow, (r, w, l), digests, ref_line))
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_sequences_and_reduce_max():
    world, port = 2, 29000 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, gathered, slow, env, digests, ref_line in res:
        assert env == (rank, world, rank)
        assert slow == 11.0                                      # max over ranks, identical on every rank
        flat = [s for g in gathered for s in g]
        assert len(set(flat)) == len(flat) == 8                  # disjoint sequences
        assert digests[0] != digests[1]
        assert ref_line is None


def test_window_is_the_same_for_both_arms():
    """the driver's --steps 20 --warmup 5: 125 untimed + 500 timed scans per lane, starting at the steady-state local map"""
    sys.path.insert(0, ROOT)
    import bench
    class A: sensor = "vlp16"; steps = 20; warmup = 5; speed = 4.0; scans_per_step = 0
    w = bench.Window(A, "vlp16")
    assert (w.S, w.first_timed, w.T, w.prime) == (25, 125, 625, 0)
    assert w.describe()["timed_scan_range"] == [125, 624]
    w = bench.Window(A, "hdl64")                  # fewer scans per step: the window is primed up to the steady state
    assert (w.S, w.first_timed, w.T, w.prime) == (10, 125, 325, 75)
    A.steps, A.warmup = 2, 1
    w = bench.Window(A, "vlp16")
    assert w.first_timed == 125 and w.prime == 100 and w.T == 175
    assert bench.default_distinct("vlp16", 1, 592, 625, 28800) == 32
    assert bench.default_distinct("vlp16", 8, 592, 625, 28800) == 4


def test_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` as the driver launches it (no GPU involved): one JSON line with the arm's own metric, the
    window both arms time, a cpu_baseline describing the run and an e2e block that repeats the value with zero copy bytes."""
    import json
    import subprocess
    from conftest import require_ref
    require_ref("vlp16")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "3"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.strip().splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "scans_per_sec" and d["unit"] == "scans/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 3 and d["vs_baseline"] is None
    lo, hi = d["config"]["timed_scan_range"]
    assert lo >= 125 and hi - lo + 1 == d["config"]["timed_scans_per_lane"] == 25      # steady-state window, one 25-scan step
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"] == {"value": d["value"], "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_numa_placement_never_fails_the_bench():
    """The page-locked input arena is placed near the GPU when the box has several NUMA nodes; without a driver, without
    topology or on one node the helper must report why it did nothing and leave the process untouched."""
    sys.path.insert(0, ROOT)
    import bench
    before = os.sched_getaffinity(0)
    info = bench.numa_place_near_gpu(0)
    bench.numa_reset()
    assert isinstance(info, dict) and "note" in info and info["note"]
    if info["node"] is None:
        assert os.sched_getaffinity(0) == before
