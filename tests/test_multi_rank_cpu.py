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
        seeds = bench.lane_seeds(rank, 4)
        gathered = [None] * world
        dist.all_gather_object(gathered, seeds)
        slow = bench.max_over_ranks(10.0 + rank, world)         # rank 1 is the slow one
        r, w, l = bench.dist_env()
        # same synthetic generator, different seeds -> different scans on every rank
        x, ring, npts, stamps, cap = bench.gen_inputs("vlp16", seeds[:1], 1, speed=3.0)
        digest = float(x[0, 0, :npts[0, 0], :3].sum())
        digests = [None] * world
        dist.all_gather_object(digests, digest)
        class A: sensor = "vlp16"; steps = 1; warmup = 0; speed = 3.0; stages = 3; gpus = world
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
