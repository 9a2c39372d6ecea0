#!/usr/bin/env python
"""bench.py -- throughput of the LeGO-LOAM per-scan hot path (ImageProjection -> FeatureAssociation ->
mapOptimization scan-to-map) on B200, in scans/s, on synthetic scans of the named sensor shape.

Workload = BASELINE.json configs[1]: full pipeline on long synthetic ground-vehicle trajectories.  Every lane
(independent sequence) drives one trajectory from scan 0; a "step" is `scans_per_step` consecutive scans of every lane
(25 for the 16-ring sensor), so the driver's `--steps 20 --warmup 5` is 125 untimed + 500 timed scans per lane.  The
timed window never starts before scan `ceil(surroundingKeyframeSearchRadius / (speed * scanPeriod))` (= 125 at 4 m/s):
from there on the local map of extractSurroundingKeyFrames (mapOptmization.cpp:956-1065) holds every key frame
within 50 m and no longer grows.  BOTH arms (this one and --impl reference) process scans 0 .. T-1 and time exactly
the same scan indices; config.timed_scan_range states them.

value = lanes * timed scans * n_gpus / device time with the inputs resident in HBM; e2e = the same through the C ABI
(lego_pipeline_step_lanes) from pinned HOST buffers, the H2D copy of every scan of every lane and the D2H of every
scan's poses inside the timed region.  Lanes replay one of `distinct_sequences` generated trajectories (592 lanes x
625 scans of private input would be 150 GB); every lane still owns its state, and in the e2e arm every lane's scan
crosses PCIe separately.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--lanes L] [--sensor vlp16] [--impl reference]
                    [--workload drive|loop]

--impl reference times the reference's own CPU code (oracle/_ref, the reference sources compiled verbatim) on the
host cores, one independent sequence per core, over the same scan range.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

METRIC = "scans_per_sec"
UNIT = "scans/s"
SHAPES = {"vlp16": "VLP-16 16x1800", "vlp16_noring": "VLP-16 16x1800 (no ring channel)", "hdl32": "HDL-32E 32x1800",
          "hdl64": "64-beam 64x2083", "vls128": "VLS-128 128x1800", "os1_16": "OS1-16 16x1024", "os1_64": "OS1-64 64x1024"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--lanes", type=int, default=0, help="independent sequences per GPU stepped in lock step (0 = 592 for 16 rings, 296 for 32/64, 148 above)")
    ap.add_argument("--handles", type=int, default=4, help="pipeline handles (CUDA streams) the lanes of one GPU are split over")
    ap.add_argument("--stagger", type=int, default=0, help="1: handle h runs h scans ahead of handle 0 so the mapping calls of the handles do not coincide")
    ap.add_argument("--sensor", default="vlp16")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="drive", choices=["drive", "loop"],
                    help="drive: street trajectories, loop closure off (BASELINE config 2); loop: closed circles with loopClosureEnableFlag (config 5)")
    ap.add_argument("--stages", type=int, default=7, help="bit0 ImageProjection, bit1 FeatureAssociation, bit2 mapOptimization")
    ap.add_argument("--speed", type=float, default=4.0)
    ap.add_argument("--scans-per-step", type=int, default=0, help="scans of every lane in one step (0 = 25 for 16 rings, 10 for 32/64, 5 above)")
    ap.add_argument("--distinct", type=int, default=0, help="distinct generated sequences per GPU the lanes replay (0 = auto)")
    ap.add_argument("--input", default="xyz", choices=["xyz", "xyzi"], help="point format on the wire: xyz = 12 B + ring, xyzi = 16 B + ring")
    ap.add_argument("--secondary", default="auto", help="second sensor measured inside the same line at N=1 ('auto' = hdl64 when --sensor vlp16, 'none')")
    ap.add_argument("--threads", type=int, default=-1, help="host threads submitting the handles' steps (-1 = one per handle for --workload loop, whose mapping "
                    "calls end with a stream synchronisation, else 1; ctypes releases the GIL inside the C ABI)")
    ap.add_argument("--host-copies", type=int, default=3, help="e2e arm: 3 = the scans are slots of one host arena, neighbouring lanes share a cudaMemcpyAsync (src_on_device 3); "
                    "0 = one cudaMemcpyAsync per lane (ring channel directly behind the points); 2 = packing kernel over page-locked memory")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--profile-json", default="", help="write the per-kernel CUDA-event breakdown to this file")
    return ap.parse_args()


def default_lanes(sensor):
    """Multiples of the 148 SMs: 4 sequences per SM for 16 rings, 2 for 32 and 64, 1 for 128 rings."""
    if sensor in ("vlp16", "vlp16_noring", "os1_16"):
        return 592
    return 296 if sensor in ("hdl32", "hdl64", "os1_64") else 148


def default_scans_per_step(sensor):
    if sensor in ("vlp16", "vlp16_noring", "os1_16"):
        return 25
    return 10 if sensor in ("hdl32", "hdl64", "os1_64") else 5


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


class Window:
    """Which scans of a trajectory are primed / warmed up / timed.  Identical for both arms."""

    def __init__(self, args, sensor, workload="drive"):
        self.S = args.scans_per_step if args.scans_per_step > 0 and sensor == args.sensor else default_scans_per_step(sensor)
        self.W, self.K = args.warmup, args.steps
        if workload == "loop":
            self.steady = 0
        else:
            # local map complete: every key frame within surroundingKeyframeSearchRadius (50 m, utility.h:129) lies behind the vehicle
            self.steady = int(math.ceil(50.0 / (args.speed * 0.1)))
        self.first_timed = max(self.W * self.S, self.steady)
        self.prime = self.first_timed - self.W * self.S
        self.T = self.first_timed + self.K * self.S

    def describe(self):
        return {"scans_per_step": self.S, "scans_per_lane": self.T, "timed_scan_range": [self.first_timed, self.T - 1],
                "timed_scans_per_lane": self.K * self.S, "untimed_scans_per_lane": self.first_timed}


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ sharding
def rank_seeds(rank, world, n_distinct):
    """Sequences are the only axis the path shards on: rank r generates seeds r*1000 .. r*1000+D-1 and its
    lanes replay them (lane l -> sequence l % D); no sequence is shared between ranks, nothing is exchanged on the data path."""
    assert n_distinct <= 1000
    return [1000 * rank + d for d in range(n_distinct)]


def default_distinct(sensor, world, lanes, T, cap):
    """Distinct sequences per rank: 32 per box for the 16-ring sensor (generation time and pinned memory are the limits:
    T scans x cap points x 14 B each), never more than the lanes, at least 4 per rank."""
    per_seq = T * cap * 14
    budget = 10e9 / max(world, 1)
    d = int(max(4, min(32 // max(world, 1), budget // max(per_seq, 1))))
    return max(1, min(d, lanes))


def max_over_ranks(x, world, device=None):
    """Timing reduction of the contract: the slowest rank defines the step time."""
    if world <= 1:
        return float(x)
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(x)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ------------------------------------------------------------------------------------------------ inputs
LOOP_RADIUS = 8.0


def make_sequence(sensor, seed, workload, speed, n_scans):
    import lego_loam_b200  # noqa: F401
    from lego_loam_b200.synth import SynthSequence
    if workload == "loop":
        # closed circle in the room scene: the vehicle re-visits its own key frames every lap (loop closure, config 5)
        return SynthSequence(sensor, seed=seed, scene="room", speed=speed, n_scans=max(n_scans, 64), circle_radius=LOOP_RADIUS)
    return SynthSequence(sensor, seed=seed, scene="street", speed=speed, n_scans=max(n_scans, 64))


# ------------------------------------------------------------------------------------------------ host memory placement
_NUMA = {"node": None, "note": "not tried"}


def numa_place_near_gpu(device_index):
    """Page-locked input buffers should live in the DRAM of the socket the GPU hangs on: with one rank per GPU on a two-socket
    box, round 1's 8-GPU end-to-end run moved ~25 GB/s per GPU where one GPU alone moves 55 (every rank's buffers on one
    socket).  Reads the GPU's NUMA node from sysfs, makes it the preferred node of this thread's allocations
    (set_mempolicy(MPOL_PREFERRED): falls back to other nodes by itself) and moves the thread onto that node's cores when its
    affinity mask allows.  Anything missing (one node, a VM without topology, no permission) leaves the defaults in place.
    Returns what was done, for the bench line."""
    import ctypes
    import platform
    info = {"node": None, "note": ""}
    try:
        import torch
        pr = torch.cuda.get_device_properties(device_index)
        dom, bus, dev = getattr(pr, "pci_domain_id", 0), getattr(pr, "pci_bus_id", None), getattr(pr, "pci_device_id", 0)
        if bus is None:
            info["note"] = "no PCI id from torch"; return info
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        with open(path) as f:
            node = int(f.read().strip())
        with open("/sys/devices/system/node/online") as f:
            online = f.read().strip()
        if node < 0 or online in ("0", ""):
            info["note"] = "single NUMA node (gpu node %d, online %s)" % (node, online); return info
        info["node"] = node
        nr = {"x86_64": 238, "aarch64": 237}.get(platform.machine())
        if nr is not None and node < 64:
            libc = ctypes.CDLL(None, use_errno=True)
            mask = ctypes.c_ulong(1 << node)
            rc = libc.syscall(nr, 1, ctypes.byref(mask), 65)    # MPOL_PREFERRED
            info["mempolicy"] = "preferred node %d" % node if rc == 0 else "set_mempolicy errno %d" % ctypes.get_errno()
        cpus = set()
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            for part in f.read().strip().split(","):
                if part:
                    a, _, b = part.partition("-")
                    cpus.update(range(int(a), int(b or a) + 1))
        mine = os.sched_getaffinity(0) & cpus
        if mine:
            os.sched_setaffinity(0, mine); info["cpus"] = len(mine)
        else:
            info["cpus"] = 0
        info["note"] = "gpu %d on NUMA node %d" % (device_index, node)
    except Exception as e:   # placement is an optimisation: never fail the bench over it
        info["note"] = "left alone: %s" % str(e)[:120]
    return info


def numa_reset():
    import ctypes
    import platform
    nr = {"x86_64": 238, "aarch64": 237}.get(platform.machine())
    if nr is not None and _NUMA.get("mempolicy", "").startswith("preferred"):
        try:
            ctypes.CDLL(None).syscall(nr, 0, None, 0)             # MPOL_DEFAULT
        except OSError:
            pass


class Inputs:
    """Pinned host copy of D sequences x T scans, one slot of cap * (4 fpp + 2) bytes per (scan, sequence) so that one scan of
    all sequences is contiguous: a slot holds the scan's npts points (fpp floats each) and, directly behind them, its ring
    channel (npts uint16) -- the layout with which lego_pipeline_step_lanes moves a lane with one copy."""

    def __init__(self, sensor, seeds, T, workload, speed, fpp, pinned=True):
        import torch
        seqs = [make_sequence(sensor, sd, workload, speed, T) for sd in seeds]
        self.D, self.T, self.fpp = len(seeds), T, fpp
        self.cap = cap = seqs[0].cap
        D = self.D
        self.slot = slot = cap * (4 * fpp + 2)
        t0 = time.perf_counter()
        blob = torch.zeros((T, D, slot), dtype=torch.uint8, pin_memory=pinned)
        self._keep = blob
        self.blob = blob.numpy()
        self.npts = np.zeros((T, D), np.int32)
        # generate into scratch (points | ring at the END of the slot), then slide every scan's ring channel up behind its points
        xyz = self.blob[:, :, :cap * 4 * fpp].view(np.float32)
        ring = self.blob[:, :, cap * 4 * fpp:].view(np.uint16)
        for d, s in enumerate(seqs):
            s.batch_into(0, T, xyz[0, d], ring[0, d], self.npts[0, d:], fpp, D * slot // 4, D, ring_stride=D * slot // 2)
        for t in range(T):
            for d in range(D):
                n = int(self.npts[t, d])
                self.blob[t, d, n * 4 * fpp:n * (4 * fpp + 2)] = self.blob[t, d, cap * 4 * fpp:cap * 4 * fpp + 2 * n].copy()
        self.gen_s = time.perf_counter() - t0
        self.stamps = np.arange(T, dtype=np.float64) * 0.1
        self.scan_bytes = D * slot

    def scan_xyz(self, t, d):
        """(npts, fpp) float32 view of one scan's points"""
        n = int(self.npts[t, d])
        return self.blob[t, d, :n * 4 * self.fpp].view(np.float32).reshape(n, self.fpp)

    def tables(self, base, lanes):
        """per-lane slot addresses of scan 0 (uint64) and the sequence every lane replays; scan t adds t * scan_bytes, the ring
        channel of a lane sits npts * 4 * fpp bytes behind its points"""
        seq = np.arange(lanes, dtype=np.uint64) % np.uint64(self.D)
        return np.uint64(base) + seq * np.uint64(self.slot), seq.astype(np.int64)


# ------------------------------------------------------------------------------------------------ roofline
def algorithmic_bytes(name, c):
    """Compulsory HBM bytes of ONE launch of kernel `name` (every input read once, every output written once;
    formulas of SURVEY.md 8d / DESIGN.md, evaluated with the run's own counts, summed over the lanes of the launch).
    c: dict of per-step totals over lanes (P, N, S, O, Fs, Fl, Ff, Flf, R, C, G, L, ...)."""
    N, P, S, O, L = c["N"], c["P"], c["S"], c["O"], c["L"]
    F3 = c["Fs"] + c["Fl"] + c["Ff"]
    t = {
        "memset_owner": 4 * N,
        "ip_project_kernel": 18 * P + 4 * P,
        "ip_gather_kernel": 4 * N + 16 * P + 21 * N,   # owner in, the winners' points in; fullCloud, rangeMat, groundMat out
        "ip_ground_kernel": 16 * (c["G"] + 1) * c["C"] * L + N,
        "ip_ccl_runs_kernel": 17 * N,       # rangeMat, owner, groundMat in; parent, runLen out
        "ip_ccl_union_kernel": 13 * N,      # rangeMat, owner, groundMat, runLen in; parent updated in place at the roots
        "ip_ccl_stats_kernel": 17 * N,      # parent in/out, runLen, owner, groundMat in
        "ip_ccl_tile_kernel": 14 * N,       # rangeMat, owner, groundMat in; parent, marker out (statistics only at tile roots)
        "ip_ccl_border_kernel": 0,          # a few hundred cells per image
        "ip_ccl_finish_kernel": 9 * N,      # marker, parent in; parent out
        "ip_row_count_kernel": 10 * N,      # owner, groundMat, parent in; cellClass out
        "ip_compact_kernel": 10 * N + 41 * S + 16 * O + 8 * c["R"] * L,
        "fa_prepare_kernel": 20 * S,
        "fa_distort_kernel": 32 * S,
        "fa_sort_sectors_kernel": 13 * S + 4 * S,          # curvature, column, ground flag in; candidate lists out
        "fa_select_kernel": 4 * S + 8 * S + 32 * F3,       # candidate lists, cloudNeighborPicked in/out; picked points in/out
        "fa_voxel_ring_kernel": 4 * S + 16 * S + 16 * c["Flf"],   # cloudLabel, points in; less-flat cloud out
        "fa_gather_features_kernel": 32 * (F3 + c["Flf"]),
        "fa_odometry_kernel": c.get("odo_bytes", 0),
        "fa_clouds_last_kernel": 32 * (c["Fl"] + c["Flf"]) + 32 * O,
        "fa_finish_scan_kernel": 64 * L,
        "vh_build(cornerLast)": 36 * c["Fl"],
        "vh_build(surfLast)": 36 * c["Flf"],
    }
    t.update(c.get("mo_bytes", {}))
    return float(t.get(name, 0))


def gather_counts(pl, params, lanes):
    """Per-step workload counts summed over lanes, read back from the device counters after the run."""
    c = dict(N=params.N_SCAN * params.Horizon_SCAN * lanes, R=params.N_SCAN, C=params.Horizon_SCAN, G=params.groundScanInd, L=lanes,
             P=0, S=0, O=0, Fs=0, Fl=0, Ff=0, Flf=0)
    odo = 0
    for l in range(lanes):
        ipc = pl.download("ipCounts", l)
        fc = pl.download("featCounts", l)
        lc = pl.download("lastCounts", l)
        st = pl.download("lmState", l)
        c["S"] += int(ipc[0]); c["O"] += int(ipc[1])
        c["Fs"] += int(fc[0]); c["Fl"] += int(fc[1]); c["Ff"] += int(fc[2]); c["Flf"] += int(fc[3])
        # K10 + K11 of SURVEY.md 8d: per correspondence refresh 24*Q + 16*M, per LM iteration 96 bytes per row
        it_s, it_c = max(int(st[6]), 1), max(int(st[7]), 1)
        odo += ((it_s + 4) // 5) * (24 * int(fc[2]) + 16 * int(lc[1])) + it_s * 96 * int(fc[2])
        odo += ((it_c + 4) // 5) * (24 * int(fc[0]) + 16 * int(lc[0])) + it_c * 96 * int(fc[0])
    c["odo_bytes"] = odo
    # mapOptimization (counts of the last mapping call of every lane; K13..K18 of SURVEY.md 8d)
    mc = np.zeros(4, np.int64); sc = np.zeros(8, np.int64); s2m = 0; rerank = 0; search = 0
    for l in range(lanes):
        m_, s_, st = pl.download("mapCounts", l), pl.download("scanCounts", l), pl.download("moState", l)
        mc += m_; sc += s_
        s2m += int(st[3]) * (128 * (int(s_[0]) + int(s_[4])) + 32 * int(st[4]))
        # the 5-NN of an LM iteration: iteration 0 searches every query (query 16 + neighbours 80 + candidate cache 48 + result 32),
        # iterations >= 1 re-rank the cached candidates (query 16 + cache 48 + 8 candidates 128 + result 32); the searches of
        # the queries whose cache could not prove the answer are not counted (lower bound)
        search += 176 * (int(s_[0]) + int(s_[4]))
        rerank += max(int(st[3]) - 1, 0) * 224 * (int(s_[0]) + int(s_[4]))
    arr, n_map, n_ds = int(sc[0] + sc[1] + sc[2]), int(mc[0] + mc[1]), int(mc[2] + mc[3])
    c["mo_bytes"] = {
        "mo_begin_kernel": 200 * lanes,
        # incremental local map (csrc/incremental_map.cu): n = points of the sorted local map, arr = points of the key frames that
        # arrived in this call (about one key frame = this scan's down-sampled clouds), nDS = points of the down-sampled map
        "mo_assemble_map_kernel": 32 * arr,
        "im_sort_arrivals": 20 * arr,
        "im_gather_new_kernel": 48 * arr,
        "im_mark_kernel": 4 * n_map,
        "im_merge_old_kernel": 40 * n_map,
        "im_merge_new_kernel": 48 * arr,
        "im_commit_kernel": 0.0,
        "im_reduce_kernel": 16 * n_map + 16 * n_ds,
        "im_compact_kernel": 32 * n_ds,
        "vg_filter(localMap)": 16 * int(mc.sum()),        # only with LEGO_INCR_MAP=0 (whole map assembled and filtered per call)
        "vg_filter(currentScan)": 16 * (c["Fl"] + c["Flf"] + c["O"] + int(sc[0] + sc[1] + sc[2])),
        "mo_concat_kernel": 32 * int(sc[3]),
        "vg_filter(surfTotal)": 16 * int(sc[3] + sc[4]),
        "vh_build(cornerMap)": 36 * int(mc[2]),
        "vh_build(surfMap)": 36 * int(mc[3]),
        "mo_lm_begin_kernel": 40 * lanes,
        "mo_rerank_kernel": rerank / 9.0,   # averages over the launches of a mapping call (converged lanes do nothing)
        "mo_search_kernel": search / 10.0,
        "mo_residual_kernel": s2m / 10.0,
        "mo_solve_kernel": 0.0,
        "mo_lm_persistent_kernel": rerank + search + s2m,   # one launch = all iterations of a mapping call
        "mo_save_keyframe_kernel": 32 * int(sc[0] + sc[1] + sc[2]),
    }
    c["map"] = {"cornerFromMap": int(mc[0]), "surfFromMap": int(mc[1]), "cornerFromMapDS": int(mc[2]), "surfFromMapDS": int(mc[3]),
                "cornerLastDS": int(sc[0]), "surfLastDS": int(sc[1]), "outlierLastDS": int(sc[2]), "surfTotalLastDS": int(sc[4])}
    return c


def ncu_traffic(kernel, lanes, sensor):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/ncu_traffic.json), scaled to
    the lanes of one launch.  None when there is no capture of this kernel for this sensor at the steady-state local map
    (entries carry "sensor" and, for the kernels whose traffic grows with the map, "steady": true)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            e = json.load(f).get(kernel)
        if not e or e.get("sensor", "vlp16") != sensor:
            return None
        map_dependent = kernel.startswith("mo_") or "Map" in kernel
        if map_dependent and not e.get("steady", False):
            return None
        return float(e["bytes_per_lane"]) * lanes
    except (OSError, ValueError, KeyError):
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def build_roofline(prof_h0, prof_all, counts_h0, counts_all, dom_timed, lanes_h0, sensor, step_ms, scans_profiled, scans_per_step):
    """roofline of the dominant kernel from ONE handle's launches (bytes, traffic and launch time all describe the lanes_h0
    lanes one launch covers) + the whole pipeline's algorithmic bytes per step against the step time."""
    peak, peak_src = load_peaks()
    tot_ms = sum(v[0] for v in prof_h0.values()) or 1.0
    dominant = dom_timed["kernel"]
    abytes = algorithmic_bytes(dominant, counts_h0)
    avg_ms = dom_timed["ms"] / max(dom_timed["n"], 1)
    achieved = abytes / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    traffic = ncu_traffic(dominant, lanes_h0, sensor)
    frac = achieved / peak
    assert frac <= 1.0, "roofline.frac > 1: bytes and launch time do not describe the same lanes"
    # pipeline aggregate: sum over kernels and handles of (launches in the profiled step) x (algorithmic bytes per launch)
    pipe_bytes = 0.0
    for prof, counts in zip(prof_all, counts_all):
        for k, (_, n) in prof.items():
            pipe_bytes += n * algorithmic_bytes(k, counts)
    pipe_bytes *= scans_per_step / float(max(scans_profiled, 1))
    pipe_gbs = pipe_bytes / (step_ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": frac,
            "traffic": traffic, "peak_source": peak_src, "lanes_per_launch": lanes_h0,
            "algorithmic_bytes_per_launch": abytes, "avg_launch_ms": avg_ms,
            "share_of_step": prof_h0[dominant][0] / tot_ms if dominant in prof_h0 else None,
            "pipeline": {"algorithmic_bytes_per_step": pipe_bytes, "achieved": pipe_gbs, "frac": pipe_gbs / peak,
                         "note": "sum over all kernels and handles of launches x algorithmic bytes per launch in one step, / ms_per_step; the profile is flat, "
                                 "so this, not the dominant kernel, says how far the whole step is from the HBM roof"}}


# ------------------------------------------------------------------------------------------------ B200 arm
def run_b200(args, sensor, lanes, workload):
    """One sensor through both arms (device-resident `value`, host-buffer `e2e`) on this rank's GPU."""
    import torch
    import torch.distributed as dist
    import lego_loam_b200  # noqa: F401
    from lego_loam_b200.capi import Pipeline
    from lego_loam_b200.params import make_params

    rank, world, local = dist_env()
    win = Window(args, sensor, workload)
    S, W, K = win.S, win.W, win.K
    L = lanes
    H = max(1, min(args.handles, L))
    fpp = 3 if args.input == "xyz" else 4
    # 2000-scan sequences save ~500 key frames (one per mapping call at 4 m/s): size the pool for the run, with head room
    params = make_params(sensor, device=local, loop_closure=(workload == "loop"), max_keyframes=max(640, win.T // 3 + 64))
    phase = [(h if args.stagger else 0) for h in range(H)]
    Ttot = win.T + max(phase)
    cap0 = params.N_SCAN * params.Horizon_SCAN
    D = args.distinct if args.distinct > 0 else default_distinct(sensor, world, L, Ttot, cap0)
    _NUMA.update(numa_place_near_gpu(local))
    inp = Inputs(sensor, rank_seeds(rank, world, D), Ttot, workload, args.speed, fpp)
    numa_reset()
    cap = inp.cap
    bounds = [(h * L) // H for h in range(H + 1)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def make_group():
        return [Pipeline(params, lanes=bounds[h + 1] - bounds[h], max_points=cap) for h in range(H)]

    def close_group(group):
        for pl in group:
            pl.close()

    def make_scan_fn(base, on_device):
        xt, seq = inp.tables(base, L)
        psz = np.uint64(4 * fpp)

        def scan(h, pl, s):
            l0, l1 = bounds[h], bounds[h + 1]
            n = inp.npts[s][seq[l0:l1]]
            px = xt[l0:l1] + np.uint64(s * inp.scan_bytes)
            pl.step_lanes(px, px + n.astype(np.uint64) * psz, n, np.full(l1 - l0, inp.stamps[s]), floats_per_point=fpp,
                          stage_mask=args.stages, on_device=on_device)
        return scan, seq

    # ---- device-resident inputs (value) -------------------------------------------------------------
    # The lanes are split over `handles` pipeline handles = CUDA streams: while one handle sits in its latency-bound LM
    # kernels (one CTA per lane) the others' streaming kernels fill the rest of the GPU.
    d_x = torch.from_numpy(inp.blob).cuda()
    torch.cuda.synchronize()
    dev_scan, seq = make_scan_fn(d_x.data_ptr(), True)
    group = make_group()

    n_threads = args.threads if args.threads >= 0 else (H if workload == "loop" else 1)
    pool = None
    if n_threads > 1:
        from concurrent.futures import ThreadPoolExecutor
        pool = ThreadPoolExecutor(max_workers=min(n_threads, H))

    def run_scans(scan_fn, s0, s1):
        """scans s0 .. s1-1 for every handle.  With a thread pool every handle is driven by its own host thread (one handle = one
        host thread is the C ABI's rule), so a handle that blocks -- loop-closure mode synchronises its stream in every mapping
        call -- does not keep the others from submitting."""
        if pool is None:
            for s in range(s0, s1):
                for h, pl in enumerate(group):
                    scan_fn(h, pl, s + phase[h])
            return

        def drive(h, pl):
            for s in range(s0, s1):
                scan_fn(h, pl, s + phase[h])
        for f in [pool.submit(drive, h, pl) for h, pl in enumerate(group)]:
            f.result()

    for h, pl in enumerate(group):           # handle h runs phase[h] scans ahead
        for s in range(phase[h]):
            dev_scan(h, pl, s)
    run_scans(dev_scan, 0, win.prime + max(W - 1, 0) * S)
    # last warm-up step: every kernel bracketed with events -> per-kernel shares, launches per step, the dominant kernel
    prof_all = [{} for _ in group]
    scans_profiled = 0
    if W >= 1:
        for pl in group:
            pl.sync()
        for pl in group:
            pl.profile(1)
        run_scans(dev_scan, win.first_timed - S, win.first_timed)
        scans_profiled = S
        prof_all = [pl.profile_read() for pl in group]
        for pl in group:
            pl.profile(0)
    prof_sum = {}
    for prof in prof_all:
        for k, (ms_k, n_k) in prof.items():
            a = prof_sum.get(k, (0.0, 0))
            prof_sum[k] = (a[0] + ms_k, a[1] + n_k)
    dominant = max(prof_sum, key=lambda k: prof_sum[k][0]) if prof_sum else "fa_odometry_kernel"
    for pl in group:
        pl.sync()
        pl.profile(2, dominant)
    launches0 = sum(pl.launch_count() for pl in group)
    clk = ClockSampler(local)
    clk.start()
    barrier()
    for pl in group:
        pl.timer_record(0)
    t_host0 = time.perf_counter()
    submit_scan_ms = []      # host time to enqueue one scan of every handle; the first ones find the launch queues empty
    if pool is None:
        for s in range(win.first_timed, win.T):
            t_s = time.perf_counter()
            run_scans(dev_scan, s, s + 1)
            submit_scan_ms.append((time.perf_counter() - t_s) * 1e3)
    else:
        run_scans(dev_scan, win.first_timed, win.T)       # every handle's thread runs through the whole timed window on its own
        submit_scan_ms.append((time.perf_counter() - t_host0) * 1e3 / max(win.T - win.first_timed, 1))
    host_submit_s = time.perf_counter() - t_host0
    for pl in group:
        pl.timer_record(1)
    for pl in group:
        pl.sync()
    barrier()
    ms = max(group[0].timer_elapsed_to(0, pl, 1) for pl in group)
    clocks = clk.stop()
    launches = sum(pl.launch_count() for pl in group) - launches0
    d_ms, d_n = group[0].profile_read().get(dominant, (0.0, 0))
    for pl in group:
        pl.profile(0)
    dom_timed = {"kernel": dominant, "ms": d_ms, "n": d_n}
    counts_all = []
    for h, pl in enumerate(group):
        ch = gather_counts(pl, params, bounds[h + 1] - bounds[h])
        ch["P"] = int(inp.npts[win.T - 1][seq[bounds[h]:bounds[h + 1]]].sum())
        counts_all.append(ch)
    kf_total, kf_in_map = [], []
    for pl in group:
        for l in range(0, pl.lanes, max(1, pl.lanes // 16)):     # a sample of the lanes
            kf_total.append(int(pl.download("moState", l)[1]))
            kf_in_map.append(len(pl.download("surroundingExistingKeyPosesID", l)))
    poses_dev = np.concatenate([pl.poses()[0] for pl in group])
    close_group(group)
    del d_x
    torch.cuda.empty_cache()
    ms = max_over_ranks(ms, world, "cuda")
    # ---- host buffers through the C ABI (e2e) --------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        group = make_group()
        # the scans sit in page-locked host memory, every lane's ring channel directly behind its points: one cudaMemcpyAsync per
        # lane (input mode 0).  --host-copies 2 = input mode 2, one packing kernel per scan pulls the lanes over PCIe (measured
        # slower: SM-initiated PCIe reads reach ~23 GB/s on this box, the copy engine ~2x that with one copy per lane)
        # at 2-3 us of copy-engine time per copy, 592 copies per scan cap the link at ~34 GB/s).  --host-copies 3 (default) = input
        # mode 3: the caller declares the lane blocks slots of one arena, lanes in neighbouring slots travel in one copy with the
        # unused tail of each slot (counted in h2d_bytes_per_step)
        host_scan, _ = make_scan_fn(inp.blob.ctypes.data, args.host_copies if args.host_copies in (2, 3) else 0)
        # Streaming use of the C ABI: scan t is submitted from pinned host memory (one H2D copy per lane of exactly its
        # points on the handle's copy stream), its pose read-back is enqueued behind it, and the result of scan t-4 is
        # collected while scans t-3 .. t are in flight (a mapping scan lasts as long as several uploads, so less slack
        # leaves the PCIe link idle behind it).  Every scan's inputs cross PCIe and every scan's poses are read on the host
        # inside the timed region.
        DEPTH = 5   # scans in flight per handle: the poses of scan k are collected after scan k+4 was submitted

        def submit(s):
            def one(h, pl):
                host_scan(h, pl, s + phase[h])
                pl.poses_async(s % 8)
            if pool is None:
                for h, pl in enumerate(group):
                    one(h, pl)
            else:
                for f in [pool.submit(one, h, pl) for h, pl in enumerate(group)]:
                    f.result()

        def collect(s):
            return np.concatenate([pl.poses_wait(s % 8)[0] for pl in group])

        for h, pl in enumerate(group):
            for s in range(phase[h]):
                host_scan(h, pl, s)
        for s in range(win.first_timed):
            submit(s)
            if s - (DEPTH - 1) >= 0:
                collect(s - (DEPTH - 1))
        for s in range(max(0, win.first_timed - (DEPTH - 1)), win.first_timed):
            collect(s)
        barrier()
        t0 = time.perf_counter()
        poses_host = None
        for s in range(win.first_timed, win.T):
            submit(s)
            if s - (DEPTH - 1) >= win.first_timed:
                poses_host = collect(s - (DEPTH - 1))
        for s in range(max(win.first_timed, win.T - (DEPTH - 1)), win.T):
            poses_host = collect(s)
        barrier()
        e2e_s = time.perf_counter() - t0
        close_group(group)
        e2e_s = max_over_ranks(e2e_s, world, "cuda")
        same = bool(np.allclose(poses_host, poses_dev, atol=1e-6))
        assert same, "host-buffer and device-buffer runs disagree"
        # what the PCIe link of this box gives a plain pinned->device copy (context for E: the streaming loop cannot beat
        # h2d_bytes_per_step / this rate)
        n_probe = max(1, min(inp.T, int(256e6 // max(inp.scan_bytes, 1))))
        src = torch.from_numpy(inp.blob[:n_probe])
        probe = torch.empty(src.shape, dtype=torch.uint8, device="cuda")
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        probe.copy_(src, non_blocking=True); torch.cuda.synchronize()
        ev0.record()
        for _ in range(3):
            probe.copy_(src, non_blocking=True)
        ev1.record(); torch.cuda.synchronize()
        h2d_gbs = 3 * src.numel() / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
        del probe
        pts_timed = int(sum(int(inp.npts[s][seq].sum()) for s in range(win.first_timed, win.T)))
        moved = pts_timed * (fpp * 4 + 2)
        copies_per_scan = L
        if args.host_copies == 3:
            # a run = lanes of one handle in consecutive slots; it moves whole slots except for its last lane
            last = np.ones(L, bool)
            last[:-1] = seq[1:] != seq[:-1] + 1
            last[np.asarray(bounds[1:], np.int64) - 1] = True
            copies_per_scan = int(last.sum())
            moved = int(sum(int(inp.npts[s][seq[last]].sum()) for s in range(win.first_timed, win.T))) * (fpp * 4 + 2) \
                + int((~last).sum()) * inp.slot * (win.T - win.first_timed)
        e2e = {"value": L * K * S * world / e2e_s, "unit": UNIT,
               "h2d_bytes_per_step": int(moved / K + S * L * (16 + 4)),
               "h2d_copies_per_scan": copies_per_scan,
               "d2h_bytes_per_step": int(S * L * 19 * 4),
               "h2d_bytes_per_scan": round(moved / float(K * S * L), 1),
               "input_format": "x y z float32 + ring uint16 (14 B/point), only the npts points of each lane" if fpp == 3 else
                               "x y z intensity float32 + ring uint16 (18 B/point), only the npts points of each lane",
               "h2d_path": "packing kernel reading page-locked host memory (lego_pipeline_step_lanes, src_on_device = 2)" if args.host_copies == 2 else
                           "the lanes' scans are fixed-size slots of one pinned arena, ring channel directly behind the points: lanes in neighbouring slots "
                           "share one cudaMemcpyAsync, unused slot tails included in the byte count (lego_pipeline_step_lanes, src_on_device = 3)"
                           if args.host_copies == 3 else
                           "one cudaMemcpyAsync per lane: the ring channel sits directly behind the lane's points (lego_pipeline_step_lanes, src_on_device = 0)",
               "h2d_gbs_achieved": round(moved / e2e_s / 1e9, 1),
               "seconds": e2e_s, "pcie_h2d_gbs_plain_copy": round(h2d_gbs, 1), "poses_equal_device_arm": same,
               "host_memory": dict(_NUMA),
               "mode": "streaming: up to 5 scans in flight per handle, the poses of scan k are collected after scan k+4 was submitted"}

    prof_h0 = prof_all[0]
    step_ms = ms / K
    roofline = build_roofline(prof_h0, prof_all, counts_all[0], counts_all, dom_timed, bounds[1] - bounds[0], sensor, step_ms, scans_profiled, S) \
        if prof_sum else None
    cmap = {k: sum(c["map"][k] for c in counts_all) / float(L) for k in counts_all[0]["map"]}
    local_map = {"keyframes_saved": float(np.mean(kf_total)), "keyframes_in_local_map": float(np.mean(kf_in_map)),
                 "cornerFromMap": cmap["cornerFromMap"], "surfFromMap": cmap["surfFromMap"],
                 "cornerFromMapDS": cmap["cornerFromMapDS"], "surfFromMapDS": cmap["surfFromMapDS"],
                 "note": "mean per lane at the last mapping call of the run"}
    return dict(rank=rank, world=world, ms=ms, L=L, K=K, W=W, S=S, win=win, e2e=e2e, clocks=clocks, launches=launches,
                prof_all=prof_all, prof_sum=prof_sum, counts_all=counts_all, roofline=roofline, cap=cap, params=params, D=D,
                host_submit_ms=host_submit_s * 1e3 / K, submit_scan_ms=submit_scan_ms, local_map=local_map, gen_s=inp.gen_s, fpp=fpp, H=H,
                mean_points=float(inp.npts[win.first_timed:win.T].mean()), inputs=inp, threads=(min(n_threads, H) if pool else 1))


def measure_latency(args, sensor, workload, inp, win, device):
    """One lane alone through the full pipeline, host buffers in, poses out, one scan at a time (the reference's own
    shape: one callback per scan at 10 Hz, imageProjection.cpp:181-197): wall ms from the call to the poses on the host."""
    import lego_loam_b200  # noqa: F401
    from lego_loam_b200.capi import Pipeline
    from lego_loam_b200.params import make_params
    params = make_params(sensor, device=device, loop_closure=(workload == "loop"), max_keyframes=max(640, win.T // 3 + 64))
    pl = Pipeline(params, lanes=1, max_points=inp.cap)
    xt, _ = inp.tables(inp.blob.ctypes.data, 1)
    n_timed = min(200, win.T - win.first_timed)
    lat, mapped = [], []
    for s in range(win.first_timed + n_timed):
        t0 = time.perf_counter()
        px = xt + np.uint64(s * inp.scan_bytes)
        pl.step_lanes(px, px + np.uint64(int(inp.npts[s][0]) * 4 * inp.fpp), inp.npts[s][:1],
                      inp.stamps[s:s + 1], floats_per_point=inp.fpp, stage_mask=args.stages)
        _, flags = pl.poses()
        dt = (time.perf_counter() - t0) * 1e3
        if s >= win.first_timed:
            lat.append(dt); mapped.append(bool(flags[0] & 8))
    launches = pl.launch_count()
    graph_launches = pl.graph_launch_count()
    pl.close()
    lat = np.array(lat); mapped = np.array(mapped)
    out = {"lanes": 1, "scans": int(len(lat)), "scan_range": [win.first_timed, win.first_timed + n_timed - 1],
           "ms_per_scan": {"mean": float(lat.mean()), "p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)), "max": float(lat.max())},
           "launches_per_scan": launches / float(win.first_timed + n_timed),
           "graph_launches_per_scan": graph_launches / float(win.first_timed + n_timed),
           "submission": ("CUDA graphs: the front end and the mapping body of a scan are one cudaGraphLaunch each (csrc/engine.cu runGraphed; "
                          "launches_per_scan counts the kernels inside them)") if graph_launches else "plain kernel launches (LEGO_GRAPH=0)",
           "what": "wall time of lego_pipeline_step_lanes (pinned host scan in) + lego_pipeline_get_poses (poses on the host), one scan at a time"}
    if mapped.any() and (~mapped).any():
        out["ms_per_scan"]["p50_mapping_scan"] = float(np.percentile(lat[mapped], 50))
        out["ms_per_scan"]["p50_other_scan"] = float(np.percentile(lat[~mapped], 50))
    return out


# ------------------------------------------------------------------------------------------------ CPU reference
def ref_cfg(sensor, workload):
    return sensor + "_loop" if workload == "loop" else sensor


def _ref_worker(q, sensor, workload, seed, n_scans, first_timed, speed, stages, pin_core):
    """One process = one sequence through the reference's own code (oracle/_ref): scans 0 .. n_scans-1 are fed, the scans
    from first_timed on are timed (scan generation happens in chunks outside the clock)."""
    try:
        if pin_core is not None and hasattr(os, "sched_setaffinity"):
            os.sched_setaffinity(0, {pin_core})
        import refpy
        seq = make_sequence(sensor, seed, workload, speed, n_scans)
        ref = refpy.RefPipeline(ref_cfg(sensor, workload), snapshots=False)
        dt, CH = 0.0, 25
        t_stage0 = None
        for k0 in range(0, n_scans, CH):
            cnt = min(CH, n_scans - k0)
            x, r, n = seq.batch(k0, cnt, threads=1)
            for j in range(cnt):
                k = k0 + j
                if k == first_timed:
                    t_stage0 = ref.timers()
                t0 = time.perf_counter()
                ref.feed(x[j, :n[j]], r[j, :n[j]], k * 0.1, stages)
                if k >= first_timed:
                    dt += time.perf_counter() - t0
        timers, scans, mo_calls = ref.timers()
        if t_stage0 is not None:
            timers = {k: v - t_stage0[0].get(k, 0.0) for k, v in timers.items()}
            scans -= t_stage0[1]; mo_calls -= t_stage0[2]
        ref.close()
        q.put((n_scans - first_timed, dt, {k: v / max(scans, 1) for k, v in timers.items()}, mo_calls))
    except Exception as ex:  # noqa: BLE001
        q.put((0, 0.0, {"error": repr(ex)}, 0))


def run_reference_cpu(sensor, workload, n_procs, n_scans, first_timed, speed, stages, seed0=0, pin=False):
    """-> (timed scans/s over all processes, per-stage ms per scan of process 0 over the timed scans, its mapping calls,
    the slowest process's timed seconds)"""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ref_worker, args=(q, sensor, workload, seed0 + i, n_scans, first_timed, speed, stages, i if pin else None))
             for i in range(n_procs)]
    for p in procs:
        p.start()
    res = [q.get() for _ in procs]
    for p in procs:
        p.join()
    errs = [r[2] for r in res if r[0] == 0]
    if errs:
        raise RuntimeError("reference worker failed: %s" % errs[0])
    total = sum(r[0] for r in res)
    slowest = max(r[1] for r in res) or 1e-9
    return total / slowest, res[0][2], res[0][3], slowest


def reference_available(sensor, workload="drive"):
    import refpy
    return refpy.available(ref_cfg(sensor, workload))


def workload_name(args, sensor, workload, win):
    st = "+".join(n for b, n in ((1, "imageProjection"), (2, "featureAssociation"), (4, "mapOptimization")) if args.stages & b)
    if workload == "loop":
        return "%s synthetic closed-circle trajectories (radius %.0f m, %.1f m/s) with loopClosureEnableFlag, %s incl. loop-closure ICP + pose graph, " \
               "%d-scan sequences, scans %d..%d timed" % (SHAPES[sensor], LOOP_RADIUS, args.speed, st, win.T, win.first_timed, win.T - 1)
    return "%s synthetic street trajectories (%.1f m/s), %s, %d-scan sequences, scans %d..%d timed (local map at its steady-state size: " \
           "every key frame within 50 m)" % (SHAPES[sensor], args.speed, st, win.T, win.first_timed, win.T - 1)


def cpu_baseline_block(args, sensor, workload, win):
    rate, stages_ms, mo_calls, secs = run_reference_cpu(sensor, workload, 1, win.T, win.first_timed, args.speed, args.stages, pin=False)
    return {"value": rate, "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": "one sequence (seed 0), scans 0..%d fed, scans %d..%d timed (%.1f s): the same scan range as the GPU arm; "
                      "reference sources compiled verbatim (oracle/_ref)" % (win.T - 1, win.first_timed, win.T - 1, secs),
            "ms_per_scan": 1000.0 / rate if rate > 0 else None, "mapping_calls": int(mo_calls),
            "per_stage_ms": {k: round(v, 4) for k, v in stages_ms.items() if isinstance(v, float)}}


def run_reference_arm(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return None
    cores = os.cpu_count() or 1
    win = Window(args, args.sensor, args.workload)
    rate, stages_ms, mo_calls, secs = run_reference_cpu(args.sensor, args.workload, cores, win.T, win.first_timed, args.speed, args.stages)
    cfg = {"workload": workload_name(args, args.sensor, args.workload, win), "sensor": args.sensor, "stages": args.stages,
           "note": "reference sources compiled verbatim (oracle/_ref), one independent sequence per host core, every process feeds scans 0..T-1 "
                   "and times exactly the scans the GPU arm times"}
    cfg.update(win.describe())
    return {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * secs / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "%d processes x %d timed scans (scans %d..%d of a %d-scan sequence each), all stages"
                                   % (cores, win.T - win.first_timed, win.first_timed, win.T - 1, win.T),
                         "per_stage_ms": {k: round(v, 4) for k, v in stages_ms.items() if isinstance(v, float)}},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


# ------------------------------------------------------------------------------------------------ main
def result_block(args, r, sensor, workload):
    """the measured numbers of one sensor as they appear in the JSON line (top level for the primary, under
    `secondary.<sensor>` for the second one)"""
    L, K, S, world = r["L"], r["K"], r["S"], r["world"]
    win = r["win"]
    cfg = {"workload": workload_name(args, sensor, workload, win), "sensor": sensor, "lanes_per_gpu": L, "handles_per_gpu": r["H"],
           "handle_stagger_scans": args.stagger, "host_threads": r["threads"], "stages": args.stages,
           "distinct_sequences_per_gpu": r["D"],
           "lane_inputs": "lane l replays generated sequence l %% %d of its GPU (own state per lane; own H2D copy per lane in the e2e arm)" % r["D"],
           "points_per_scan": round(r["mean_points"], 1), "input_format": "xyz+ring (14 B/point)" if r["fpp"] == 3 else "xyzi+ring (18 B/point)",
           "local_map": r["local_map"],
           "l2": "one scan of all lanes reads %.0f MB of input and touches %.0f GB of per-lane state: far beyond the 126 MB L2 (no flush needed)"
                 % (L * r["mean_points"] * (r["fpp"] * 4 + 2) / 1e6, L * 0.115 if sensor.startswith("vlp16") else L * 0.3)}
    cfg.update(win.describe())
    blk = {"value": L * K * S * world / (r["ms"] * 1e-3), "unit": UNIT, "ms_per_step": r["ms"] / K, "timed_region_s": r["ms"] * 1e-3,
           "config": cfg, "roofline": r["roofline"], "clocks": r["clocks"], "gpu_launches": int(r["launches"]),
           "gpu_launches_per_scan_step": round(r["launches"] / float(K * S), 1),
           "host_submit_ms_per_step": round(r["host_submit_ms"], 3),
           "host_submit_ms_per_scan": {"first_two_scans_queue_empty": [round(v, 3) for v in r["submit_scan_ms"][:2]],
                                       "min": round(min(r["submit_scan_ms"]), 3), "median": round(float(np.median(r["submit_scan_ms"])), 3),
                                       "note": "one scan of all lanes = every handle's launches; once the launch queues are full the host blocks and tracks the device"},
           "input_generation_s": round(r["gen_s"], 1)}
    if r["e2e"]:
        blk["e2e"] = r["e2e"]
    return blk


def main():
    args = parse_args()
    rank, world, local = dist_env()
    if args.impl == "reference":
        if not reference_available(args.sensor, args.workload):
            # the oracle restatement always exists in the repo; only the verbatim-compiled .so can be missing
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/liblego_ref_%s.so not built" % ref_cfg(args.sensor, args.workload)}))
            return
        line = run_reference_arm(args)
        if line is not None:
            print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    lanes = args.lanes if args.lanes > 0 else default_lanes(args.sensor)
    r = run_b200(args, args.sensor, lanes, args.workload)
    blk = result_block(args, r, args.sensor, args.workload)
    line = {"metric": METRIC, "value": blk["value"], "unit": UNIT, "n_gpus": world, "steps": r["K"], "warmup": r["W"],
            "ms_per_step": blk["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic"}
    for k in ("config", "roofline", "clocks", "gpu_launches", "gpu_launches_per_scan_step", "host_submit_ms_per_step", "host_submit_ms_per_scan", "timed_region_s",
              "input_generation_s", "e2e"):
        if k in blk:
            line[k] = blk[k]
    if rank == 0 and world == 1 and not args.no_latency:
        line["latency"] = measure_latency(args, args.sensor, args.workload, r["inputs"], r["win"], local)
    profile_dump = None
    if args.profile_json and rank == 0:
        profile_dump = {"per_kernel_ms_over_profiled_step_handle0": {k: {"ms": v[0], "launches": v[1], "algorithmic_bytes_per_launch": algorithmic_bytes(k, r["counts_all"][0])}
                                                                      for k, v in r["prof_all"][0].items()},
                        "per_kernel_ms_all_handles": {k: {"ms": v[0], "launches": v[1]} for k, v in r["prof_sum"].items()},
                        "profiled_scans": r["S"], "lanes_handle0": r["L"] // r["H"],
                        "counts_handle0": {k: v for k, v in r["counts_all"][0].items() if k != "mo_bytes"}}
    r["inputs"] = None
    win = r["win"]
    if rank == 0 and not args.no_cpu_baseline and reference_available(args.sensor, args.workload):
        line["cpu_baseline"] = cpu_baseline_block(args, args.sensor, args.workload, win)
        if "latency" in line:
            line["latency"]["reference_ms_per_scan_one_core"] = line["cpu_baseline"]["ms_per_scan"]

    # second sensor inside the same line (north-star target shape 64 x 2083), N = 1 only
    sec = args.secondary
    if sec == "auto":
        sec = "hdl64" if (args.sensor == "vlp16" and args.workload == "drive") else "none"
    if sec != "none" and world == 1:
        import gc
        del r
        gc.collect()
        torch.cuda.empty_cache()
        r2 = run_b200(args, sec, default_lanes(sec), "drive")
        b2 = result_block(args, r2, sec, "drive")
        win2 = r2["win"]
        del r2
        gc.collect()
        if not args.no_cpu_baseline and reference_available(sec):
            b2["cpu_baseline"] = cpu_baseline_block(args, sec, "drive", win2)
            b2["vs_one_core"] = {"device": b2["value"] / b2["cpu_baseline"]["value"],
                                 "e2e": (b2["e2e"]["value"] / b2["cpu_baseline"]["value"]) if "e2e" in b2 else None,
                                 "north_star_target": ">= 100x the reference's single-core rate at 64 x 2083"}
        line["secondary"] = {sec: b2}
    if rank != 0:
        return
    if profile_dump is not None:
        profile_dump["line"] = line
        with open(args.profile_json, "w") as f:
            json.dump(profile_dump, f, indent=1)
    print(json.dumps(line))


def _shutdown_process_group():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()
    except Exception:   # noqa: BLE001  (best effort at exit)
        pass


if __name__ == "__main__":
    try:
        main()
    finally:
        sys.stdout.flush()
        _shutdown_process_group()
