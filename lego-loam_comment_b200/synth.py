"""ctypes wrapper around tools/liblego_synth.so (deterministic synthetic scans)."""
import ctypes
import os
import subprocess

import numpy as np

from .params import SENSOR_GEOMETRY, make_params

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "tools", "liblego_synth.so")
_SRC = os.path.join(_ROOT, "tools", "synth_scan.cpp")


class SynthCfg(ctypes.Structure):
    _fields_ = [("n_scan", ctypes.c_int), ("horizon_scan", ctypes.c_int),
                ("elev_bottom_deg", ctypes.c_double), ("elev_step_deg", ctypes.c_double),
                ("ang_res_x_deg", ctypes.c_double), ("sensor_height", ctypes.c_double),
                ("scene", ctypes.c_int),
                ("speed", ctypes.c_double), ("yaw_amp", ctypes.c_double), ("yaw_period", ctypes.c_double), ("sway_amp", ctypes.c_double),
                ("noise_sigma", ctypes.c_double), ("max_range", ctypes.c_double), ("min_range", ctypes.c_double),
                ("dropout_prob", ctypes.c_double), ("dup_col_prob", ctypes.c_double), ("jitter_frac", ctypes.c_double),
                ("scan_period", ctypes.c_double),
                ("traj", ctypes.c_int), ("circle_radius", ctypes.c_double)]


def build_synth(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(["g++", "-std=c++14", "-O2", "-fPIC", "-shared", _SRC, "-o", _SO, "-lpthread"])
    return _SO


_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(build_synth())
        lib.lego_synth_create.restype = ctypes.c_void_p
        lib.lego_synth_create.argtypes = [ctypes.POINTER(SynthCfg), ctypes.c_uint64, ctypes.c_double]
        lib.lego_synth_destroy.argtypes = [ctypes.c_void_p]
        lib.lego_synth_scan.restype = ctypes.c_int
        lib.lego_synth_scan.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        lib.lego_synth_batch.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.lego_synth_batch_fmt.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_int]
        lib.lego_synth_batch_strided.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                 ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_long, ctypes.c_int]
        lib.lego_synth_pose.argtypes = [ctypes.POINTER(SynthCfg), ctypes.c_double, ctypes.c_void_p]
        _lib = lib
    return _lib


class SynthSequence:
    """One seeded synthetic trajectory for a sensor config. scan(k) -> (xyzi float32[P,4], ring uint16[P])."""

    def __init__(self, sensor="vlp16", seed=0, scene="street", speed=2.0, n_scans=1000, noise_sigma=0.01,
                 dropout=0.0, dup_cols=0.0, yaw_amp=0.15, yaw_period=20.0, sway_amp=1.0, circle_radius=0.0):
        p = make_params(sensor)
        eb, es, hgt = SENSOR_GEOMETRY[sensor]
        c = SynthCfg()
        c.n_scan, c.horizon_scan = p.N_SCAN, p.Horizon_SCAN
        c.elev_bottom_deg, c.elev_step_deg = eb, es
        c.ang_res_x_deg = float(p.ang_res_x)
        c.sensor_height = hgt
        c.scene = 0 if scene == "room" else 1
        c.speed, c.yaw_amp, c.yaw_period, c.sway_amp = speed, yaw_amp, yaw_period, sway_amp
        c.noise_sigma, c.max_range, c.min_range = noise_sigma, 100.0, 0.5
        c.dropout_prob, c.dup_col_prob, c.jitter_frac = dropout, dup_cols, 0.3
        c.scan_period = 0.1
        c.traj, c.circle_radius = (1, circle_radius) if circle_radius > 0 else (0, 0.0)   # closed circle (room scene) for loop closure
        self.cfg, self.sensor, self.seed, self.n_scans = c, sensor, seed, n_scans
        self.cap = p.N_SCAN * p.Horizon_SCAN + (p.N_SCAN * 64 if dup_cols > 0 else 0)
        self._h = _load().lego_synth_create(ctypes.byref(c), seed, n_scans * 0.1 + 1.0)

    def __del__(self):
        if getattr(self, "_h", None):
            _load().lego_synth_destroy(self._h)
            self._h = None

    def scan(self, k):
        xyzi = np.empty((self.cap, 4), np.float32)
        ring = np.empty(self.cap, np.uint16)
        n = _load().lego_synth_scan(self._h, k, xyzi.ctypes.data, ring.ctypes.data, self.cap)
        return xyzi[:n].copy(), ring[:n].copy()

    def batch(self, k0, count, threads=None):
        """-> xyzi float32[count,cap,4], ring uint16[count,cap], npts int32[count] (padded to cap)."""
        threads = threads or min(32, os.cpu_count() or 1)
        xyzi = np.zeros((count, self.cap, 4), np.float32)
        ring = np.zeros((count, self.cap), np.uint16)
        npts = np.zeros(count, np.int32)
        _load().lego_synth_batch(self._h, k0, count, xyzi.ctypes.data, ring.ctypes.data, npts.ctypes.data, self.cap, threads)
        return xyzi, ring, npts

    def batch_into(self, k0, count, xyz, ring, npts, floats_per_point, scan_stride_points, npts_stride=1, threads=None, ring_stride=None):
        """scans k0 .. k0+count-1 written straight into caller memory (e.g. one sequence's slots of a pinned
        [scan][sequence][cap] buffer): xyz / ring / npts are numpy views whose first element is scan k0's slot.  Scan s goes
        to xyz + s * scan_stride_points * floats_per_point floats and ring + s * scan_stride_points uint16 -- or, with
        ring_stride given, scan_stride_points counts FLOATS of the xyz channel and ring_stride uint16 of the ring channel."""
        threads = threads or min(32, os.cpu_count() or 1)
        if ring_stride is None:
            _load().lego_synth_batch_fmt(self._h, k0, count, xyz.ctypes.data, ring.ctypes.data, npts.ctypes.data, self.cap, threads,
                                         int(floats_per_point), int(scan_stride_points), int(npts_stride))
        else:
            _load().lego_synth_batch_strided(self._h, k0, count, xyz.ctypes.data, ring.ctypes.data, npts.ctypes.data, self.cap, threads,
                                             int(floats_per_point), int(scan_stride_points), int(ring_stride), int(npts_stride))

    def pose(self, t):
        out = np.zeros(6)
        _load().lego_synth_pose(ctypes.byref(self.cfg), t, out.ctypes.data)
        return out

    def stamp(self, k):
        return k * 0.1

    # ---- synthetic /imu/data (optional IMU path, featureAssociation.cpp:431-459) ----
    IMU_RATE = 100.0
    IMU_OFFSET = 0.003   # keeps IMU stamps off the scan stamps

    def _rot(self, t):
        p = self.pose(t)
        cr, sr, cp, sp, cy, sy = np.cos(p[3]), np.sin(p[3]), np.cos(p[4]), np.sin(p[4]), np.cos(p[5]), np.sin(p[5])
        return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                         [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                         [-sp, cp * sr, cp * cr]]), p

    def imu_sample(self, j):
        """j-th IMU message -> (stamp, orientation xyzw, angular velocity, linear acceleration): an ideal IMU rigidly
        mounted on the lidar (x forward, y left, z up), orientation = world_from_body, acceleration = specific force."""
        t = self.IMU_OFFSET + j / self.IMU_RATE
        h = 1e-3
        R, p = self._rot(t)
        Rm, pm = self._rot(t - h)
        Rp, pp = self._rot(t + h)
        acc_w = (pp[:3] - 2 * p[:3] + pm[:3]) / (h * h)
        f_b = R.T @ (acc_w + np.array([0.0, 0.0, 9.81]))
        dR = Rm.T @ Rp
        w_b = np.array([dR[2, 1] - dR[1, 2], dR[0, 2] - dR[2, 0], dR[1, 0] - dR[0, 1]]) / (4 * h)
        hr, hp, hy = p[3] / 2, p[4] / 2, p[5] / 2
        q = np.array([np.sin(hr) * np.cos(hp) * np.cos(hy) - np.cos(hr) * np.sin(hp) * np.sin(hy),
                      np.cos(hr) * np.sin(hp) * np.cos(hy) + np.sin(hr) * np.cos(hp) * np.sin(hy),
                      np.cos(hr) * np.cos(hp) * np.sin(hy) - np.sin(hr) * np.sin(hp) * np.cos(hy),
                      np.cos(hr) * np.cos(hp) * np.cos(hy) + np.sin(hr) * np.sin(hp) * np.sin(hy)])
        return t, q, w_b, f_b

    def imu_until(self, j0, t_end):
        """messages j0, j0+1, ... with stamp <= t_end -> (list of samples, next j)"""
        out = []
        j = j0
        while self.IMU_OFFSET + j / self.IMU_RATE <= t_end:
            out.append(self.imu_sample(j)); j += 1
        return out, j
