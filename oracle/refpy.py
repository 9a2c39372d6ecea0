"""TEST INFRASTRUCTURE ONLY.  ctypes binding for oracle/_ref/liblego_ref_<cfg>.so (the reference's own sources
compiled verbatim against oracle/shim; see oracle/ref_harness.cpp).  Used by tests/, smoke() and bench.py's
cpu_baseline / --impl reference arm only."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def ref_so(cfg):
    return os.path.join(HERE, "_ref", "liblego_ref_%s.so" % cfg)


def available(cfg):
    return os.path.exists(ref_so(cfg))


class RefPipeline:
    IP, FA, MO = 1, 2, 4

    def __init__(self, cfg="vlp16", snapshots=True):
        self.lib = lib = ctypes.CDLL(ref_so(cfg))
        lib.ref_create.restype = ctypes.c_void_p
        lib.ref_create.argtypes = [ctypes.c_int]
        lib.ref_destroy.argtypes = [ctypes.c_void_p]
        lib.ref_set_verbatim_drivers.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.ref_feed_scan.restype = ctypes.c_int
        lib.ref_feed_scan.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                      ctypes.c_double, ctypes.c_int]
        for fn in (lib.ref_get_f32, lib.ref_get_i32):
            fn.restype = ctypes.c_long
            fn.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long]
        lib.ref_get_poses.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.ref_get_timers.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.ref_timer_name.restype = ctypes.c_char_p
        lib.ref_timer_name.argtypes = [ctypes.c_int]
        lib.ref_params.argtypes = [ctypes.c_void_p]
        self.N_SCAN, self.Horizon_SCAN = lib.ref_n_scan(), lib.ref_horizon_scan()
        self.groundScanInd, self.useCloudRing = lib.ref_ground_scan_ind(), lib.ref_use_cloud_ring()
        self.h = lib.ref_create(1 if snapshots else 0)

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    __del__ = close

    def params(self):
        out = np.zeros(8, np.float32)
        self.lib.ref_params(out.ctypes.data)
        return dict(zip(["ang_res_x", "ang_res_y", "ang_bottom", "segmentAlphaX", "segmentAlphaY", "segmentTheta",
                         "sensorMinimumRange", "sensorMountAngle"], out))

    def set_verbatim(self, on=True):
        self.lib.ref_set_verbatim_drivers(self.h, 1 if on else 0)

    def feed(self, xyzi, ring, stamp, stages=7, is_dense=True):
        xyzi = np.ascontiguousarray(xyzi, np.float32)
        ring_p = None
        if ring is not None:
            ring = np.ascontiguousarray(ring, np.uint16)
            ring_p = ring.ctypes.data
        return self.lib.ref_feed_scan(self.h, xyzi.ctypes.data, ring_p, int(xyzi.shape[0]), 1 if is_dense else 0,
                                      float(stamp), stages)

    def imu(self, stamp, orientation_xyzw, angular_velocity, linear_acceleration):
        """one sensor_msgs/Imu message to FeatureAssociation::imuHandler and mapOptimization::imuHandler"""
        f = self.lib.ref_feed_imu
        f.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        q = np.ascontiguousarray(orientation_xyzw, np.float64); w = np.ascontiguousarray(angular_velocity, np.float64)
        a = np.ascontiguousarray(linear_acceleration, np.float64)
        f(self.h, float(stamp), q.ctypes.data, w.ctypes.data, a.ctypes.data)

    def get(self, name):
        b = name.encode()
        n = self.lib.ref_get_f32(self.h, b, None, 0)
        if n >= 0:
            out = np.empty(n, np.float32)
            self.lib.ref_get_f32(self.h, b, out.ctypes.data, n)
            return out
        n = self.lib.ref_get_i32(self.h, b, None, 0)
        if n >= 0:
            out = np.empty(n, np.int32)
            self.lib.ref_get_i32(self.h, b, out.ctypes.data, n)
            return out
        raise KeyError(name)

    def cloud(self, name):
        return self.get(name).reshape(-1, 4)

    def global_map(self):
        """publishGlobalMap + the shutdown half of visualizeGlobalMapThread (mapOptmization.cpp:724-800); afterwards
        cloud("globalMapKeyFramesDS"), cloud("pcd:finalCloud" | "pcd:cornerMap" | "pcd:surfaceMap" | "pcd:trajectory")
        and pcd_text(name) are available.  Returns the number of PCD files the reference wrote."""
        self.lib.ref_global_map.argtypes = [ctypes.c_void_p]
        return self.lib.ref_global_map(self.h)

    def pcd_text(self, name):
        f = self.lib.ref_get_pcd_text
        f.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long]
        f.restype = ctypes.c_long
        n = f(self.h, name.encode(), None, 0)
        if n < 0:
            raise KeyError(name)
        buf = ctypes.create_string_buffer(n)
        f(self.h, name.encode(), buf, n)
        return buf.raw[:n]

    def poses(self):
        out = np.zeros(18, np.float32)
        self.lib.ref_get_poses(self.h, out.ctypes.data)
        return out[:6].copy(), out[6:12].copy(), out[12:].copy()

    def timers(self):
        n = self.lib.ref_num_timers()
        ms = np.zeros(n, np.float64)
        scans, mo = ctypes.c_long(0), ctypes.c_long(0)
        self.lib.ref_get_timers(self.h, ms.ctypes.data, ctypes.byref(scans), ctypes.byref(mo))
        return {self.lib.ref_timer_name(i).decode(): ms[i] for i in range(n)}, scans.value, mo.value


# ---------------------------------------------------------------- third-party restatements (oracle/port.cpp)
_port = None


def port():
    """oracle/_build/liblego_oracle.so (built by `make -C oracle port`; needs no /root/reference)."""
    global _port
    if _port is None:
        so = os.path.join(HERE, "_build", "liblego_oracle.so")
        if not os.path.exists(so):
            import subprocess
            subprocess.check_call(["make", "-s", "-C", HERE, "port"])
        lib = ctypes.CDLL(so)
        lib.oracle_voxel_grid.restype = ctypes.c_long
        lib.oracle_voxel_grid.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_float, ctypes.c_void_p, ctypes.c_long]
        lib.oracle_knn.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_long, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_solve_qr.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.oracle_eigen_sym.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        _port = lib
    return _port


def voxel_grid(xyzi, leaf):
    xyzi = np.ascontiguousarray(xyzi, np.float32).reshape(-1, 4)
    out = np.empty((max(len(xyzi), 1), 4), np.float32)
    m = port().oracle_voxel_grid(xyzi.ctypes.data, len(xyzi), leaf, out.ctypes.data, len(out))
    return out[:m].copy()


def knn(pts, queries, k):
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 4)
    queries = np.ascontiguousarray(queries, np.float32).reshape(-1, 4)
    idx = np.empty((len(queries), k), np.int32); d2 = np.empty((len(queries), k), np.float32)
    port().oracle_knn(pts.ctypes.data, len(pts), queries.ctypes.data, len(queries), k, idx.ctypes.data, d2.ctypes.data)
    return idx, d2


def icp(source, target, max_dist=100.0, max_iter=100, eps_t=1e-6, eps_f=1e-6):
    """oracle restatement of pcl::IterativeClosestPoint -> (T float32[4,4], fitness, converged, iterations)"""
    source = np.ascontiguousarray(source, np.float32).reshape(-1, 4)
    target = np.ascontiguousarray(target, np.float32).reshape(-1, 4)
    lib = port()
    lib.oracle_icp.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_long, ctypes.c_double, ctypes.c_int, ctypes.c_double,
                               ctypes.c_double, ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
    T = np.zeros(16, np.float32)
    fit, conv, it = ctypes.c_double(0), ctypes.c_int(0), ctypes.c_int(0)
    lib.oracle_icp(source.ctypes.data, len(source), target.ctypes.data, len(target), max_dist, max_iter, eps_t, eps_f, T.ctypes.data,
                   ctypes.byref(fit), ctypes.byref(conv), ctypes.byref(it))
    return T.reshape(4, 4), fit.value, conv.value, it.value


def pose_graph(poses, factors, meas, variances, steps):
    """oracle Gauss-Newton on a Pose3 graph; see oracle/port.cpp::oracle_pose_graph for the layouts."""
    poses = np.ascontiguousarray(poses, np.float64); factors = np.ascontiguousarray(factors, np.int32)
    meas = np.ascontiguousarray(meas, np.float64); variances = np.ascontiguousarray(variances, np.float64)
    out = np.zeros_like(poses)
    lib = port()
    lib.oracle_pose_graph.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    lib.oracle_pose_graph(poses.ctypes.data, len(poses), factors.ctypes.data, meas.ctypes.data, variances.ctypes.data, len(factors), steps, out.ctypes.data)
    return out


# ---------------------------------------------------------------- topics (reference nodes and B200 node shells)
TOPICS_CLOUD = ["/segmented_cloud", "/outlier_cloud", "/laser_cloud_sharp", "/laser_cloud_less_sharp", "/laser_cloud_flat", "/laser_cloud_less_flat",
                "/laser_cloud_corner_last", "/laser_cloud_surf_last", "/outlier_cloud_last"]
TOPICS_ODOMETRY = ["/laser_odom_to_init", "/aft_mapped_to_init", "/integrated_to_init"]
TOPIC_INFO = "/segmented_cloud_info"


def _topic_dump(fn, topic, nvalid=0):
    """-> dict(values float64[n], stamp, frame_id, child_frame_id, count) of the last message on `topic`, or None"""
    fn.restype = ctypes.c_long
    fn.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_long, ctypes.POINTER(ctypes.c_double), ctypes.c_char_p,
                   ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_long)]
    stamp, count = ctypes.c_double(0), ctypes.c_long(0)
    frame, child = ctypes.create_string_buffer(64), ctypes.create_string_buffer(64)
    n = fn(topic.encode(), None, 0, nvalid, ctypes.byref(stamp), frame, child, 64, ctypes.byref(count))
    if n < 0:
        return None
    out = np.zeros(max(n, 1), np.float64)
    fn(topic.encode(), out.ctypes.data, n, nvalid, ctypes.byref(stamp), frame, child, 64, ctypes.byref(count))
    return {"values": out[:n], "stamp": stamp.value, "frame_id": frame.value.decode(), "child_frame_id": child.value.decode(), "count": count.value}


def ref_topic(ref, topic, nvalid=0):
    """last message the REFERENCE's nodes published on `topic` (ref: RefPipeline)"""
    return _topic_dump(ref.lib.ref_topic_dump, topic, nvalid)


def ref_subscribe_all(ref, on=True):
    ref.lib.ref_subscribe_all.argtypes = [ctypes.c_int]
    ref.lib.ref_subscribe_all(1 if on else 0)


def shells_so(cfg):
    return os.path.join(HERE, "_ref", "liblego_shells_%s.so" % cfg)


class ShellPipeline:
    """The four B200 ROS node shells (ros/lego_loam_b200/src) compiled against oracle/shim and linked to liblego_b200.so,
    driven synchronously (oracle/shell_harness.cpp).  Needs a CUDA device: the shells call the C ABI."""

    def __init__(self, cfg="vlp16", subscribe_all=True):
        self.lib = lib = ctypes.CDLL(shells_so(cfg))
        lib.shell_create.restype = ctypes.c_void_p
        lib.shell_create.argtypes = [ctypes.c_int]
        lib.shell_destroy.argtypes = [ctypes.c_void_p]
        lib.shell_feed_scan.restype = ctypes.c_int
        lib.shell_feed_scan.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double]
        self.h = lib.shell_create(1 if subscribe_all else 0)
        if not self.h:
            raise RuntimeError("node shells could not be created (no CUDA device?)")

    def feed(self, xyzi, ring, stamp, is_dense=True):
        xyzi = np.ascontiguousarray(xyzi, np.float32)
        ring = None if ring is None else np.ascontiguousarray(ring, np.uint16)
        return self.lib.shell_feed_scan(self.h, xyzi.ctypes.data, None if ring is None else ring.ctypes.data, int(xyzi.shape[0]),
                                        1 if is_dense else 0, float(stamp))

    def topic(self, topic, nvalid=0):
        return _topic_dump(self.lib.shell_topic_dump, topic, nvalid)

    def errors(self):
        return self.lib.shell_errors()

    def close(self):
        if self.h:
            self.lib.shell_destroy(self.h)
            self.h = None

    __del__ = close
