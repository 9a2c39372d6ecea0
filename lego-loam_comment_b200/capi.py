"""ctypes binding of include/lego_b200.h.  No computation happens here: every call goes straight into
liblego_b200.so; if the library is missing it is built with nvcc, and if that fails the import raises."""
import ctypes
import os

import numpy as np

from .params import LegoParams

_HERE = os.path.dirname(os.path.abspath(__file__))
# LEGO_B200_LIB: a differently compiled build of the same library (compile-time knob experiments under tools/)
LIB_PATH = os.environ.get("LEGO_B200_LIB") or os.path.join(_HERE, "liblego_b200.so")

# every extern "C" symbol declared in include/lego_b200.h (checked by tests/test_abi.py)
EXPORTED = [
    "lego_last_error", "lego_device_count",
    "lego_ip_create", "lego_ip_destroy", "lego_ip_process",
    "lego_fa_create", "lego_fa_destroy", "lego_fa_process",
    "lego_mo_create", "lego_mo_destroy", "lego_mo_process",
    "lego_pipeline_create", "lego_pipeline_destroy", "lego_pipeline_step", "lego_pipeline_step_lanes", "lego_pipeline_input_wait", "lego_pipeline_get_poses", "lego_pipeline_poses_async", "lego_pipeline_poses_wait",
    "lego_pipeline_sync", "lego_pipeline_launch_count", "lego_pipeline_graph_launch_count", "lego_pipeline_download", "lego_pipeline_upload",
    "lego_pipeline_timer_record", "lego_pipeline_timer_elapsed", "lego_pipeline_timer_elapsed_between", "lego_pipeline_profile", "lego_pipeline_profile_read",
    "lego_tf_create", "lego_tf_destroy", "lego_tf_laser_odometry", "lego_tf_aft_mapped",
    "lego_selftest_atan2f", "lego_selftest_std_sort", "lego_selftest_warp_sort", "lego_selftest_voxel_grid",
    "lego_selftest_sincosf", "lego_icp_align", "lego_pipeline_imu", "lego_fa_imu", "lego_mo_imu",
    "lego_pipeline_global_map", "lego_mo_global_map", "lego_write_pcd_ascii",
    "lego_voxel_grid_noalloc", "lego_voxel_grid_noalloc_max_points",
    "lego_pg_create", "lego_pg_destroy", "lego_pg_add_prior", "lego_pg_add_between", "lego_pg_insert", "lego_pg_update", "lego_pg_size", "lego_pg_estimate",
]

_lib = None


class LegoError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            from .build import build
            build()
        lib = ctypes.CDLL(LIB_PATH)
        lib.lego_last_error.restype = ctypes.c_char_p
        lib.lego_pipeline_create.argtypes = [ctypes.POINTER(LegoParams), ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
        lib.lego_pipeline_destroy.argtypes = [ctypes.c_void_p]
        lib.lego_pipeline_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                           ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.lego_pipeline_step_lanes.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                 ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.lego_pipeline_input_wait.argtypes = [ctypes.c_void_p]
        lib.lego_pipeline_sync.argtypes = [ctypes.c_void_p]
        lib.lego_pipeline_launch_count.argtypes = [ctypes.c_void_p]
        lib.lego_pipeline_launch_count.restype = ctypes.c_long
        lib.lego_pipeline_graph_launch_count.argtypes = [ctypes.c_void_p]
        lib.lego_pipeline_graph_launch_count.restype = ctypes.c_long
        lib.lego_pipeline_download.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long,
                                               ctypes.POINTER(ctypes.c_int)]
        lib.lego_pipeline_download.restype = ctypes.c_long
        lib.lego_pipeline_upload.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long]
        lib.lego_pipeline_upload.restype = ctypes.c_long
        lib.lego_pipeline_get_poses.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.lego_pipeline_poses_async.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.lego_pipeline_poses_wait.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        lib.lego_pipeline_timer_record.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.lego_pipeline_timer_elapsed.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
        lib.lego_pipeline_profile.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p]
        lib.lego_pipeline_profile_read.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
        lib.lego_pipeline_profile_read.restype = ctypes.c_long
        _lib = lib
    return _lib


def _check(rc, what):
    if rc < 0:
        raise LegoError("%s failed (%d): %s" % (what, rc, load().lego_last_error().decode()))
    return rc


MAP_GLOBAL, MAP_CORNER, MAP_SURFACE, MAP_TRAJECTORY = 0, 1, 2, 3


def _global_map(fn, head, what):
    fn.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * (len(head) - 1) + [ctypes.c_void_p, ctypes.c_long, ctypes.POINTER(ctypes.c_long)]
    n = ctypes.c_long(0)
    _check(fn(*head, None, 0, ctypes.byref(n)), what)
    out = np.zeros((n.value, 4), np.float32)
    if n.value:
        _check(fn(*head, out.ctypes.data, n.value, ctypes.byref(n)), what)
    return out[:n.value]


def write_pcd_ascii(path, xyzi):
    """pcl::io::savePCDFileASCII for an (n, 4) x y z intensity cloud (lego_write_pcd_ascii)"""
    c = np.ascontiguousarray(xyzi, np.float32).reshape(-1, 4)
    f = load().lego_write_pcd_ascii
    f.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long]
    _check(f(os.fsencode(path), c.ctypes.data if len(c) else None, len(c)), "lego_write_pcd_ascii")


_DTYPES = {("f", 4): np.float32, ("i", 4): np.int32, ("i", 1): np.int8, ("u", 1): np.uint8, ("f", 16): np.float32}
_INT_NAMES = {"groundMat": np.int8, "segGroundFlag": np.uint8}


def selftest_atan2f(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(y)
    lib = load()
    lib.lego_selftest_atan2f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
    _check(lib.lego_selftest_atan2f(y.ctypes.data, x.ctypes.data, out.ctypes.data, y.size), "lego_selftest_atan2f")
    return out


def selftest_sincosf(x):
    x = np.ascontiguousarray(x, np.float32)
    s = np.empty_like(x); c = np.empty_like(x)
    lib = load()
    lib.lego_selftest_sincosf.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
    _check(lib.lego_selftest_sincosf(x.ctypes.data, s.ctypes.data, c.ctypes.data, x.size), "lego_selftest_sincosf")
    return s, c


def selftest_std_sort(values, offsets):
    values = np.ascontiguousarray(values, np.float32); offsets = np.ascontiguousarray(offsets, np.int32)
    out = np.empty(values.size, np.int32)
    lib = load()
    lib.lego_selftest_std_sort.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    _check(lib.lego_selftest_std_sort(values.ctypes.data, offsets.ctypes.data, offsets.size - 1, out.ctypes.data), "lego_selftest_std_sort")
    return out


def selftest_warp_sort(values, offsets):
    values = np.ascontiguousarray(values, np.float32); offsets = np.ascontiguousarray(offsets, np.int32)
    out = np.empty(values.size, np.int32)
    lib = load()
    lib.lego_selftest_warp_sort.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    _check(lib.lego_selftest_warp_sort(values.ctypes.data, offsets.ctypes.data, offsets.size - 1, out.ctypes.data), "lego_selftest_warp_sort")
    return out


def selftest_voxel_grid(clouds, leaf):
    """clouds: list of float32[n_i, 4] -> list of filtered float32[m_i, 4] (device VoxelGrid, one segment per cloud)."""
    offsets = np.concatenate([[0], np.cumsum([len(c) for c in clouds])]).astype(np.int32)
    xyzi = np.ascontiguousarray(np.concatenate(clouds) if len(clouds) else np.zeros((0, 4)), np.float32)
    out = np.empty((max(int(offsets[-1]), 1), 4), np.float32)
    counts = np.zeros(len(clouds), np.int32)
    lib = load()
    lib.lego_selftest_voxel_grid.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p]
    _check(lib.lego_selftest_voxel_grid(xyzi.ctypes.data, offsets.ctypes.data, len(clouds), leaf, out.ctypes.data, counts.ctypes.data),
           "lego_selftest_voxel_grid")
    res, w = [], 0
    for c in counts:
        res.append(out[w:w + c].copy()); w += c
    return res


def voxel_grid_noalloc(d_xyzi_ptr, n, leaf, d_out_ptr, out_cap, d_n_out_ptr, stream=0):
    """lego_voxel_grid_noalloc: raw DEVICE addresses in, kernels enqueued on `stream` (a cudaStream_t as int, 0 = default
    stream), nothing allocated, nothing synchronised.  Returns the C return code (LEGO_OK = 0, negative = error)."""
    lib = load()
    lib.lego_voxel_grid_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    return lib.lego_voxel_grid_noalloc(d_xyzi_ptr, int(n), float(leaf), d_out_ptr, int(out_cap), d_n_out_ptr, stream or None)


class Pipeline:
    """`lanes` independent sequences stepped in lock step on one GPU (lego_pipeline_* in include/lego_b200.h)."""
    IP, FA, MO = 1, 2, 4

    def __init__(self, params, lanes=1, max_points=0):
        self.lib = load()
        self.params = params
        self.lanes = lanes
        self.h = ctypes.c_void_p()
        _check(self.lib.lego_pipeline_create(ctypes.byref(params), lanes, max_points, ctypes.byref(self.h)), "lego_pipeline_create")
        self.N = params.N_SCAN * params.Horizon_SCAN

    def close(self):
        if getattr(self, "h", None):
            self.lib.lego_pipeline_destroy(self.h)
            self.h = None

    __del__ = close

    def step(self, xyzi, ring, npts, stamps, stage_mask=7, on_device=False, stride_points=None):
        """xyzi float32[lanes, stride, 4], ring uint16[lanes, stride], npts int32[lanes], stamps float64[lanes]
        (numpy host arrays, or raw device pointers as ints when on_device)."""
        npts = np.ascontiguousarray(npts, np.int32)
        stamps = np.ascontiguousarray(stamps, np.float64)
        if on_device:
            xp, rp = xyzi, ring
        else:
            xyzi = np.ascontiguousarray(xyzi, np.float32)
            ring = None if ring is None else np.ascontiguousarray(ring, np.uint16)
            stride_points = xyzi.shape[1]
            xp, rp = xyzi.ctypes.data, (None if ring is None else ring.ctypes.data)
        _check(self.lib.lego_pipeline_step(self.h, xp, rp, npts.ctypes.data, int(stride_points), stamps.ctypes.data,
                                           1 if on_device else 0, stage_mask), "lego_pipeline_step")

    def step_lanes(self, xyz_ptrs, ring_ptrs, npts, stamps, floats_per_point=4, stage_mask=7, on_device=False):
        """lego_pipeline_step_lanes: xyz_ptrs / ring_ptrs = uint64[lanes] tables of per-lane addresses; on_device False / 0 =
        host memory, True / 1 = device memory, 2 = page-locked device-mapped host memory (packing kernel), 3 = host memory whose
        lane blocks are slots of one allocation (neighbouring lanes share a copy, the bytes between them are read too).
        Otherwise only npts[l] points of lane l are read."""
        npts = np.ascontiguousarray(npts, np.int32)
        stamps = np.ascontiguousarray(stamps, np.float64)
        xp = np.ascontiguousarray(xyz_ptrs, np.uint64)
        rp = None if ring_ptrs is None else np.ascontiguousarray(ring_ptrs, np.uint64)
        assert xp.shape == (self.lanes,) and npts.shape == (self.lanes,) and stamps.shape == (self.lanes,)
        _check(self.lib.lego_pipeline_step_lanes(self.h, xp.ctypes.data, None if rp is None else rp.ctypes.data, npts.ctypes.data,
                                                 int(floats_per_point), stamps.ctypes.data, int(on_device), stage_mask),
               "lego_pipeline_step_lanes")

    def input_wait(self):
        _check(self.lib.lego_pipeline_input_wait(self.h), "lego_pipeline_input_wait")

    def imu(self, lane, stamp, orientation_xyzw, angular_velocity, linear_acceleration):
        """one sensor_msgs/Imu message for one lane (lego_pipeline_imu)"""
        f = self.lib.lego_pipeline_imu
        f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        q = np.ascontiguousarray(orientation_xyzw, np.float64); w = np.ascontiguousarray(angular_velocity, np.float64)
        a = np.ascontiguousarray(linear_acceleration, np.float64)
        _check(f(self.h, lane, float(stamp), q.ctypes.data, w.ctypes.data, a.ctypes.data), "lego_pipeline_imu")

    def global_map(self, which=MAP_GLOBAL, lane=0):
        """publishGlobalMap's cloud (MAP_GLOBAL) or one of the clouds the reference saves at shutdown (MAP_CORNER,
        MAP_SURFACE, MAP_TRAJECTORY), mapOptmization.cpp:724-800; (n, 4) float32"""
        return _global_map(self.lib.lego_pipeline_global_map, (self.h, int(lane), int(which)), "lego_pipeline_global_map")

    def sync(self):
        _check(self.lib.lego_pipeline_sync(self.h), "lego_pipeline_sync")

    def launch_count(self):
        return self.lib.lego_pipeline_launch_count(self.h)

    def graph_launch_count(self):
        return self.lib.lego_pipeline_graph_launch_count(self.h)

    def timer_record(self, slot):
        _check(self.lib.lego_pipeline_timer_record(self.h, slot), "lego_pipeline_timer_record")

    def timer_elapsed_ms(self, a, b):
        ms = ctypes.c_float(0)
        _check(self.lib.lego_pipeline_timer_elapsed(self.h, a, b, ctypes.byref(ms)), "lego_pipeline_timer_elapsed")
        return ms.value

    def timer_elapsed_to(self, a, other, b):
        """device ms from this handle's event a to `other`'s event b"""
        ms = ctypes.c_float(0)
        f = self.lib.lego_pipeline_timer_elapsed_between
        f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
        _check(f(self.h, a, other.h, b, ctypes.byref(ms)), "lego_pipeline_timer_elapsed_between")
        return ms.value

    def profile(self, mode, name_filter=""):
        _check(self.lib.lego_pipeline_profile(self.h, mode, name_filter.encode()), "lego_pipeline_profile")

    def profile_read(self, max_kernels=128):
        names = ctypes.create_string_buffer(8192)
        ms = np.zeros(max_kernels, np.float64)
        counts = np.zeros(max_kernels, np.int64)
        n = _check(self.lib.lego_pipeline_profile_read(self.h, names, 8192, ms.ctypes.data, counts.ctypes.data, max_kernels), "profile_read")
        keys = names.value.decode().split("\n")[:n]
        return {k: (float(ms[i]), int(counts[i])) for i, k in enumerate(keys)}

    def poses(self):
        poses = np.zeros((self.lanes, 18), np.float32)
        flags = np.zeros(self.lanes, np.int32)
        _check(self.lib.lego_pipeline_get_poses(self.h, poses.ctypes.data, flags.ctypes.data), "lego_pipeline_get_poses")
        return poses, flags

    def poses_async(self, slot):
        """Enqueue the read-back of the poses after the last step into pinned slot 0..7 (returns at once)."""
        _check(self.lib.lego_pipeline_poses_async(self.h, int(slot)), "lego_pipeline_poses_async")

    def poses_wait(self, slot):
        poses = np.zeros((self.lanes, 18), np.float32)
        flags = np.zeros(self.lanes, np.int32)
        _check(self.lib.lego_pipeline_poses_wait(self.h, int(slot), poses.ctypes.data, flags.ctypes.data), "lego_pipeline_poses_wait")
        return poses, flags

    def download(self, name, lane=0):
        es = ctypes.c_int(0)
        n = _check(self.lib.lego_pipeline_download(self.h, lane, name.encode(), None, 0, ctypes.byref(es)), "download " + name)
        e = es.value
        if e == 16:
            out = np.empty((n, 4), np.float32)
        elif e == 1:
            out = np.empty(n, _INT_NAMES.get(name, np.uint8))
        elif e == 8:
            out = np.empty(n, np.float64)
        elif name in _FLOAT_NAMES:
            out = np.empty(n, np.float32)
        else:
            out = np.empty(n, np.int32)
        if n:
            _check(self.lib.lego_pipeline_download(self.h, lane, name.encode(), out.ctypes.data, n, ctypes.byref(es)), "download " + name)
        return out

    def upload(self, name, arr, lane=0):
        arr = np.ascontiguousarray(arr)
        cnt = arr.shape[0]
        _check(self.lib.lego_pipeline_upload(self.h, lane, name.encode(), arr.ctypes.data, cnt), "upload " + name)


_FLOAT_NAMES = {"rangeMat", "segRange", "orientation", "cloudCurvature", "transformCur", "transformSum", "slot4val",
                "matP", "moTransformSum", "transformTobeMapped", "transformAftMapped", "transformBefMapped",
                "transformMapped", "integratedToInit", "tfTransformAftMapped", "tfTransformBefMapped",
                "cloudKeyPoses6D", "icpTransform", "icpFitness", "imuState"}


def icp_align(source, target, max_correspondence_distance=100.0, max_iterations=100, transformation_epsilon=1e-6,
              euclidean_fitness_epsilon=1e-6, device=0):
    """lego_icp_align: pcl::IterativeClosestPoint as configured at mapOptmization.cpp:892-905, on the device.
    -> (T float32[4,4], fitness, converged, iterations)"""
    source = np.ascontiguousarray(source, np.float32).reshape(-1, 4)
    target = np.ascontiguousarray(target, np.float32).reshape(-1, 4)
    T = np.zeros(16, np.float32)
    fit, conv, it = ctypes.c_double(0), ctypes.c_int(0), ctypes.c_int(0)
    lib = load()
    lib.lego_icp_align.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_int,
                                   ctypes.c_double, ctypes.c_double, ctypes.c_void_p, ctypes.POINTER(ctypes.c_double),
                                   ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
    _check(lib.lego_icp_align(device, source.ctypes.data, len(source), target.ctypes.data, len(target), max_correspondence_distance,
                              max_iterations, transformation_epsilon, euclidean_fitness_epsilon, T.ctypes.data, ctypes.byref(fit),
                              ctypes.byref(conv), ctypes.byref(it)), "lego_icp_align")
    return T.reshape(4, 4), fit.value, conv.value, it.value


class PoseGraph:
    """lego_pg_*: the Pose3 graph mapOptimization keeps in iSAM2 (mapOptmization.cpp:1375-1409, 929-942). Host code."""

    def __init__(self):
        self.lib = lib = load()
        P = ctypes.POINTER(ctypes.c_double)
        lib.lego_pg_create.argtypes = [ctypes.POINTER(ctypes.c_void_p)]
        lib.lego_pg_destroy.argtypes = [ctypes.c_void_p]
        lib.lego_pg_add_prior.argtypes = [ctypes.c_void_p, ctypes.c_int, P, P]
        lib.lego_pg_add_between.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, P, P, P]
        lib.lego_pg_insert.argtypes = [ctypes.c_void_p, ctypes.c_int, P]
        lib.lego_pg_update.argtypes = [ctypes.c_void_p]
        lib.lego_pg_size.argtypes = [ctypes.c_void_p]
        lib.lego_pg_estimate.argtypes = [ctypes.c_void_p, ctypes.c_int, P]
        self.h = ctypes.c_void_p()
        _check(lib.lego_pg_create(ctypes.byref(self.h)), "lego_pg_create")

    @staticmethod
    def _v(x):
        return (ctypes.c_double * 6)(*[float(v) for v in x])

    def add_prior(self, key, pose, variances):
        _check(self.lib.lego_pg_add_prior(self.h, key, self._v(pose), self._v(variances)), "lego_pg_add_prior")

    def add_between(self, k1, k2, pose_from, pose_to, variances):
        _check(self.lib.lego_pg_add_between(self.h, k1, k2, self._v(pose_from), self._v(pose_to), self._v(variances)), "lego_pg_add_between")

    def insert(self, key, pose):
        _check(self.lib.lego_pg_insert(self.h, key, self._v(pose)), "lego_pg_insert")

    def update(self):
        _check(self.lib.lego_pg_update(self.h), "lego_pg_update")

    def estimate(self):
        n = self.lib.lego_pg_size(self.h)
        out = np.zeros((n, 6))
        buf = (ctypes.c_double * 6)()
        for k in range(n):
            _check(self.lib.lego_pg_estimate(self.h, k, buf), "lego_pg_estimate")
            out[k] = list(buf)
        return out

    def close(self):
        if getattr(self, "h", None):
            self.lib.lego_pg_destroy(self.h)
            self.h = None

    __del__ = close


# ---------------------------------------------------------------------------------------------- per-stage handles
class CloudView(ctypes.Structure):
    _fields_ = [("dev", ctypes.c_void_p), ("host", ctypes.c_void_p), ("count", ctypes.c_int)]

    def numpy(self):
        if not self.count:
            return np.zeros((0, 4), np.float32)
        return np.ctypeslib.as_array(ctypes.cast(self.host, ctypes.POINTER(ctypes.c_float)), (self.count, 4)).copy()


class CloudInfoView(ctypes.Structure):
    """cloud_msgs/cloud_info (cloud_msgs/msg/cloud_info.msg:1-11)"""
    _fields_ = [("startRingIndex", ctypes.c_void_p), ("endRingIndex", ctypes.c_void_p),
                ("startOrientation", ctypes.c_float), ("endOrientation", ctypes.c_float), ("orientationDiff", ctypes.c_float),
                ("segmentedCloudGroundFlag", ctypes.c_void_p), ("segmentedCloudColInd", ctypes.c_void_p),
                ("segmentedCloudRange", ctypes.c_void_p), ("count", ctypes.c_int), ("on_device", ctypes.c_int)]


class FaOutputs(ctypes.Structure):
    _fields_ = [("cornerPointsSharp", CloudView), ("cornerPointsLessSharp", CloudView), ("surfPointsFlat", CloudView),
                ("surfPointsLessFlat", CloudView), ("laserCloudCornerLast", CloudView), ("laserCloudSurfLast", CloudView),
                ("outlierCloudLast", CloudView), ("odometry_valid", ctypes.c_int), ("published", ctypes.c_int),
                ("transformCur", ctypes.c_float * 6)]


class _Stage:
    _prefix = ""

    def __init__(self, params):
        self.lib = load()
        self.h = ctypes.c_void_p()
        create = getattr(self.lib, self._prefix + "_create")
        create.argtypes = [ctypes.POINTER(LegoParams), ctypes.POINTER(ctypes.c_void_p)]
        _check(create(ctypes.byref(params), ctypes.byref(self.h)), self._prefix + "_create")

    def close(self):
        if getattr(self, "h", None):
            d = getattr(self.lib, self._prefix + "_destroy")
            d.argtypes = [ctypes.c_void_p]
            d(self.h)
            self.h = None

    __del__ = close


class ImageProjection(_Stage):
    """lego_ip_* : drop-in for ImageProjection::cloudHandler (imageProjection.cpp:181-197)."""
    _prefix = "lego_ip"

    def process(self, xyzi, ring, stamp, is_dense=True):
        xyzi = np.ascontiguousarray(xyzi, np.float32)
        ring = None if ring is None else np.ascontiguousarray(ring, np.uint16)
        info, seg, outl = CloudInfoView(), CloudView(), CloudView()
        f = self.lib.lego_ip_process
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                      ctypes.POINTER(CloudInfoView), ctypes.POINTER(CloudView), ctypes.POINTER(CloudView)]
        rc = f(self.h, xyzi.ctypes.data, None if ring is None else ring.ctypes.data, len(xyzi), 1 if is_dense else 0, stamp,
               ctypes.byref(info), ctypes.byref(seg), ctypes.byref(outl))
        _check(rc, "lego_ip_process")
        return info, seg, outl


class FeatureAssociation(_Stage):
    """lego_fa_* : drop-in for FeatureAssociation::runFeatureAssociation (featureAssociation.cpp:1817-1860)."""
    _prefix = "lego_fa"

    def process(self, info, seg, outl, stamp):
        out = FaOutputs()
        ts = (ctypes.c_float * 6)()
        f = self.lib.lego_fa_process
        f.argtypes = [ctypes.c_void_p, ctypes.POINTER(CloudInfoView), ctypes.POINTER(CloudView), ctypes.POINTER(CloudView),
                      ctypes.c_double, ctypes.c_void_p, ctypes.POINTER(FaOutputs)]
        _check(f(self.h, ctypes.byref(info), ctypes.byref(seg), ctypes.byref(outl), stamp, ts, ctypes.byref(out)), "lego_fa_process")
        return np.array(ts[:], np.float32), out


class MapOptimization(_Stage):
    """lego_mo_* : drop-in for mapOptimization::run (mapOptmization.cpp:1487-1522)."""
    _prefix = "lego_mo"

    def process(self, corner_last, surf_last, outlier_last, transform_sum, stamp):
        aft, bef = (ctypes.c_float * 6)(), (ctypes.c_float * 6)()
        done = ctypes.c_int(0)
        ts = np.ascontiguousarray(transform_sum, np.float32)
        f = self.lib.lego_mo_process
        f.argtypes = [ctypes.c_void_p, ctypes.POINTER(CloudView), ctypes.POINTER(CloudView), ctypes.POINTER(CloudView), ctypes.c_void_p,
                      ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        _check(f(self.h, ctypes.byref(corner_last), ctypes.byref(surf_last), ctypes.byref(outlier_last), ts.ctypes.data, stamp, aft, bef,
                 ctypes.byref(done)), "lego_mo_process")
        return np.array(aft[:], np.float32), np.array(bef[:], np.float32), bool(done.value)


    def global_map(self, which=MAP_GLOBAL):
        return _global_map(self.lib.lego_mo_global_map, (self.h, int(which)), "lego_mo_global_map")


class TransformFusion(_Stage):
    """lego_tf_* : drop-in for the two handlers of the TransformFusion node (transformFusion.cpp:181-241)."""
    _prefix = "lego_tf"

    def laser_odometry(self, transform_sum):
        ts = np.ascontiguousarray(transform_sum, np.float32)
        mapped, integrated = np.zeros(6, np.float32), np.zeros(7, np.float32)
        f = self.lib.lego_tf_laser_odometry
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _check(f(self.h, ts.ctypes.data, mapped.ctypes.data, integrated.ctypes.data), "lego_tf_laser_odometry")
        return mapped, integrated

    def aft_mapped(self, transform_aft_mapped, transform_bef_mapped):
        a = np.ascontiguousarray(transform_aft_mapped, np.float32); b = np.ascontiguousarray(transform_bef_mapped, np.float32)
        f = self.lib.lego_tf_aft_mapped
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _check(f(self.h, a.ctypes.data, b.ctypes.data), "lego_tf_aft_mapped")
