"""Sensor / algorithm parameters, field names 1:1 with the reference's compile-time constants
(LeGO-LOAM/include/utility.h:60-136).  The reference changes sensor by editing the header and rebuilding
(README.md:60-86); here the same names are runtime fields of `lego_params` (include/lego_b200.h)."""
import ctypes
import math

import numpy as np


class LegoParams(ctypes.Structure):
    """ctypes mirror of `lego_params` in include/lego_b200.h (keep field order identical)."""
    _fields_ = [
        ("N_SCAN", ctypes.c_int), ("Horizon_SCAN", ctypes.c_int),
        ("ang_res_x", ctypes.c_float), ("ang_res_y", ctypes.c_float), ("ang_bottom", ctypes.c_float),
        ("groundScanInd", ctypes.c_int), ("useCloudRing", ctypes.c_int),
        ("sensorMinimumRange", ctypes.c_float), ("sensorMountAngle", ctypes.c_float),
        ("segmentTheta", ctypes.c_float), ("segmentValidPointNum", ctypes.c_int), ("segmentValidLineNum", ctypes.c_int),
        ("segmentAlphaX", ctypes.c_float), ("segmentAlphaY", ctypes.c_float),
        ("scanPeriod", ctypes.c_float),
        ("edgeFeatureNum", ctypes.c_int), ("surfFeatureNum", ctypes.c_int), ("sectionsTotal", ctypes.c_int),
        ("edgeThreshold", ctypes.c_float), ("surfThreshold", ctypes.c_float),
        ("nearestFeatureSearchSqDist", ctypes.c_float),
        ("surroundingKeyframeSearchRadius", ctypes.c_float),
        ("mappingProcessInterval", ctypes.c_double),
        ("loopClosureEnableFlag", ctypes.c_int),
        ("device", ctypes.c_int),
        ("surroundingKeyframeSearchNum", ctypes.c_int), ("historyKeyframeSearchRadius", ctypes.c_float),
        ("historyKeyframeSearchNum", ctypes.c_int), ("historyKeyframeFitnessScore", ctypes.c_float),
        ("maxKeyFrames", ctypes.c_int),
    ]


def _f32(x):
    return float(np.float32(x))


# name -> (N_SCAN, Horizon_SCAN, ang_res_x, ang_res_y, ang_bottom, groundScanInd, useCloudRing); every value is
# evaluated exactly as the C expression in the header would be (double arithmetic, then narrowed to float).
_SENSORS = {
    "vlp16": (16, 1800, 0.2, 2.0, 15.0 + 0.1, 7, 1),                                   # utility.h:63-68
    "vlp16_noring": (16, 1800, 0.2, 2.0, 15.0 + 0.1, 7, 0),                            # utility.h:60 = false
    "hdl32": (32, 1800, 360.0 / _f32(1800), 41.333 / _f32(31), 30.666666, 20, 1),      # README.md:73-82
    "hdl64": (64, 2083, 360.0 / _f32(2083), 26.9 / _f32(63), 24.9, 50, 1),             # SURVEY.md 8c (not in reference)
    "vls128": (128, 1800, 0.2, 0.3, 25.0, 10, 1),                                      # utility.h:79-84
    "os1_16": (16, 1024, 360.0 / _f32(1024), 33.2 / _f32(15), 16.6 + 0.1, 7, 1),       # utility.h:88-94
    "os1_64": (64, 1024, 360.0 / _f32(1024), 33.2 / _f32(63), 16.6 + 0.1, 15, 1),      # utility.h:96-102
}

# beam geometry used only by the synthetic generator: (elev_bottom_deg, elev_step_deg, sensor_height)
SENSOR_GEOMETRY = {
    "vlp16": (-15.0, 2.0, 0.7), "vlp16_noring": (-15.0, 2.0, 0.7),
    "hdl32": (-30.67, 41.33 / 31.0, 1.2), "hdl64": (-24.9, 26.9 / 63.0, 1.73), "vls128": (-25.0, 0.3, 1.8),
    "os1_16": (-16.6, 33.2 / 15.0, 0.9), "os1_64": (-16.6, 33.2 / 63.0, 1.5),
}


def make_params(sensor="vlp16", device=0, loop_closure=False, max_keyframes=0):
    if sensor.endswith("_loop"):
        sensor, loop_closure = sensor[:-5], True
    n, h, rx, ry, bottom, gsi, ring = _SENSORS[sensor]
    p = LegoParams()
    p.N_SCAN, p.Horizon_SCAN = n, h
    p.ang_res_x, p.ang_res_y, p.ang_bottom = rx, ry, bottom
    p.groundScanInd, p.useCloudRing = gsi, ring
    p.sensorMinimumRange = 1.0                      # utility.h:111
    p.sensorMountAngle = 0.0                        # utility.h:112
    p.segmentTheta = 60.0 / 180.0 * math.pi         # utility.h:113
    p.segmentValidPointNum = 5                      # utility.h:114
    p.segmentValidLineNum = 3                       # utility.h:115
    p.segmentAlphaX = float(p.ang_res_x) / 180.0 * math.pi   # utility.h:116 (float / double math, narrowed)
    p.segmentAlphaY = float(p.ang_res_y) / 180.0 * math.pi   # utility.h:117
    p.scanPeriod = 0.1                              # utility.h:107
    p.edgeFeatureNum, p.surfFeatureNum, p.sectionsTotal = 2, 4, 6   # utility.h:120-122
    p.edgeThreshold, p.surfThreshold = 0.1, 0.1     # utility.h:123-124
    p.nearestFeatureSearchSqDist = 25.0             # utility.h:125
    p.surroundingKeyframeSearchRadius = 50.0        # utility.h:129
    p.mappingProcessInterval = 0.3                  # utility.h:105
    p.loopClosureEnableFlag = 1 if loop_closure else 0   # utility.h:104 (false as shipped)
    p.device = device
    p.surroundingKeyframeSearchNum = 50             # utility.h:130
    p.historyKeyframeSearchRadius = 7.0             # utility.h:132
    p.historyKeyframeSearchNum = 25                 # utility.h:133
    p.historyKeyframeFitnessScore = 0.3             # utility.h:134
    p.maxKeyFrames = max_keyframes                  # handle capacity, 0 = default (640)
    return p


SENSORS = tuple(_SENSORS)
