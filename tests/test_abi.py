"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol that
include/lego_b200.h declares, and its entry points fail loudly (error code + message, no crash, no CPU fallback) when
no CUDA device is present."""
import ctypes
import os
import re

import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from lego_loam_b200 import capi
from lego_loam_b200.params import LegoParams, make_params

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "lego_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lego_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported():
    lib = capi.load()
    declared = _declared_symbols()
    assert len(declared) >= 25
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, "declared in include/lego_b200.h but not exported: %s" % missing
    assert sorted(capi.EXPORTED) == declared, "capi.EXPORTED out of sync with the header"


def test_params_struct_matches_header():
    """ctypes mirror of lego_params: same field names, same order as the C struct."""
    src = open(os.path.join(ROOT, "include", "lego_b200.h")).read()
    body = src[src.index("typedef struct lego_params {"):src.index("} lego_params;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.replace("typedef struct lego_params {", "").strip()
        if not decl:
            continue
        first, *rest = decl.split(",")
        names.append(first.split()[-1])
        names.extend(r.strip() for r in rest)
    assert names == [f[0] for f in LegoParams._fields_]


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    lib = capi.load()
    assert lib.lego_device_count() == 0
    h = ctypes.c_void_p()
    p = make_params("vlp16")
    rc = lib.lego_pipeline_create(ctypes.byref(p), 1, 0, ctypes.byref(h))
    assert rc == -2 and not h.value          # LEGO_ERR_CUDA: no CPU fallback
    assert lib.lego_last_error()
    for create in (lib.lego_ip_create, lib.lego_fa_create, lib.lego_mo_create):
        create.argtypes = [ctypes.POINTER(LegoParams), ctypes.POINTER(ctypes.c_void_p)]
        assert create(ctypes.byref(p), ctypes.byref(h)) == -2
    with pytest.raises(capi.LegoError):
        capi.Pipeline(p, lanes=1)


def test_argument_errors():
    lib = capi.load()
    h = ctypes.c_void_p()
    p = make_params("vlp16")
    assert lib.lego_pipeline_create(None, 1, 0, ctypes.byref(h)) == -1
    assert lib.lego_pipeline_create(ctypes.byref(p), 0, 0, ctypes.byref(h)) == -1
    p.N_SCAN = 1
    assert lib.lego_pipeline_create(ctypes.byref(p), 1, 0, ctypes.byref(h)) == -1
    assert lib.lego_pipeline_destroy(None) == -1


def test_sensor_parameter_tables():
    """utility.h values: segmentAlphaX/Y derive from ang_res_x/y exactly as the header's expressions do."""
    for s in ("vlp16", "hdl32", "hdl64", "vls128"):
        p = make_params(s)
        assert p.N_SCAN * p.Horizon_SCAN > 0
        assert np.float32(p.segmentAlphaX) == np.float32(float(np.float32(p.ang_res_x)) / 180.0 * np.pi)
        assert np.float32(p.segmentAlphaY) == np.float32(float(np.float32(p.ang_res_y)) / 180.0 * np.pi)
    p = make_params("vlp16")
    assert (p.N_SCAN, p.Horizon_SCAN, p.groundScanInd) == (16, 1800, 7)
    assert abs(p.ang_bottom - 15.1) < 1e-6 and p.segmentValidPointNum == 5 and p.segmentValidLineNum == 3
