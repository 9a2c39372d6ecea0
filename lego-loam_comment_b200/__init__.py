"""lego-loam_comment_b200: B200-native (sm_100a) per-scan hot path of LeGO-LOAM behind a C ABI.

Python here is glue only: ctypes bindings over `liblego_b200.so` (include/lego_b200.h), parameter structs that
mirror the reference's utility.h names, and the synthetic scan generator used by tests and bench."""
from .params import LegoParams, make_params, SENSORS  # noqa: F401
