"""starrocks_b200 — B200-native scan -> filter -> hash-aggregate / hash-join operators behind the StarRocks BE
pipeline Operator surface. The compute lives in csrc/ (hand-written sm_100a CUDA behind the C ABI of
include/srb/srb_c_api.h); this package is the ctypes binding and the host-side mirror of the operator interface."""
from . import capi, plan  # noqa: F401

__all__ = ["capi", "plan"]
