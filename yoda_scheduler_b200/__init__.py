"""yoda_scheduler_b200 -- B200-native Filter/Score hot path of the Yoda scheduler plugin.

The product is libyoda_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/yoda_b200.h).  This package is the thin Python host layer used by the tests and
bench: a ctypes binding, a mirror of the plugin surface (plugin.py) and the synthetic
snapshot generator.  There is no CPU fallback: evaluating without the CUDA library or
without a GPU raises.
"""
from .snapshot import Snapshot, NODE_REC, POD_REQ, POD_RES  # noqa: F401
from .labels import go_atoi, str_to_uint64, pods_to_requests  # noqa: F401
