"""CPU oracle for the encoded-vector hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package (see qz_oracle.c header).  `oracle.py` is a ctypes + numpy wrapper around
liboracle.so (our restatement) and _ref/libsimd_utils_ref.so (the reference's own C scorers).
"""
from .oracle import *  # noqa: F401,F403
