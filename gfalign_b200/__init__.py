"""gfalign_b200 — B200-native engine for gfalign's path-scoring hot path.

Host-side Python here is plumbing only (ctypes binding of the C ABI in include/gfalign_b200.h,
synthetic workload generators, torch.distributed glue for alignment shards).  All scoring runs in
the hand-written sm_100a kernels of libgfalign_b200.so; importing this package without that
library, or creating an Engine without a CUDA device, raises — there is no CPU fallback.
"""
from .capi import (Engine, GfeError, lib_path, load_library, SearchRow,  # noqa: F401
                   search_with_scorer)

__all__ = ["Engine", "GfeError", "lib_path", "load_library", "SearchRow", "search_with_scorer"]
