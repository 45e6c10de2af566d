"""kmeans_b200 — B200-native Lloyd iteration behind a C ABI (libkmeans.so).

Only what the hot path needs lives here: csrc/ (CUDA kernels for sm_100a + the C ABI),
the ctypes binding, the host-side mirror of the JS facade, and shard arithmetic.
"""
from . import _lib
from . import board_json
from .api import (KMeansError, KMeansState, assign, export_board, host_alloc, host_free, import_board, init, run, step,
                  update)
from .shard import shard_range

__all__ = ["KMeansError", "KMeansState", "init", "assign", "update", "step", "run", "host_alloc", "host_free",
           "shard_range", "_lib", "export_board", "import_board", "board_json"]
