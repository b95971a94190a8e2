"""reth_b200: B200-native prioritized experience replay behind the reth_buffer surface.

The compute path is the CUDA library reth_b200/_lib/libreth_replay.so (C ABI in
include/reth_replay.h); this package is the host-side mirror of the reference's
Python interface.  There is no CPU fallback.
"""
from ._cabi import ReplayError, build  # noqa: F401
from .shard import ReplayShard  # noqa: F401

__all__ = ["ReplayShard", "ReplayError", "build"]
