"""Device-resident dedup index: Xet chunk hash -> 64-bit location (include/zest_gpu.h, zg_dedup_*).

Stands in, on the seed/dedup path, for zest's per-hash probes of the chunk cache (storage.hasChunk,
src/storage.zig:137) and the XorbRegistry scan (src/storage.zig:145-228): one launch answers a whole batch.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native
from ._native import check
from .context import Context, default_context

MISSING = np.uint64(0xFFFFFFFFFFFFFFFF)


class DedupIndex:
    """`max_entries` distinct hashes at most.  Per hash the index keeps the smallest location ever inserted."""

    def __init__(self, max_entries: int, ctx: Context | None = None):
        self.ctx = ctx or default_context()
        self._h = C.c_void_p()
        check(_native.lib().zg_dedup_create(self.ctx.handle, int(max_entries), C.byref(self._h)))

    @property
    def handle(self):
        return self._h

    def insert(self, hashes, locations) -> np.ndarray:
        """-> was_new (bool per input): the hash was absent before this call and this occurrence's location was kept."""
        h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
        loc = np.ascontiguousarray(locations, dtype=np.uint64)
        if loc.size != h.shape[0]:
            raise ValueError("hashes and locations differ in length")
        new = np.zeros(h.shape[0], np.uint8)
        check(_native.lib().zg_dedup_insert(self._h, h.ctypes.data, loc.ctypes.data, h.shape[0], new.ctypes.data))
        return new.astype(bool)

    def lookup(self, hashes) -> np.ndarray:
        """-> locations (u64), MISSING where the hash is not in the index."""
        h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
        out = np.empty(h.shape[0], np.uint64)
        check(_native.lib().zg_dedup_lookup(self._h, h.ctypes.data, h.shape[0], out.ctypes.data))
        return out

    def __len__(self) -> int:
        n = C.c_uint64(0)
        check(_native.lib().zg_dedup_count(self._h, C.byref(n)))
        return int(n.value)

    def close(self):
        if self._h:
            _native.lib().zg_dedup_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def shard_of(hashes, n_shards: int) -> np.ndarray:
    """Which of n_shards GPUs owns each hash when the index is split across GPUs: the hash's third u64 modulo
    n_shards (the first two words pick slot and tag inside a shard, so the split leaves them uniform)."""
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    return (h[:, 16:24].copy().view(np.uint64).reshape(-1) % np.uint64(n_shards)).astype(np.int32)
