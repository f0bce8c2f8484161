"""Mirror of zig-xet's `xet.chunking` (Gearhash CDC; parameters /root/reference/src/config.zig:12,
DESIGN.md:267-273).  zest has no call site today; the seed path would feed
storage.writeChunk (src/storage.zig:102-107) with the chunks and hashes returned here."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native
from ._native import check
from .context import Context, default_context
from .hashing import _u8

MIN_CHUNK = 8192
TARGET_CHUNK = 65536
MAX_CHUNK = 131072


def cdc(data, ctx: Context | None = None) -> np.ndarray:
    """Chunk end offsets (exclusive, ascending, last == len(data)) as uint64."""
    ctx = ctx or default_context()
    d = _u8(data)
    cap = d.size // MIN_CHUNK + 2
    ends = np.empty(cap, np.uint64)
    n = C.c_size_t()
    check(_native.lib().zg_cdc(ctx.handle, d.ctypes.data, d.size, ends.ctypes.data, cap, C.byref(n)))
    return ends[:n.value].copy()


def cdc_hash(data, want_file_hash: bool = True, ctx: Context | None = None):
    """-> (ends u64[n], chunk hashes (n, 32) u8, file hash bytes | None) in one device pass."""
    ctx = ctx or default_context()
    d = _u8(data)
    cap = d.size // MIN_CHUNK + 2
    ends = np.empty(cap, np.uint64)
    hashes = np.empty((cap, 32), np.uint8)
    fh = np.zeros(32, np.uint8)
    n = C.c_size_t()
    check(_native.lib().zg_cdc_hash(ctx.handle, d.ctypes.data, d.size, ends.ctypes.data, cap, C.byref(n),
                                    hashes.ctypes.data, fh.ctypes.data if want_file_hash else None))
    return ends[:n.value].copy(), hashes[:n.value].copy(), (fh.tobytes() if want_file_hash else None)
