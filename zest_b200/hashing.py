"""Mirror of zig-xet's `xet.hashing` as zest uses it (hashToHex at /root/reference/src/xet_bridge.zig:154 and
src/server.zig:201-203; chunk / Merkle hashes per CONTRIBUTING.md:144-146).  All hashing runs on the GPU."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native
from ._native import check
from .context import Context, default_context

PLAIN_BLAKE3 = "plain"


def _u8(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(a, dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def hash_to_hex(h) -> str:
    """32-byte hash -> 64 hex chars, each 8-byte group printed as a little-endian u64 (NOT bytes.hex())."""
    hh = _u8(h)
    if hh.size != 32:
        raise ValueError("hash must be 32 bytes")
    buf = C.create_string_buffer(65)
    _native.lib().zg_hash_to_hex(hh.ctypes.data, buf)
    return buf.value.decode()


def hex_to_hash(s: str) -> bytes:
    out = np.empty(32, np.uint8)
    check(_native.lib().zg_hex_to_hash(s.encode(), out.ctypes.data))
    return out.tobytes()


def hash_chunks(data, offsets, key=None, ctx: Context | None = None) -> np.ndarray:
    """Xet chunk hashes of data[offsets[i]:offsets[i+1]] -> (n, 32) uint8.  key=None: Xet data key;
    key=32 bytes: that BLAKE3 key; key=PLAIN_BLAKE3: unkeyed BLAKE3 (what `zest bench --synthetic` hashes)."""
    ctx = ctx or default_context()
    d = _u8(data)
    off = np.ascontiguousarray(offsets, dtype=np.uint64)
    n = off.size - 1
    if n < 0:
        raise ValueError("offsets needs at least one entry")
    out = np.empty((n, 32), np.uint8)
    if key is None:
        kp = None
    elif isinstance(key, str) and key == PLAIN_BLAKE3:
        kp = C.c_void_p(1)
    else:
        kb = _u8(key)
        if kb.size != 32:
            raise ValueError("key must be 32 bytes")
        kp = C.c_void_p(kb.ctypes.data)
    check(_native.lib().zg_hash_chunks(ctx.handle, d.ctypes.data, off.ctypes.data, n, kp, out.ctypes.data))
    return out


def chunk_hash(data, ctx: Context | None = None) -> bytes:
    d = _u8(data)
    return hash_chunks(d, [0, d.size], ctx=ctx)[0].tobytes()


def blake3(data, ctx: Context | None = None) -> bytes:
    d = _u8(data)
    return hash_chunks(d, [0, d.size], key=PLAIN_BLAKE3, ctx=ctx)[0].tobytes()


def merkle_root(hashes, lens, ctx: Context | None = None) -> bytes:
    """Xorb hash of (chunk hash, chunk length) pairs."""
    return _merkle(hashes, lens, 0, ctx)


def file_hash(hashes, lens, ctx: Context | None = None) -> bytes:
    """Xet file hash: zero-keyed BLAKE3 of the Merkle root over all of a file's chunks."""
    return _merkle(hashes, lens, 1, ctx)


def range_hashes(chunk_hashes, bounds, ctx: Context | None = None) -> np.ndarray:
    """Verification hashes of reconstruction terms: term t covers chunks bounds[t] .. bounds[t+1]-1 of the list.
    -> (n_terms, 32) u8.  Compare with the range hash a Xet shard stores beside each term."""
    ctx = ctx or default_context()
    h = np.ascontiguousarray(chunk_hashes, dtype=np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(bounds, dtype=np.uint32)
    if b.size and int(b[-1]) > h.shape[0]:
        raise ValueError("bounds run past the chunk-hash list")
    out = np.empty((max(b.size - 1, 0), 32), np.uint8)
    if out.shape[0]:
        check(_native.lib().zg_range_hashes(ctx.handle, h.ctypes.data, b.ctypes.data, out.shape[0], out.ctypes.data))
    return out


def _merkle(hashes, lens, file_mode, ctx):
    ctx = ctx or default_context()
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    l = np.ascontiguousarray(lens, dtype=np.uint64)
    if h.shape[0] != l.size:
        raise ValueError("hashes and lens differ in length")
    out = np.empty(32, np.uint8)
    check(_native.lib().zg_merkle_root(ctx.handle, h.ctypes.data, l.ctypes.data, l.size, file_mode, out.ctypes.data))
    return out.tobytes()
