"""Mirror of zig-xet's `xet.xorb.XorbReader` at the two places zest calls it
(/root/reference/src/xet_bridge.zig:256-257, src/parallel_download.zig:65-66):

    var reader = xorb_mod.XorbReader.init(allocator, result.data);
    const chunk_data = try reader.extractChunkRange(result.local_start, result.local_end);

Same names, same argument meaning (chunk indices relative to the blob actually fetched), errors raised instead
of returned.  The decode runs on the GPU through zg_extract_chunk_range."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _native
from ._native import check
from .context import Context, default_context
from .hashing import _u8


@dataclass
class XorbIndex:
    hdr_off: np.ndarray  # u64: byte offset of each chunk's 8-byte header
    clen: np.ndarray     # u32: compressed length
    ulen: np.ndarray     # u32: uncompressed length
    scheme: np.ndarray   # u8: 0 none, 1 lz4, 2 bg4-lz4
    total_ulen: int

    @property
    def n_chunks(self) -> int:
        return int(self.clen.size)


def xorb_index(blob) -> XorbIndex:
    """Host walk of the chunk stream (zg_xorb_index)."""
    b = _u8(blob)
    L = _native.lib()
    n = C.c_uint32()
    tot = C.c_uint64()
    check(L.zg_xorb_index(b.ctypes.data, b.size, 0, C.byref(n), None, None, None, None, C.byref(tot)))
    k = n.value
    ho, cl, ul, sc = np.empty(k, np.uint64), np.empty(k, np.uint32), np.empty(k, np.uint32), np.empty(k, np.uint8)
    check(L.zg_xorb_index(b.ctypes.data, b.size, k, C.byref(n), ho.ctypes.data, cl.ctypes.data, ul.ctypes.data,
                          sc.ctypes.data, C.byref(tot)))
    return XorbIndex(ho, cl, ul, sc, int(tot.value))


class XorbReader:
    """XorbReader.init(allocator, data) -- the allocator becomes the (optional) device context."""

    def __init__(self, data, ctx: Context | None = None):
        self.data = _u8(data)
        self.ctx = ctx or default_context()

    @classmethod
    def init(cls, data, ctx: Context | None = None) -> "XorbReader":
        return cls(data, ctx)

    def extract_chunk_range(self, start: int, end: int) -> bytes:
        L = _native.lib()
        need = C.c_uint64()
        check(L.zg_chunk_range_size(self.data.ctypes.data, self.data.size, start, end, C.byref(need)))
        out = np.empty(max(int(need.value), 1), np.uint8)
        got = C.c_uint64()
        check(L.zg_extract_chunk_range(self.ctx.handle, self.data.ctypes.data, self.data.size, start, end,
                                       out.ctypes.data, int(need.value), C.byref(got)))
        return out[:got.value].tobytes()

    extractChunkRange = extract_chunk_range  # the reference's spelling


def xorb_footer(xorb_hash, chunk_hashes, clen, ulen) -> bytes:
    """The XETBLOB footer xet-core appends to a stored xorb (zg_xorb_footer)."""
    h = _u8(xorb_hash)
    ch = np.ascontiguousarray(chunk_hashes, dtype=np.uint8).reshape(-1, 32)
    cl = np.ascontiguousarray(clen, dtype=np.uint32)
    ul = np.ascontiguousarray(ulen, dtype=np.uint32)
    n = cl.size
    L = _native.lib()
    out = np.empty(L.zg_xorb_footer_bytes(n), np.uint8)
    got = C.c_size_t()
    check(L.zg_xorb_footer(h.ctypes.data, ch.ctypes.data, cl.ctypes.data, ul.ctypes.data, n, out.ctypes.data, out.size,
                           C.byref(got)))
    return out[:got.value].tobytes()


def seal_xorb(stream, ctx: Context | None = None) -> tuple[bytes, bytes]:
    """Seed path: chunk stream -> (XorbCache-format blob = stream + XETBLOB footer, xorb hash).  The chunk hashes and
    the Merkle root come from the GPU (verify_xorbs without an expected hash)."""
    from .verify import verify_xorbs
    b = _u8(stream)
    ix = xorb_index(b)
    res = verify_xorbs([b], None, want_chunk_hashes=True, ctx=ctx)
    if res.verdict[0] < 0:
        raise _native.ZestGpuError(int(res.verdict[0]), "chunk stream does not decode")
    root = res.roots[0].tobytes()
    return b.tobytes() + xorb_footer(root, res.chunk_hashes, ix.clen, ix.ulen), root
