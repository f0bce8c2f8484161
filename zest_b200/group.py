"""All GPUs of one box behind one call (zg_group_*, include/zest_gpu.h).

zest is a single process -- one XetBridge reconstructs a file (/root/reference/src/xet_bridge.zig:231-264) and one task
group fans its terms out (src/parallel_download.zig:91-134) -- so the drop-in for a multi-GPU box is one object that
owns a context and a worker thread per device and splits each verify batch over them (greedy longest-first by
compressed bytes, no data moves between GPUs, results come back in the caller's order)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native
from ._native import check
from .context import Context
from .hashing import _u8
from .verify import VerifyResult


def partition_native(sizes, n_parts: int) -> list[list[int]]:
    """zg_partition_xorbs: the split zg_group_verify_xorbs uses (same rule as verify.partition_xorbs)."""
    n = len(sizes)
    lens = (C.c_size_t * max(n, 1))(*[int(s) for s in sizes])
    part = np.zeros(max(n, 1), np.int32)
    check(_native.lib().zg_partition_xorbs(lens, n, n_parts, part.ctypes.data))
    out = [[] for _ in range(n_parts)]
    for i in range(n):
        out[int(part[i])].append(i)
    return out


class Group:
    def __init__(self, devices=None, n_lanes: int = 0):
        L = _native.lib()
        self._h = C.c_void_p()

        class Opts(C.Structure):
            _fields_ = [("struct_size", C.c_uint32), ("n_lanes", C.c_uint32), ("reserved", C.c_uint64 * 3)]

        o = Opts(C.sizeof(Opts), n_lanes)
        if devices is None:
            check(L.zg_group_init(None, 0, C.byref(o), C.byref(self._h)))
        else:
            arr = (C.c_int * len(devices))(*devices)
            check(L.zg_group_init(arr, len(devices), C.byref(o), C.byref(self._h)))
        self.size = int(L.zg_group_size(self._h))

    def context(self, i: int) -> Context:
        """The i-th device's context (owned by the group: do not close it)."""
        h = _native.lib().zg_group_ctx(self._h, i)
        if not h:
            raise IndexError(i)
        return Context.borrowed(C.c_void_p(h))

    def verify_xorbs(self, blobs, expected=None, want_chunk_hashes: bool = False) -> VerifyResult:
        """verify.verify_xorbs over every device of the group; same result object, caller's order."""
        arrs = [_u8(b) for b in blobs]
        n = len(arrs)
        ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
        lens = (C.c_size_t * max(n, 1))(*[a.size for a in arrs])
        verdict = np.zeros(n, np.int8)
        roots = np.zeros((n, 32), np.uint8)
        exp = None
        if expected is not None:
            if isinstance(expected, (list, tuple)):
                expected = np.frombuffer(b"".join(bytes(e) for e in expected), np.uint8)
            exp = np.ascontiguousarray(expected, dtype=np.uint8).reshape(n, 32)
        ch, cap = None, 0
        if want_chunk_hashes:
            from .xorb import xorb_index
            for a in arrs:
                try:
                    cap += xorb_index(a).n_chunks
                except _native.ZestGpuError:
                    pass
            ch = np.zeros((cap, 32), np.uint8)
        nb = C.c_uint64()
        check(_native.lib().zg_group_verify_xorbs(self._h, ptrs, lens, exp.ctypes.data if exp is not None else None, n,
                                                  verdict.ctypes.data, roots.ctypes.data,
                                                  ch.ctypes.data if ch is not None and cap else None, cap, C.byref(nb)))
        return VerifyResult(verdict, roots, int(nb.value), ch)

    def close(self):
        if self._h:
            _native.lib().zg_group_shutdown(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
