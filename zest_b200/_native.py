"""ctypes binding of libzest_gpu.so (the C ABI declared in include/zest_gpu.h).

This is the only place the shared library is loaded.  There is no Python or CPU implementation behind it:
if the library is missing, or no CUDA device is usable, the calls raise -- they never fall back.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ZEST_GPU_LIB") or os.path.join(_PKG, "libzest_gpu.so")  # override: developer A/B builds
CSRC = os.path.join(_PKG, "csrc")

ERR_NAMES = {
    0: "OK", -1: "INVALID_ARG", -2: "CUDA", -3: "TRUNCATED", -4: "BAD_HEADER", -5: "SCHEME", -6: "LZ4_FRAME",
    -7: "LZ4_BLOCK", -8: "SIZE_MISMATCH", -9: "RANGE", -10: "CAPACITY", -11: "NO_DEVICE", -12: "OOM", -13: "LIMIT",
}

# every symbol include/zest_gpu.h declares (tests check the library exports exactly these)
SYMBOLS = [
    "zg_init", "zg_shutdown", "zg_strerror", "zg_last_error", "zg_pinned_alloc", "zg_pinned_free", "zg_hash_to_hex",
    "zg_hex_to_hash", "zg_xorb_index", "zg_chunk_range_size", "zg_extract_chunk_range", "zg_verify_xorbs",
    "zg_hash_chunks", "zg_range_hashes", "zg_merkle_root", "zg_cdc", "zg_cdc_hash", "zg_dev_hash_scratch_bytes", "zg_dev_hash_chunks",
    "zg_dev_decode_scratch_bytes", "zg_dev_decode_chunks", "zg_dev_merkle_scratch_bytes", "zg_dev_merkle",
    "zg_dev_verify_scratch_bytes", "zg_dev_verify_xorbs", "zg_dev_cdc_scratch_bytes", "zg_dev_cdc", "zg_get_stats",
    "zg_chunk_order", "zg_dev_hash_chunks_static", "zg_xorb_footer_bytes", "zg_xorb_footer", "zg_dev_int_pipe_peak",
    "zg_dedup_create", "zg_dedup_destroy", "zg_dedup_insert", "zg_dedup_lookup", "zg_dedup_count", "zg_dev_dedup_insert",
    "zg_dev_dedup_lookup", "zg_group_init", "zg_group_shutdown", "zg_group_size", "zg_group_ctx", "zg_partition_xorbs",
    "zg_group_verify_xorbs", "zg_dev_decode_for_hash", "zg_dev_hash_decoded_scratch_bytes", "zg_dev_hash_decoded",
]


class ZestGpuError(RuntimeError):
    """A non-zero return of the C ABI.  `.code` is the ZG_ERR_* value (mirrors a Zig error at the call site)."""

    def __init__(self, code: int, msg: str = ""):
        super().__init__(f"zest_gpu error {code} ({ERR_NAMES.get(code, '?')}): {msg}")
        self.code = code


class ChunkDesc(C.Structure):
    _fields_ = [("src_off", C.c_uint64), ("dst_off", C.c_uint64), ("clen", C.c_uint32), ("ulen", C.c_uint32),
                ("scheme", C.c_uint32), ("order", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("calls", "kernels_launched", "h2d_bytes", "d2h_bytes", "bytes_decoded",
                                          "chunks_hashed", "ns_h2d", "ns_decode", "ns_hash", "ns_merkle", "ns_d2h")]


def build(verbose: bool = False) -> str:
    """Compile libzest_gpu.so in-tree with nvcc for sm_100a (zest_b200/csrc/Makefile)."""
    subprocess.check_call(["make", "-C", CSRC, "-j8"], stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None
_lock = threading.Lock()


def lib() -> C.CDLL:
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(nvcc, sm_100a).  zest_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        vp, sz, u64, u32, i32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_uint32, C.c_int
        L.zg_init.argtypes = [i32, vp, C.POINTER(vp)]
        L.zg_shutdown.argtypes = [vp]
        L.zg_shutdown.restype = None
        L.zg_strerror.argtypes = [i32]
        L.zg_strerror.restype = C.c_char_p
        L.zg_last_error.restype = C.c_char_p
        L.zg_pinned_alloc.argtypes = [vp, sz, C.POINTER(vp)]
        L.zg_pinned_free.argtypes = [vp, vp]
        L.zg_hash_to_hex.argtypes = [vp, C.c_char_p]
        L.zg_hash_to_hex.restype = None
        L.zg_hex_to_hash.argtypes = [C.c_char_p, vp]
        L.zg_xorb_index.argtypes = [vp, sz, u32, C.POINTER(u32), vp, vp, vp, vp, C.POINTER(u64)]
        L.zg_chunk_range_size.argtypes = [vp, sz, u32, u32, C.POINTER(u64)]
        L.zg_extract_chunk_range.argtypes = [vp, vp, sz, u32, u32, vp, sz, C.POINTER(u64)]
        L.zg_verify_xorbs.argtypes = [vp, vp, vp, vp, sz, vp, vp, vp, sz, C.POINTER(u64)]
        L.zg_hash_chunks.argtypes = [vp, vp, vp, sz, vp, vp]
        L.zg_merkle_root.argtypes = [vp, vp, vp, sz, i32, vp]
        L.zg_range_hashes.argtypes = [vp, vp, vp, sz, vp]
        L.zg_dedup_create.argtypes = [vp, sz, C.POINTER(vp)]
        L.zg_dedup_destroy.argtypes = [vp]
        L.zg_dedup_destroy.restype = None
        L.zg_dedup_insert.argtypes = [vp, vp, vp, sz, vp]
        L.zg_dedup_lookup.argtypes = [vp, vp, sz, vp]
        L.zg_dedup_count.argtypes = [vp, vp]
        L.zg_dev_dedup_insert.argtypes = [vp, vp, vp, sz, vp, vp]
        L.zg_dev_dedup_lookup.argtypes = [vp, vp, sz, vp, vp]
        L.zg_cdc.argtypes = [vp, vp, sz, vp, sz, C.POINTER(sz)]
        L.zg_cdc_hash.argtypes = [vp, vp, sz, vp, sz, C.POINTER(sz), vp, vp]
        L.zg_dev_hash_scratch_bytes.argtypes = [u64, sz]
        L.zg_dev_hash_scratch_bytes.restype = sz
        L.zg_dev_hash_chunks.argtypes = [vp, vp, u64, vp, vp, sz, vp, vp, vp, sz, vp]
        L.zg_dev_decode_scratch_bytes.argtypes = [sz, u64]
        L.zg_dev_decode_scratch_bytes.restype = sz
        L.zg_dev_decode_chunks.argtypes = [vp, vp, u64, vp, sz, vp, u64, vp, vp, sz, vp]
        L.zg_dev_decode_for_hash.argtypes = [vp, vp, u64, vp, sz, vp, u64, vp, vp, sz, vp]
        L.zg_dev_hash_decoded_scratch_bytes.argtypes = [u64, sz]
        L.zg_dev_hash_decoded_scratch_bytes.restype = sz
        L.zg_dev_hash_decoded.argtypes = [vp, vp, u64, vp, sz, vp, vp, vp, sz, vp]
        L.zg_dev_merkle_scratch_bytes.argtypes = [sz, sz]
        L.zg_dev_merkle_scratch_bytes.restype = sz
        L.zg_dev_merkle.argtypes = [vp, vp, vp, vp, sz, sz, vp, vp, vp, vp, vp, sz, vp]
        L.zg_dev_verify_scratch_bytes.argtypes = [sz, sz, u64]
        L.zg_dev_verify_scratch_bytes.restype = sz
        L.zg_dev_verify_xorbs.argtypes = [vp, vp, u64, vp, sz, vp, sz, vp, u64, vp, vp, vp, vp, sz, vp]
        L.zg_dev_cdc_scratch_bytes.argtypes = [u64]
        L.zg_dev_cdc_scratch_bytes.restype = sz
        L.zg_dev_cdc.argtypes = [vp, vp, u64, vp, u64, vp, vp, sz, vp]
        L.zg_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.zg_dev_hash_chunks_static.argtypes = [vp, u64, vp, vp, sz, vp, vp, vp]
        L.zg_xorb_footer_bytes.argtypes = [sz]
        L.zg_xorb_footer_bytes.restype = sz
        L.zg_xorb_footer.argtypes = [vp, vp, vp, vp, sz, vp, sz, C.POINTER(sz)]
        L.zg_dev_int_pipe_peak.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), vp]
        L.zg_group_init.argtypes = [vp, i32, vp, C.POINTER(vp)]
        L.zg_group_shutdown.argtypes = [vp]
        L.zg_group_shutdown.restype = None
        L.zg_group_size.argtypes = [vp]
        L.zg_group_ctx.argtypes = [vp, i32]
        L.zg_group_ctx.restype = vp
        L.zg_partition_xorbs.argtypes = [vp, sz, i32, vp]
        L.zg_group_verify_xorbs.argtypes = [vp, vp, vp, vp, sz, vp, vp, vp, sz, C.POINTER(u64)]
        L.zg_chunk_order.argtypes = [vp, sz]
        L.zg_chunk_order.restype = None
        _lib = L
        return L


def check(code: int):
    if code != 0:
        raise ZestGpuError(code, lib().zg_last_error().decode(errors="replace"))
