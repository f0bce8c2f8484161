"""ctypes front end of the CPU oracle (oracle/xet_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product package zest_b200 never does.  It also hosts the synthetic-input generators
(Zipf text, Gaussian bf16, xorb builder) that tests and bench.py share, because they need the oracle's
CDC and LZ4 framing to make valid xorbs.

Algorithm and provenance: see the header of xet_oracle.c (reference call sites
src/xet_bridge.zig:256-257, src/parallel_download.zig:65-66, src/bench.zig:207-223).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libxet_oracle.so")

ERR_NAMES = {
    0: "OK", -3: "TRUNCATED", -4: "BAD_HEADER", -5: "SCHEME", -6: "LZ4_FRAME", -7: "LZ4_BLOCK",
    -8: "SIZE_MISMATCH", -9: "RANGE", -10: "CAPACITY", -20: "NOLZ4",
}


class OracleError(Exception):
    def __init__(self, code: int):
        super().__init__(f"oracle error {code} ({ERR_NAMES.get(code, '?')})")
        self.code = code


def build(force: bool = False) -> str:
    """Compile libxet_oracle.so with the committed Makefile if it is missing or stale."""
    src = os.path.join(_HERE, "xet_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, u64p, u32p = C.c_void_p, C.c_void_p, C.c_void_p
        L.oz_chunk_hash.argtypes = [u8p, C.c_size_t, u8p]
        L.oz_blake3_keyed.argtypes = [u8p, u8p, C.c_size_t, u8p]
        L.oz_blake3.argtypes = [u8p, C.c_size_t, u8p]
        L.oz_blake3_compressions.argtypes = [C.c_uint64]
        L.oz_blake3_compressions.restype = C.c_uint64
        L.oz_hash_to_hex.argtypes = [u8p, C.c_char_p]
        L.oz_merkle_root.argtypes = [u8p, u64p, C.c_size_t, u8p]
        L.oz_file_hash.argtypes = [u8p, u64p, C.c_size_t, u8p]
        L.oz_range_hashes.argtypes = [u8p, u32p, C.c_size_t, u8p]
        L.oz_xxh32.argtypes = [u8p, C.c_size_t, C.c_uint32]
        L.oz_xxh32.restype = C.c_uint32
        L.oz_xorb_index.argtypes = [u8p, C.c_size_t, C.c_size_t, u64p, u32p, u32p, u8p, u64p]
        L.oz_xorb_index.restype = C.c_int64
        L.oz_decode_chunk.argtypes = [u8p, C.c_uint32, C.c_uint8, u8p, C.c_uint32]
        L.oz_extract_chunk_range.argtypes = [u8p, C.c_size_t, C.c_uint32, C.c_uint32, u8p, C.c_size_t]
        L.oz_extract_chunk_range.restype = C.c_int64
        L.oz_cdc.argtypes = [u8p, C.c_size_t, u64p, C.c_size_t]
        L.oz_cdc.restype = C.c_int64
        L.oz_hash_chunks_mt.argtypes = [u8p, u64p, C.c_size_t, u8p, C.c_int]
        L.oz_verify_xorbs_mt.argtypes = [C.c_void_p, C.c_void_p, u8p, C.c_size_t, C.c_void_p, u8p, u8p, C.c_int]
        L.oz_verify_xorbs_mt.restype = C.c_int64
        L.oz_build_xorb.argtypes = [u8p, u64p, C.c_size_t, C.c_int, C.c_size_t, C.c_int, u8p, C.c_size_t]
        L.oz_build_xorb.restype = C.c_int64
        L.oz_synth_bf16_q16.argtypes = [u8p, C.c_size_t, C.c_uint64, C.c_float]
        L.oz_synth_bf16_q16.restype = None
        L.oz_set_lz4_decoder.argtypes = [C.c_int]
        _lib = L
    return _lib


def _u8(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(a, dtype=np.uint8)
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


def _p(a: np.ndarray):
    return a.ctypes.data


# ---------------------------------------------------------------- hashing
def chunk_hash(data) -> bytes:
    d = _u8(data)
    out = np.empty(32, np.uint8)
    lib().oz_chunk_hash(_p(d), d.size, _p(out))
    return out.tobytes()


def blake3_keyed(key: bytes, data) -> bytes:
    d, k = _u8(data), _u8(key)
    out = np.empty(32, np.uint8)
    lib().oz_blake3_keyed(_p(k), _p(d), d.size, _p(out))
    return out.tobytes()


def blake3(data) -> bytes:
    d = _u8(data)
    out = np.empty(32, np.uint8)
    lib().oz_blake3(_p(d), d.size, _p(out))
    return out.tobytes()


def blake3_compressions(length: int) -> int:
    return int(lib().oz_blake3_compressions(length))


def hash_to_hex(h: bytes) -> str:
    buf = C.create_string_buffer(65)
    hh = _u8(h)
    lib().oz_hash_to_hex(_p(hh), buf)
    return buf.value.decode()


def hex_to_hash(s: str) -> bytes:
    assert len(s) == 64
    return b"".join(int(s[16 * g:16 * g + 16], 16).to_bytes(8, "little") for g in range(4))


def hash_chunks(data, offsets, threads: int = 1) -> np.ndarray:
    """Keyed chunk hashes of data[offsets[i]:offsets[i+1]] -> (n, 32) uint8."""
    d = _u8(data)
    off = np.ascontiguousarray(offsets, dtype=np.uint64)
    n = off.size - 1
    out = np.empty((max(n, 0), 32), np.uint8)
    if n > 0:
        lib().oz_hash_chunks_mt(_p(d), _p(off), n, _p(out), threads)
    return out


def merkle_root(hashes, lens) -> bytes:
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    l = np.ascontiguousarray(lens, dtype=np.uint64)
    assert h.shape[0] == l.size
    out = np.empty(32, np.uint8)
    lib().oz_merkle_root(_p(h), _p(l), l.size, _p(out))
    return out.tobytes()


def file_hash(hashes, lens) -> bytes:
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    l = np.ascontiguousarray(lens, dtype=np.uint64)
    out = np.empty(32, np.uint8)
    lib().oz_file_hash(_p(h), _p(l), l.size, _p(out))
    return out.tobytes()


def range_hashes(hashes, bounds) -> np.ndarray:
    """Term verification hashes: one per [bounds[t], bounds[t+1]) run of chunk hashes -> (n_terms, 32) u8."""
    h = np.ascontiguousarray(np.asarray(hashes, np.uint8).reshape(-1, 32))
    b = np.ascontiguousarray(np.asarray(bounds, np.uint32))
    out = np.empty((max(b.size - 1, 0), 32), np.uint8)
    if out.shape[0]:
        lib().oz_range_hashes(_p(h), _p(b), out.shape[0], _p(out))
    return out


# ---------------------------------------------------------------- xorb stream
@dataclass
class XorbIndex:
    hdr_off: np.ndarray  # u64, offset of each chunk's 8-byte header
    clen: np.ndarray     # u32
    ulen: np.ndarray     # u32
    scheme: np.ndarray   # u8
    total_ulen: int

    @property
    def n(self) -> int:
        return int(self.clen.size)


def xorb_index(blob) -> XorbIndex:
    b = _u8(blob)
    n = lib().oz_xorb_index(_p(b), b.size, 0, None, None, None, None, None)
    if n < 0:
        raise OracleError(int(n))
    ho, cl, ul, sc = (np.empty(n, np.uint64), np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.uint8))
    tot = C.c_uint64(0)
    r = lib().oz_xorb_index(_p(b), b.size, n, _p(ho), _p(cl), _p(ul), _p(sc), C.addressof(tot))
    assert r == n
    return XorbIndex(ho, cl, ul, sc, int(tot.value))


def extract_chunk_range(blob, start: int, end: int) -> bytes:
    b = _u8(blob)
    idx = xorb_index(b)
    if end > idx.n or start > end:
        raise OracleError(-9)
    cap = int(idx.ulen[start:end].sum())
    out = np.empty(max(cap, 1), np.uint8)
    r = lib().oz_extract_chunk_range(_p(b), b.size, start, end, _p(out), cap)
    if r < 0:
        raise OracleError(int(r))
    return out[:r].tobytes()


def verify_xorbs(blobs, expected=None, threads: int = 1, want_chunk_hashes: bool = False):
    """-> (verdict int8[n], roots (n,32) u8, bytes_verified[, chunk_hashes])."""
    arrs = [_u8(b) for b in blobs]
    n = len(arrs)
    ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
    lens = (C.c_size_t * max(n, 1))(*[a.size for a in arrs])
    verdict = np.zeros(max(n, 1), np.int8)
    roots = np.zeros((max(n, 1), 32), np.uint8)
    exp = None
    if expected is not None:
        if isinstance(expected, (list, tuple)):
            expected = np.frombuffer(b"".join(bytes(e) for e in expected), np.uint8)
        exp = np.ascontiguousarray(expected, dtype=np.uint8).reshape(n, 32)
    ch = None
    if want_chunk_hashes:
        total = 0
        for a in arrs:
            try:
                total += xorb_index(a).n
            except OracleError:
                pass
        ch = np.zeros((max(total, 1), 32), np.uint8)
    r = lib().oz_verify_xorbs_mt(ptrs, lens, _p(exp) if exp is not None else None, n, _p(verdict), _p(roots),
                                 _p(ch) if ch is not None else None, threads)
    res = (verdict[:n], roots[:n], int(r))
    return res + (ch,) if want_chunk_hashes else res


# ---------------------------------------------------------------- CDC
def cdc(data) -> np.ndarray:
    """Chunk end offsets (exclusive, cumulative) as u64."""
    d = _u8(data)
    cap = d.size // 8192 + 2
    ends = np.empty(cap, np.uint64)
    n = lib().oz_cdc(_p(d), d.size, _p(ends), cap)
    if n < 0:
        raise OracleError(int(n))
    return ends[:n].copy()


# ---------------------------------------------------------------- synthetic inputs + builder
XORB_MAX_BYTES = 64 * 1024 * 1024
XORB_MAX_CHUNKS = 8192


FF_CSIZE, FF_BSUM, FF_CSUM, FF_LINKED = 1, 2, 4, 8


def build_xorb(data, ends, policy: int, block_max: int = 0, fflags: int = 0) -> bytes:
    """Serialize chunks of `data` cut at `ends` into one xorb chunk stream (policy 0 none, 1 lz4, 2 bg4)."""
    d = _u8(data)
    e = np.ascontiguousarray(ends, dtype=np.uint64)
    cap = d.size + e.size * 96 + d.size // 50 + 4096
    out = np.empty(cap, np.uint8)
    r = lib().oz_build_xorb(_p(d), _p(e), e.size, policy, block_max, fflags, _p(out), cap)
    if r < 0:
        raise OracleError(int(r))
    return out[:r].tobytes()


def build_xorb_into(data, ends, policy: int, out: np.ndarray) -> int:
    """build_xorb straight into `out` (a uint8 array, e.g. a slice of pinned memory).  -> blob size."""
    d = _u8(data)
    e = np.ascontiguousarray(ends, dtype=np.uint64)
    r = lib().oz_build_xorb(_p(d), _p(e), e.size, policy, 0, 0, out.ctypes.data, out.size)
    if r < 0:
        raise OracleError(int(r))
    return int(r)


def set_lz4_decoder(use_liblz4: bool) -> bool:
    """Timed CPU arm only: hand independent LZ4 blocks to liblz4's LZ4_decompress_safe (True) or keep the plain
    decoder of xet_oracle.c (False, the default and what the parity tests use).  -> whether liblz4 is in use."""
    return bool(lib().oz_set_lz4_decoder(1 if use_liblz4 else 0))


def split_into_xorbs(ends: np.ndarray):
    """Group consecutive chunks into xorbs of <= 64 MiB and <= 8192 chunks (A.4).  -> list of (lo, hi)."""
    groups, lo, base = [], 0, 0
    for i in range(ends.size):
        if i > lo and (int(ends[i]) - base > XORB_MAX_BYTES or i - lo >= XORB_MAX_CHUNKS):
            groups.append((lo, i))
            base = int(ends[i - 1])
            lo = i
    if ends.size > lo:
        groups.append((lo, int(ends.size)))
    return groups


@dataclass
class SynthXorb:
    blob: bytes
    expected: bytes          # 32-byte xorb hash (Merkle root of the chunk hashes)
    n_chunks: int
    ulen_total: int


def make_xorbs(data, policy: int, threads: int = 1, block_max: int = 0, fflags: int = 0):
    """CDC-chunk `data`, hash, and pack into xorbs.  -> (list[SynthXorb], ends u64, chunk_hashes (n,32))."""
    d = _u8(data)
    ends = cdc(d)
    off = np.concatenate([np.zeros(1, np.uint64), ends])
    hashes = hash_chunks(d, off, threads)
    lens = np.diff(off)
    xs = []
    for lo, hi in split_into_xorbs(ends):
        b0 = int(off[lo])
        rel = (ends[lo:hi] - np.uint64(b0)).astype(np.uint64)
        blob = build_xorb(d[b0:int(ends[hi - 1])], rel, policy, block_max, fflags)
        xs.append(SynthXorb(blob, merkle_root(hashes[lo:hi], lens[lo:hi]), hi - lo, int(ends[hi - 1]) - b0))
    return xs, ends, hashes


def synth_random(nbytes: int, seed: int) -> np.ndarray:
    """Config 1: incompressible bytes."""
    return np.random.default_rng(seed).integers(0, 256, nbytes, dtype=np.uint8)


def synth_zipf_text(nbytes: int, seed: int) -> np.ndarray:
    """Config 2: space-joined words drawn Zipf(1/rank) from 4,096 random lowercase words of length 2-9
    (SURVEY.md 8d row 2); LZ4-compressible to about 0.62."""
    rng = np.random.default_rng(seed)
    V = 4096
    wl = rng.integers(2, 10, V)
    letters = rng.integers(97, 123, int(wl.sum()), dtype=np.uint8)
    starts = np.concatenate([[0], np.cumsum(wl)[:-1]])
    p = 1.0 / np.arange(1, V + 1)
    p /= p.sum()
    out = np.empty(nbytes + 16, np.uint8)
    pos = 0
    # vectorised: draw a block of words, expand to bytes with a gather
    while pos < nbytes:
        k = max(1024, (nbytes - pos) // 5)
        w = rng.choice(V, size=k, p=p)
        ln = wl[w] + 1
        cs = np.cumsum(ln)
        tot = int(cs[-1])
        seg = np.full(tot, 32, np.uint8)
        wstart = cs - ln
        idx = np.repeat(np.arange(k), wl[w])
        within = np.arange(int(wl[w].sum())) - np.repeat(np.cumsum(wl[w]) - wl[w], wl[w])
        seg[np.repeat(wstart, wl[w]) + within] = letters[np.repeat(starts[w], wl[w]) + within]
        del idx
        take = min(tot, nbytes - pos)
        out[pos:pos + take] = seg[:take]
        pos += take
    return out[:nbytes]


def synth_bf16_fast(nbytes: int, seed: int, sigma: float = 0.02, out: np.ndarray | None = None) -> np.ndarray:
    """Configs 3-5 at full size: the same kind of bytes as synth_bf16 (Gaussian float32 >> 16) from the oracle's
    table-driven generator (oz_synth_bf16_q16, ~1 ns per value, GIL released), so that 16-140 GB sets are built in
    seconds.  Different stream than synth_bf16; BG4-LZ4 ratio 0.862 for both."""
    if out is None:
        out = np.empty(nbytes, np.uint8)
    lib().oz_synth_bf16_q16(out.ctypes.data, nbytes, seed, sigma)
    return out[:nbytes]


def synth_xorb(kind: str, raw_bytes: int, seed: int, policy: int, out: np.ndarray | None = None):
    """One synthetic xorb from one raw piece (<= 64 MiB): generate, CDC, chunk hashes, Merkle root, serialize.
    kind: "bf16" | "random" | "text".  -> (blob uint8 array or size written into `out`, expected 32 bytes,
    chunk_hashes (n,32), ulens u32[n]).  Single-threaded and GIL-free in the heavy parts: run many in a thread pool."""
    assert raw_bytes <= XORB_MAX_BYTES
    if kind == "bf16":
        d = synth_bf16_fast(raw_bytes, seed)
    elif kind == "random":
        d = synth_random(raw_bytes, seed)
    else:
        d = synth_zipf_text(raw_bytes, seed)
    ends = cdc(d)
    assert ends.size <= XORB_MAX_CHUNKS
    off = np.concatenate([np.zeros(1, np.uint64), ends])
    hashes = hash_chunks(d, off, 1)
    lens = np.diff(off)
    expected = merkle_root(hashes, lens)
    if out is None:
        blob = np.frombuffer(build_xorb(d, ends, policy), np.uint8)
    else:
        blob = build_xorb_into(d, ends, policy, out)
    return blob, expected, hashes, lens.astype(np.uint32)


def synth_bf16(nbytes: int, seed: int, sigma: float = 0.02) -> np.ndarray:
    """Configs 3-5: float32 N(0, sigma) truncated to its top 16 bits."""
    rng = np.random.default_rng(seed)
    n = (nbytes + 1) // 2
    x = (rng.standard_normal(n, dtype=np.float32) * np.float32(sigma)).view(np.uint32) >> np.uint32(16)
    return x.astype(np.uint16).view(np.uint8)[:nbytes]
