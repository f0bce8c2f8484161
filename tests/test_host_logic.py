"""CPU-only checks of the product's host side: the C-ABI library loads, exports every declared symbol, and its
host-only entry points (hex form, chunk-stream walk, range sizing, sharding) agree with the oracle.  No GPU
compute is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import zest_b200 as z
from zest_b200 import _native
from conftest import ROOT, read_golden_xorb


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "zest_gpu.h")).read()
    declared = re.findall(r"^ZG_API [\w \*]+?\b(zg_\w+)\(", hdr, flags=re.M)
    assert len(declared) >= 25 and len(set(declared)) == len(declared)
    assert set(declared) == set(_native.SYMBOLS)
    L = _native.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_no_cpu_fallback_without_device():
    """On a box without a GPU the context must fail loudly (ZG_ERR_NO_DEVICE); with one it must come up."""
    import torch
    if torch.cuda.is_available():
        z.Context(0).close()
    else:
        with pytest.raises(z.ZestGpuError) as ei:
            z.Context(0)
        assert ei.value.code == -11


def test_hex_form_matches_oracle(oracle):
    rng = np.random.default_rng(3)
    for _ in range(20):
        h = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
        s = z.hash_to_hex(h)
        assert s == oracle.hash_to_hex(h)
        assert z.hex_to_hash(s) == h
    assert z.hash_to_hex(bytes(range(32)))[:16] == "0706050403020100"
    with pytest.raises(z.ZestGpuError):
        z.hex_to_hash("zz" * 32)
    with pytest.raises(z.ZestGpuError):
        z.hex_to_hash("00")


@pytest.mark.parametrize("name", ["xorb_bf16", "xorb_text", "xorb_rand", "xorb_mixed_auto", "xorb_tiny"])
def test_index_matches_oracle(oracle, vectors, name):
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
    stream, _ = read_golden_xorb(rec)
    a, b = z.xorb_index(stream), oracle.xorb_index(stream)
    assert a.n_chunks == b.n == rec["n_chunks"]
    assert (a.hdr_off == b.hdr_off).all() and (a.clen == b.clen).all() and (a.ulen == b.ulen).all()
    assert (a.scheme == b.scheme).all() and a.total_ulen == b.total_ulen
    L = _native.lib()
    need = C.c_uint64()
    arr = np.frombuffer(stream, np.uint8)
    _native.check(L.zg_chunk_range_size(arr.ctypes.data, arr.size, 1 if a.n_chunks > 1 else 0, a.n_chunks, C.byref(need)))
    assert need.value == int(a.ulen[1 if a.n_chunks > 1 else 0:].sum())
    assert L.zg_chunk_range_size(arr.ctypes.data, arr.size, 0, a.n_chunks + 1, C.byref(need)) == -9
    assert L.zg_chunk_range_size(arr.ctypes.data, arr.size, 2, 1, C.byref(need)) == -9


def test_walk_errors_match_oracle(oracle, vectors):
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == "xorb_text")
    stream, _ = read_golden_xorb(rec)
    cases = []
    b = bytearray(stream); b[0] = 7; cases.append(bytes(b))          # version
    b = bytearray(stream); b[4] = 3; cases.append(bytes(b))          # scheme 3 (FullBitsliceLZ4) rejected
    cases.append(stream[:-1]); cases.append(stream[:5]); cases.append(stream + b"\x00\x00")
    b = bytearray(stream); b[4] = 0; cases.append(bytes(b))          # scheme 0 with clen != ulen
    for c in cases:
        with pytest.raises(oracle.OracleError) as eo:
            oracle.xorb_index(c)
        with pytest.raises(z.ZestGpuError) as ez:
            z.xorb_index(c)
        assert ez.value.code == eo.value.code
    assert z.xorb_index(b"").n_chunks == 0


def test_partition_xorbs():
    sizes = [57, 3, 60, 58, 1, 59, 57, 2, 61]
    for n in (1, 2, 4, 8):
        parts = z.partition_xorbs(sizes, n)
        assert sorted(i for p in parts for i in p) == list(range(len(sizes)))
        loads = [sum(sizes[i] for i in p) for p in parts]
        assert max(loads) - min(loads) <= max(sizes)


@pytest.mark.parametrize("name", ["xorb_bf16", "xorb_text", "xorb_rand", "xorb_mixed_auto", "xorb_tiny"])
def test_xetblob_footer_matches_xet_core(vectors, name):
    """zg_xorb_footer reproduces, byte for byte, the footer xet-core wrote for the golden xorbs."""
    from conftest import footer_chunk_hashes
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
    stream, raw = read_golden_xorb(rec)
    hashes, _, xorb_hash = footer_chunk_hashes(raw)
    ix = z.xorb_index(stream)
    assert z.xorb_footer(xorb_hash, hashes, ix.clen, ix.ulen) == raw[len(stream):]


def test_dedup_shard_of_is_balanced_and_stable():
    """Owner GPU of a chunk hash when the dedup index is split across GPUs: a pure function of the hash, uniform."""
    import numpy as np
    from zest_b200.dedup import shard_of
    h = np.random.default_rng(3).integers(0, 256, (80_000, 32), dtype=np.uint8)
    for n in (1, 2, 4, 8):
        s = shard_of(h, n)
        assert s.min() >= 0 and s.max() < n
        counts = np.bincount(s, minlength=n)
        assert counts.min() > 0.9 * h.shape[0] / n
        assert (shard_of(h[::-1], n)[::-1] == s).all()


def test_native_partition_matches_the_python_mirror():
    """zg_partition_xorbs (what zg_group_verify_xorbs splits a batch with) and verify.partition_xorbs (what bench.py
    splits the global set over ranks with) are the same rule: longest first, least-loaded part, ties to the lower index."""
    import numpy as np
    import zest_b200 as z
    rng = np.random.default_rng(5)
    for n, parts in [(0, 3), (1, 1), (7, 2), (50, 8), (239, 8), (5, 8), (64, 4)]:
        sizes = [int(x) for x in rng.integers(1, 1000, n)]
        tied = list(sizes)
        if n > 3:
            tied[1] = tied[2] = tied[0]
        for s in (sizes, tied, [57_800_000] * n):
            got = z.partition_native(s, parts)
            assert got == z.partition_xorbs(s, parts)
            assert sorted(i for p in got for i in p) == list(range(n))
            if n >= parts and len(set(s)) == 1:
                assert max(map(len, got)) - min(map(len, got)) <= 1


def test_bench_descriptors_follow_the_chunk_streams():
    """bench.build_descs (the device-resident leg's chunk descriptors, built from zg_xorb_index on the host): payload
    offsets, lengths and schemes equal the oracle's walk of the same blobs; decoded chunks are laid out back to back on
    16-byte boundaries; `order` is a permutation that puts the longest chunks first (zg_chunk_order)."""
    import numpy as np
    import bench
    from oracle import xet_oracle as oz
    from zest_b200 import _native
    old = bench.XORB_RAW
    bench.XORB_RAW = 1 << 20
    try:
        xs, n_all, comp_all, _ = bench.build_xorb_set("llama8b-bg4", 3 * (1 << 20) + 777, 5, 0, 1, 2, lambda b: np.zeros(b, np.uint8),
                                                      lambda a: a, lambda m: None)
    finally:
        bench.XORB_RAW = old
    d, out_bytes = bench.build_descs(xs, _native.lib())
    k = 0
    for o, v in zip(xs.offsets, xs.views):
        ix = oz.xorb_index(v)
        m = ix.n
        assert np.array_equal(d["src_off"][k:k + m], o + ix.hdr_off + 8)
        assert np.array_equal(d["clen"][k:k + m], ix.clen) and np.array_equal(d["ulen"][k:k + m], ix.ulen)
        assert np.array_equal(d["scheme"][k:k + m], ix.scheme)
        k += m
    assert k == d.size == int(xs.first[-1])
    al = (d["ulen"].astype(np.int64) + 15) // 16 * 16
    assert np.array_equal(d["dst_off"], np.concatenate([[0], np.cumsum(al)[:-1]])) and out_bytes == int(al.sum())
    order = d["order"].astype(np.int64)
    assert sorted(order.tolist()) == list(range(d.size))
    size = d["clen"].astype(np.int64) + d["ulen"].astype(np.int64)
    assert (np.diff(size[order]) <= 0).all()
    assert bench.plan_xorbs(3 * (64 << 20) + 5) == [64 << 20] * 3 + [5] and bench.plan_xorbs(10) == [10]
