"""Parity of the CUDA path against the oracle and the golden vectors, through the C ABI (zest_b200 -> ctypes ->
libzest_gpu.so).  Integer / byte work: every comparison is bit-exact.  Run on the B200 box: pytest -m gpu."""
import hashlib
import os

import numpy as np
import pytest

import zest_b200 as z
from conftest import footer_chunk_hashes, golden_input, read_golden_xorb

pytestmark = pytest.mark.gpu

NODE_KEY = bytes([1, 126, 197, 199, 165, 71, 41, 150, 253, 148, 102, 102, 180, 138, 2, 230, 93, 221, 83, 111, 55, 199,
                  109, 210, 248, 99, 82, 230, 74, 83, 113, 63])


@pytest.fixture(scope="module")
def ctx():
    c = z.Context(0)
    yield c
    c.close()


# ------------------------------------------------------------------------------------------------ BLAKE3 / chunk hash
def test_blake3_kats(ctx, vectors):
    for v in vectors["blake3"]:
        d = bytes([v["fill"]]) * v["len"] if "fill" in v else bytes(i % 251 for i in range(v["len"]))
        assert z.blake3(d, ctx=ctx).hex() == v["plain"], v["len"]
        if "data_key" in v:
            assert z.chunk_hash(d, ctx=ctx).hex() == v["data_key"], v["len"]
            assert z.hash_chunks(d, [0, len(d)], key=NODE_KEY, ctx=ctx)[0].tobytes().hex() == v["node_key"], v["len"]


def test_hello_world(ctx, vectors):
    rec = next(r for r in vectors["xet"] if r["name"] == "hello_world")
    assert z.hash_to_hex(z.chunk_hash(b"hello world", ctx=ctx)) == rec["xorb_hash"]


@pytest.mark.parametrize("seed,total", [(1, 1_200_000), (2, 5_000_000), (3, 150_000_000)])
def test_hash_chunks_vs_oracle(ctx, oracle, seed, total):
    """Ragged chunkings: empty chunks, 1-byte chunks, leaf/block edge lengths, unaligned starts, every group size."""
    rng = np.random.default_rng(seed)
    data = rng.integers(0, 256, total, dtype=np.uint8)
    edge = [0, 0, 1, 63, 64, 65, 1023, 1024, 1025, 2047, 2048, 2049, 4096, 4097, 8191, 8192, 8193, 16384, 16385,
            65536, 65537, 131071, 131072, 3, 0, 5000, 12289]
    lens = list(edge)
    while sum(lens) < total - 200_000:
        lens.append(int(rng.integers(0, 131073)))
    lens.append(total - sum(lens))
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    got = z.hash_chunks(data, off, ctx=ctx)
    want = oracle.hash_chunks(data, off, threads=8)
    assert (got == want).all()
    # a sub-range that starts at an odd offset inside the buffer (offsets[0] != 0)
    k = len(edge)
    assert (z.hash_chunks(data, off[k:k + 9], ctx=ctx) == want[k:k + 8]).all()


def test_hash_large_chunks(ctx, oracle):
    """Chunks far above the CDC maximum (the header allows 16 MiB) take the warp-cooperative tree path."""
    data = np.random.default_rng(9).integers(0, 256, 40_000_000, dtype=np.uint8)
    off = np.array([0, 16_777_215, 16_777_215 + 3_000_001, 40_000_000], np.uint64)
    assert (z.hash_chunks(data, off, ctx=ctx) == oracle.hash_chunks(data, off, threads=4)).all()
    assert z.blake3(data[:20_000_000], ctx=ctx) == oracle.blake3(data[:20_000_000])


# ------------------------------------------------------------------------------------------------ Merkle fold
@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 8, 9, 10, 17, 100, 1060, 2047, 2048, 2049, 4100, 8192, 20011])
def test_merkle_vs_oracle(ctx, oracle, n):
    rng = np.random.default_rng(n)
    hashes = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    lens = rng.integers(0, 131073, n).astype(np.uint64)
    assert z.merkle_root(hashes, lens, ctx=ctx) == oracle.merkle_root(hashes, lens)
    assert z.file_hash(hashes, lens, ctx=ctx) == oracle.file_hash(hashes, lens)


def test_merkle_cut_patterns(ctx, oracle):
    """Force the cut rule's corners: no hash divisible by 4 (groups of 9), all divisible (groups of 3)."""
    n = 500
    rng = np.random.default_rng(0)
    for low in (1, 0):
        hashes = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        hashes[:, 24] = (hashes[:, 24] & 0xFC) | low
        lens = rng.integers(1, 10**9, n).astype(np.uint64)
        assert z.merkle_root(hashes, lens, ctx=ctx) == oracle.merkle_root(hashes, lens)


# ------------------------------------------------------------------------------------------------ decode / extract
@pytest.mark.parametrize("name", ["xorb_bf16", "xorb_text", "xorb_rand", "xorb_mixed_auto", "xorb_tiny"])
def test_extract_real_xorbs(ctx, oracle, vectors, name):
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
    stream, raw = read_golden_xorb(rec)
    gold_hashes, unc_ends, xorb_hash = footer_chunk_hashes(raw)
    reader = z.XorbReader.init(stream, ctx)
    n = rec["n_chunks"]
    full = reader.extract_chunk_range(0, n)
    assert full == oracle.extract_chunk_range(stream, 0, n)
    assert hashlib.sha256(full).hexdigest() == rec["sha256"]
    for s, e in [(0, 1), (n - 1, n), (1, n), (0, n - 1), (n // 2, n // 2), (n // 3, 2 * n // 3 + 1)]:
        if 0 <= s <= e <= n:
            assert reader.extractChunkRange(s, e) == oracle.extract_chunk_range(stream, s, e)
    res = z.verify_xorbs([stream], [xorb_hash], want_chunk_hashes=True, ctx=ctx)
    assert res.verdict[0] == 1 and res.roots[0].tobytes() == xorb_hash
    assert (res.chunk_hashes == gold_hashes).all() and res.bytes_verified == rec["n_bytes"]
    with pytest.raises(z.ZestGpuError) as ei:
        reader.extract_chunk_range(0, n + 1)
    assert ei.value.code == -9


@pytest.mark.parametrize("name", ["xorb_bf16", "xorb_text", "xorb_rand", "xorb_mixed_auto", "xorb_tiny"])
def test_seal_xorb_reproduces_xet_core_file(ctx, vectors, name):
    """Seed path: chunk stream -> GPU chunk hashes + Merkle root -> XETBLOB footer == the file xet-core stored."""
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
    stream, raw = read_golden_xorb(rec)
    blob, root = z.seal_xorb(stream, ctx)
    assert blob == raw and z.hash_to_hex(root) == rec["xorb_hash"]


@pytest.mark.parametrize("name", ["xorb_bf16", "xorb_mixed_auto", "xorb_tiny"])
def test_sealed_xorb_file_reads_back(ctx, oracle, vectors, name):
    """Writer and reader round-trip: the stored form (chunk stream + XETBLOB footer, what xet-core's local store and
    seal_xorb produce) goes straight back through index, verify and extract -- the walk ends where the footer begins."""
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
    stream, raw = read_golden_xorb(rec)
    assert len(raw) > len(stream)
    ix_s, ix_f = z.xorb_index(stream), z.xorb_index(raw)
    assert ix_f.n_chunks == ix_s.n_chunks and ix_f.total_ulen == ix_s.total_ulen and np.array_equal(ix_f.hdr_off, ix_s.hdr_off)
    res = z.verify_xorbs([raw], [z.hex_to_hash(rec["xorb_hash"])], ctx=ctx)
    assert res.verdict[0] == 1 and res.bytes_verified == rec["n_bytes"]
    n = ix_s.n_chunks
    assert z.XorbReader.init(raw, ctx).extractChunkRange(0, n) == oracle.extract_chunk_range(stream, 0, n)
    # a footer that does not fit its own length word is not a footer: the bytes are walked as chunks and rejected
    bad = bytearray(raw); bad[-2] ^= 0x40
    with pytest.raises(z.ZestGpuError):
        z.xorb_index(bytes(bad))


def test_stats_report_gpu_time_per_stage(ctx, oracle):
    """zg_get_stats: GPU time of the verify stages (copies in, decode, hash, fold, results out) from CUDA events."""
    data = oracle.synth_zipf_text(3_000_000, 77)
    xs, _, _ = oracle.make_xorbs(data, 1)
    s0 = ctx.stats()
    r = z.verify_xorbs([x.blob for x in xs], [x.expected for x in xs], ctx=ctx)
    s1 = ctx.stats()
    assert r.ok
    for k in ("ns_h2d", "ns_decode", "ns_hash", "ns_merkle", "ns_d2h"):
        assert s1[k] > s0[k], k
    assert s1["ns_decode"] - s0["ns_decode"] < 1_000_000_000


GENS = {"text": "synth_zipf_text", "bf16": "synth_bf16", "rand": "synth_random"}


@pytest.mark.parametrize("kind,policy,nbytes,block_max", [
    ("text", 1, 3_000_001, 0), ("text", 1, 1_500_000, 65536), ("text", 2, 900_000, 0), ("bf16", 2, 6_000_000, 0),
    ("bf16", 1, 1_000_003, 0), ("rand", 1, 1_000_000, 0), ("rand", 0, 777_777, 0), ("bf16", 2, 2_000_000, 4096),
])
def test_synthetic_roundtrip(ctx, oracle, kind, policy, nbytes, block_max):
    data = getattr(oracle, GENS[kind])(nbytes, 31)
    xs, ends, hashes = oracle.make_xorbs(data, policy, threads=4, block_max=block_max)
    x = xs[0]
    out = z.XorbReader(x.blob, ctx).extract_chunk_range(0, x.n_chunks)
    assert out == data.tobytes()
    res = z.verify_xorbs([x.blob], [x.expected], want_chunk_hashes=True, ctx=ctx)
    assert res.verdict[0] == 1 and (res.chunk_hashes == hashes).all() and res.bytes_verified == nbytes


@pytest.mark.parametrize("fflags", [1, 2, 4, 8, 9, 15])
@pytest.mark.parametrize("block_max", [0, 65536, 4096])
def test_frame_variants(ctx, oracle, fflags, block_max):
    """Parts of the LZ4 frame format xet-core does not emit but a decoder must accept: content-size
    field, block and content checksums (skipped), linked blocks whose matches reach into earlier blocks, many small
    blocks, stored blocks."""
    for kind, policy in (("text", 1), ("bf16", 2), ("rand", 1)):
        data = getattr(oracle, GENS[kind])(500_001, fflags)
        xs, _, hashes = oracle.make_xorbs(data, policy, block_max=block_max, fflags=fflags)
        x = xs[0]
        assert z.XorbReader(x.blob, ctx).extract_chunk_range(0, x.n_chunks) == data.tobytes()
        res = z.verify_xorbs([x.blob], [x.expected], want_chunk_hashes=True, ctx=ctx)
        assert res.ok and (res.chunk_hashes == hashes).all()
        # a wrong content-size field / a cut frame must give the oracle's verdict
        if fflags & 1:
            idx = oracle.xorb_index(x.blob)
            bad = bytearray(x.blob)
            bad[int(idx.hdr_off[1]) + 8 + 6] ^= 1          # low byte of chunk 1's content size
            bad = bytes(bad)
            got = z.verify_xorbs([bad], [x.expected], ctx=ctx)
            want, _, _ = oracle.verify_xorbs([bad], [x.expected])
            assert got.verdict[0] == want[0] and want[0] < 0


def test_crafted_streams(ctx, oracle):
    """Hand-rolled LZ4 blocks that no real compressor emits but the format allows: multi-byte length extensions,
    matches beyond 64 KiB (the record splitter), offsets of 1, 2, 3 and 65535, empty literal runs, thousands of tiny
    sequences -- as scheme 1 and (through the byte-grouping inverse) scheme 2; then random corruptions of them must
    get the oracle's verdict."""
    from conftest import chunk_record, craft_lz4_block, frame_single_block
    rng = np.random.default_rng(78)
    records, wants = [], []
    for style, n in [("tiny", 5000), ("mixed", 60000), ("long", 900000), ("long", 300001), ("tiny", 9), ("mixed", 1),
                     ("mixed", 131072), ("tiny", 70003), ("long", 2_000_000), ("edge", 400_000), ("edge", 65_537)]:
        blk, want = craft_lz4_block(rng, n, style)
        frame = frame_single_block(blk, len(want))
        records.append(chunk_record(frame, 1, len(want))); wants.append(want)
        # scheme 2: the decoded block is the grouped form of some data; regroup it to get the expected bytes
        g = np.frombuffer(want, np.uint8)
        q, r = divmod(g.size, 4)
        sizes = [q + (k < r) for k in range(4)]
        planes = np.split(g, np.cumsum(sizes)[:-1])
        inter = np.empty(g.size, np.uint8)
        for k in range(4):
            inter[k::4] = planes[k]
        records.append(chunk_record(frame, 2, len(want))); wants.append(inter.tobytes())
    stream = b"".join(records)
    n = len(records)
    assert oracle.extract_chunk_range(stream, 0, n) == b"".join(wants)
    assert z.XorbReader(stream, ctx).extract_chunk_range(0, n) == b"".join(wants)
    for i in (0, 3, 4, 9, n - 1):
        assert z.XorbReader(stream, ctx).extract_chunk_range(i, i + 1) == wants[i]
    # corruptions: same verdicts as the oracle, xorb by xorb
    blobs = []
    for _ in range(60):
        b = bytearray(records[int(rng.integers(0, n))])
        for _ in range(int(rng.integers(1, 4))):
            b[int(rng.integers(8, len(b)))] ^= 1 << int(rng.integers(0, 8))
        blobs.append(bytes(b))
    got = z.verify_xorbs(blobs, ctx=ctx)
    want_v, want_roots, _ = oracle.verify_xorbs(blobs, threads=8)
    assert list(got.verdict) == list(want_v)
    ok = want_v >= 0
    assert (got.roots[ok] == want_roots[ok]).all()


def test_special_streams(ctx, oracle):
    """Long matches (zeros, period-3, period-40), a one-byte file, an empty xorb list, an empty blob."""
    cases = [np.zeros(300_000, np.uint8), np.tile(np.frombuffer(b"abc", np.uint8), 100_000),
             np.tile(np.arange(40, dtype=np.uint8), 9000), np.frombuffer(b"x", np.uint8),
             np.concatenate([np.zeros(70_000, np.uint8), oracle.synth_random(5000, 1), np.zeros(70_000, np.uint8)])]
    for data in cases:
        for policy in (1, 2):
            xs, _, hashes = oracle.make_xorbs(data, policy)
            out = z.XorbReader(xs[0].blob, ctx).extract_chunk_range(0, xs[0].n_chunks)
            assert out == data.tobytes()
            assert z.verify_xorbs([xs[0].blob], [xs[0].expected], ctx=ctx).ok
    assert z.verify_xorbs([], ctx=ctx).verdict.size == 0
    r = z.verify_xorbs([b""], [bytes(32)], ctx=ctx)          # zero chunks: root is 32 zero bytes (A.3)
    assert r.verdict[0] == 1 and r.roots[0].tobytes() == bytes(32)


def test_corruption_verdicts_match_oracle(ctx, oracle, vectors):
    """Bit flips in header, frame and payload; truncation; wrong expected hash: same verdict as the oracle."""
    blobs, expected = [], []
    for name in ("xorb_text", "xorb_bf16", "xorb_mixed_auto"):
        rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
        stream, raw = read_golden_xorb(rec)
        _, _, xh = footer_chunk_hashes(raw)
        rng = np.random.default_rng(len(stream))
        blobs.append(stream); expected.append(xh)
        blobs.append(stream); expected.append(bytes(32))                      # wrong address
        for pos in [0, 4, 8, 12, 13, 14, 15, 19, 40] + [int(p) for p in rng.integers(0, len(stream), 40)]:
            b = bytearray(stream)
            b[pos] ^= 1 << int(rng.integers(0, 8))
            blobs.append(bytes(b)); expected.append(xh)
        blobs.append(stream[:-7]); expected.append(xh)
        blobs.append(stream[:len(stream) // 2]); expected.append(xh)
    got = z.verify_xorbs(blobs, expected, ctx=ctx)
    want_v, want_roots, _ = oracle.verify_xorbs(blobs, expected, threads=8)
    assert list(got.verdict) == list(want_v)
    ok = want_v >= 0
    assert (got.roots[ok] == want_roots[ok]).all()
    assert (got.verdict[0], got.verdict[1]) == (1, 0)
    assert (got.verdict <= 0).sum() >= len(blobs) - 3 - 6   # nearly every flip is caught (a flip in padding may not be)


def test_decode_chunk_size_sweep(ctx, oracle):
    """Chunks of every size from 1 to 700 bytes and a run of sizes around the decoder's round geometry (its
    speculative walk cuts a block into 68-byte segments, 32 per round): each must decode bit-exact under both
    compression schemes, whatever the block length is modulo the segment and the span."""
    text = oracle.synth_zipf_text(3_000_000, 11)
    bf = oracle.synth_bf16(3_000_000, 12)
    sizes = list(range(1, 701)) + list(range(2100, 2400, 7)) + list(range(4200, 4500, 13)) + [8191, 8192, 8193, 40_000]
    ends = np.cumsum(np.array(sizes, np.uint64))
    total = int(ends[-1])
    for data, policy in ((text, 1), (text, 2), (bf, 2), (bf, 1)):
        d = data[:total]
        blob = oracle.build_xorb(d, ends, policy)
        n = len(sizes)
        assert oracle.extract_chunk_range(blob, 0, n) == d.tobytes()
        assert z.XorbReader(blob, ctx).extract_chunk_range(0, n) == d.tobytes()
        off = np.concatenate([np.zeros(1, np.uint64), ends])
        root = oracle.merkle_root(oracle.hash_chunks(d, off), np.diff(off))
        assert z.verify_xorbs([blob], [root], ctx=ctx).ok
    # every size as a byte-grouped chunk (a frame flag forces the framed form even where it is not smaller): verify
    # hashes such chunks from their four planes without regrouping them first -- plane lengths n/4 + (k < n%4), planes
    # starting at any alignment, blocks and leaves that end inside a plane
    d = bf[:total]
    blob = oracle.build_xorb(d, ends, 2, fflags=1)
    assert set(oracle.xorb_index(blob).scheme.tolist()) == {2}
    gold = oracle.hash_chunks(d, off)
    res = z.verify_xorbs([blob], [oracle.merkle_root(gold, np.diff(off))], want_chunk_hashes=True, ctx=ctx)
    assert res.ok and np.array_equal(res.chunk_hashes, gold)
    assert z.XorbReader(blob, ctx).extract_chunk_range(0, len(sizes)) == d.tobytes()


def test_decode_fuzz_payload(ctx, oracle):
    """Random damage inside the LZ4 payload of real-sized chunks (text / LZ4 and bf16 / BG4-LZ4), one damaged chunk
    per xorb: the verdict must be the oracle's, and where the damaged stream still decodes (a flipped literal, a
    changed length that happens to stay consistent) the root must be the oracle's too -- i.e. the decoded bytes are
    the same even for streams no compressor would emit."""
    rng = np.random.default_rng(4242)
    blobs = []
    for data, policy, block_max, fflags in ((oracle.synth_zipf_text(400_000, 21), 1, 0, 0), (oracle.synth_bf16(400_000, 22), 2, 0, 0),
                                            (oracle.synth_zipf_text(300_000, 23), 1, 4096, 15),      # many small linked blocks, all checksums
                                            (oracle.synth_bf16(300_000, 24), 2, 65536, 9)):          # 64 KiB blocks, content size
        xs, _, _ = oracle.make_xorbs(data, policy, block_max=block_max, fflags=fflags)
        blob = xs[0].blob
        ix = oracle.xorb_index(blob)
        for c in range(min(4, ix.n)):
            lo = int(ix.hdr_off[c]); clen = int(ix.clen[c])
            rec = blob[lo:lo + 8 + clen]
            for _ in range(300):
                b = bytearray(rec)
                k = int(rng.integers(1, 5))
                for _ in range(k):
                    pos = int(rng.integers(8 + 11, len(b) - 4))      # past the frame header and first block size
                    mode = int(rng.integers(0, 4))
                    if mode == 0: b[pos] ^= 1 << int(rng.integers(0, 8))
                    elif mode == 1: b[pos] = int(rng.integers(0, 256))
                    elif mode == 2: b[pos] = 0xff
                    else: b[pos] = 0
                blobs.append(bytes(b))
    got = z.verify_xorbs(blobs, ctx=ctx)
    want_v, want_roots, _ = oracle.verify_xorbs(blobs, threads=8)
    assert list(got.verdict) == list(want_v)
    ok = want_v >= 0
    assert ok.sum() > 200 and (~ok).sum() > 1000         # both outcomes are well represented
    assert (got.roots[ok] == want_roots[ok]).all()


def test_decode_garbage_blocks(ctx, oracle):
    """LZ4 blocks made of noise behind a valid frame header: thousands of byte strings that mean nothing go through the
    hop tables, the chain walk and the execute step until some check stops them.  Nothing may hang or fault, and every
    verdict -- which error, or that the noise happened to decode -- must be the oracle's.  Three kinds of noise: uniform
    bytes, bytes biased to small values (short literal runs, long chains of sequences), and a valid stream's prefix with
    noise after it."""
    from conftest import chunk_record, craft_lz4_block, frame_single_block
    rng = np.random.default_rng(9001)
    good_blk, good_out = craft_lz4_block(rng, 60000, "mixed")
    blobs = []
    for i in range(4000):
        kind = i % 3
        n = int(rng.integers(32, 5000))
        if kind == 0:
            blk = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        elif kind == 1:
            blk = (rng.integers(0, 256, n, dtype=np.uint8) & rng.choice([0x0f, 0x1f, 0x33, 0x77], n).astype(np.uint8)).tobytes()
        else:
            cut = int(rng.integers(16, len(good_blk) - 1))
            blk = good_blk[:cut] + rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        ulen = int(rng.choice([1, 100, 8192, 65536, 131072, int(rng.integers(1, 200000))]))
        blobs.append(chunk_record(frame_single_block(blk, ulen), 1 + (i // 3) % 2, ulen))
    got = z.verify_xorbs(blobs, ctx=ctx)
    want_v, want_roots, _ = oracle.verify_xorbs(blobs, threads=8)
    assert list(got.verdict) == list(want_v)
    ok = want_v >= 0
    assert (got.roots[ok] == want_roots[ok]).all()
    assert len(set(int(v) for v in want_v)) >= 2          # more than one kind of outcome


def test_many_xorbs_batch(ctx, oracle):
    """A batch that spans several sub-batches and both pipeline lanes, mixed schemes, one corrupt in the middle."""
    blobs, expected = [], []
    for i in range(14):
        kind, policy = [("text", 1), ("bf16", 2), ("rand", 0)][i % 3]
        data = getattr(oracle, GENS[kind])(60_000_000 if i == 5 else 2_000_000 + 4099 * i, 100 + i)
        for x in oracle.make_xorbs(data, policy, threads=8)[0]:
            blobs.append(x.blob); expected.append(x.expected)
    bad = bytearray(blobs[7]); bad[len(bad) // 3] ^= 0x40; blobs[7] = bytes(bad)
    got = z.verify_xorbs(blobs, expected, ctx=ctx)
    want_v, want_roots, want_bytes = oracle.verify_xorbs(blobs, expected, threads=8)
    assert list(got.verdict) == list(want_v) and got.verdict[7] <= 0
    assert (got.roots[want_v >= 0] == want_roots[want_v >= 0]).all()


def test_concurrent_callers(ctx, oracle, vectors):
    """zest fans out up to 128 concurrent tasks per batch (src/parallel_download.zig:117,174-182), each calling
    extractChunkRange on its own term: many host threads on one context must all get oracle-exact bytes."""
    import threading
    items = []
    for name in ("xorb_text", "xorb_bf16", "xorb_mixed_auto", "xorb_rand"):
        rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
        stream, raw = read_golden_xorb(rec)
        _, _, xh = footer_chunk_hashes(raw)
        n = rec["n_chunks"]
        for (s, e) in [(0, n), (1, n), (0, n - 1), (n // 2, n)]:
            items.append((stream, s, e, oracle.extract_chunk_range(stream, s, e), xh))
    errors = []

    def worker(k):
        try:
            for rep in range(6):
                stream, s, e, want, xh = items[(k + rep) % len(items)]
                assert z.XorbReader(stream, ctx).extract_chunk_range(s, e) == want
                if rep % 3 == 0:
                    assert z.verify_xorbs([stream], [xh], ctx=ctx).ok
        except Exception as ex:  # noqa: BLE001
            errors.append(repr(ex))

    ts = [threading.Thread(target=worker, args=(k,)) for k in range(32)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors[:3]


def test_range_hashes(ctx, oracle, vectors):
    """zg_range_hashes against the range hashes in hf_xet's shards (golden) and, for random term partitions of a
    long chunk-hash list (empty terms, single-chunk terms, one term of everything), against the oracle."""
    for rec in vectors["terms"]:
        data = golden_input(rec["name"])
        ends, hashes, fh = z.cdc_hash(data, ctx=ctx)
        assert z.hash_to_hex(fh) == rec["file_hash"]
        bounds = np.concatenate([[0], np.cumsum([t["chunk_end"] - t["chunk_start"] for t in rec["terms"]])]).astype(np.uint32)
        assert int(bounds[-1]) == ends.size
        got = z.range_hashes(hashes, bounds, ctx=ctx)
        assert [g.tobytes().hex() for g in got] == [t["range_hash"] for t in rec["terms"]]
    rng = np.random.default_rng(9)
    hashes = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    for k in (1, 2, 7, 400, 4999):
        cuts = np.sort(rng.integers(0, 5001, k - 1)) if k > 1 else np.zeros(0, np.int64)
        bounds = np.concatenate([[0], cuts, [5000]]).astype(np.uint32)
        assert (z.range_hashes(hashes, bounds, ctx=ctx) == oracle.range_hashes(hashes, bounds)).all()
    assert z.range_hashes(hashes, [3, 3], ctx=ctx)[0].tobytes() == oracle.range_hashes(hashes, [3, 3])[0].tobytes()
    assert z.range_hashes(hashes, [], ctx=ctx).shape == (0, 32)
    with pytest.raises(z.ZestGpuError):
        z.range_hashes(hashes, [5, 2], ctx=ctx)


def _dict_insert(table, hashes, locs):
    """Reference semantics of DedupIndex.insert on a Python dict."""
    keys = [h.tobytes() for h in hashes]
    fresh = {k for k in keys if k not in table}
    for k, l in zip(keys, locs):
        table[k] = min(table.get(k, 1 << 64), int(l))
    return np.array([k in fresh and table[k] == int(l) for k, l in zip(keys, locs)])


def test_dedup_index_vs_dict(ctx):
    """Insert / lookup against a dict: duplicates inside a batch and across batches, absent keys, keys that share
    their slot word or their tag word with a stored key, locations that are not monotonic."""
    rng = np.random.default_rng(31)
    idx = z.DedupIndex(200_000, ctx)
    table = {}
    pool = rng.integers(0, 256, (60_000, 32), dtype=np.uint8)
    pool[1000:1100, :8] = pool[0, :8]            # same home slot as key 0: a probe chain
    pool[2000:2050, 8:16] = pool[1, 8:16]        # same tag as key 1, different slot word
    for b in range(6):
        pick = rng.integers(0, pool.shape[0], 30_000)
        hashes = pool[pick]
        locs = rng.integers(0, 1 << 40, pick.size).astype(np.uint64)
        got = idx.insert(hashes, locs)
        want = _dict_insert(table, hashes, locs)
        assert (got == want).all(), b
        assert len(idx) == len(table)
    probe = np.concatenate([pool[::7], rng.integers(0, 256, (5000, 32), dtype=np.uint8)])
    got = idx.lookup(probe)
    want = np.array([table.get(h.tobytes(), int(z.dedup.MISSING)) for h in probe], dtype=np.uint64)
    assert (got == want).all()
    assert idx.lookup(np.zeros((0, 32), np.uint8)).size == 0
    with pytest.raises(z.ZestGpuError):          # the reserved value is refused
        idx.insert(pool[:1], np.array([z.dedup.MISSING], np.uint64))
    idx.close()


def test_dedup_index_capacity_and_real_hashes(ctx, oracle):
    """Over-filling reports ZG_ERR_CAPACITY; chunk hashes of a file with repeated regions dedup as expected."""
    small = z.DedupIndex(100, ctx)
    h = np.random.default_rng(5).integers(0, 256, (400, 32), dtype=np.uint8)
    with pytest.raises(z.ZestGpuError) as ei:
        small.insert(h, np.arange(400, dtype=np.uint64))
    assert ei.value.code == -10
    small.close()
    base = oracle.synth_random(3_000_000, 8)
    data = np.concatenate([base, oracle.synth_random(500_000, 9), base])      # the second copy re-synchronises
    ends, hashes, _ = z.cdc_hash(data, ctx=ctx)
    idx = z.DedupIndex(4 * ends.size, ctx)
    new = idx.insert(hashes, np.arange(ends.size, dtype=np.uint64))
    uniq = len({x.tobytes() for x in hashes})
    assert int(new.sum()) == uniq == len(idx) and uniq < ends.size - 20
    loc = idx.lookup(hashes)
    first = {}
    for i, x in enumerate(hashes):
        first.setdefault(x.tobytes(), i)
    assert [int(v) for v in loc] == [first[x.tobytes()] for x in hashes]
    idx.close()


def test_dedup_device_calls(ctx):
    """zg_dev_dedup_insert / lookup on device-resident hashes (as they come out of zg_dev_hash_chunks) give what
    the host-buffer calls give."""
    import ctypes as C
    import torch
    from zest_b200 import _native
    L = _native.lib()
    rng = np.random.default_rng(77)
    hashes = rng.integers(0, 256, (50_000, 32), dtype=np.uint8)
    hashes[25_000:] = hashes[:25_000]                       # every hash twice
    locs = np.arange(50_000, dtype=np.uint64)[::-1].copy()  # descending: the later occurrence has the smaller location
    dev = torch.device("cuda", ctx.device)
    d_h = torch.from_numpy(hashes).to(dev)
    d_l = torch.from_numpy(locs.view(np.int64)).to(dev)
    d_new = torch.zeros(50_000, dtype=torch.uint8, device=dev)
    d_out = torch.zeros(50_000, dtype=torch.int64, device=dev)
    idx = z.DedupIndex(60_000, ctx)
    sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _native.check(L.zg_dev_dedup_insert(idx.handle, d_h.data_ptr(), d_l.data_ptr(), 50_000, d_new.data_ptr(), sp))
    _native.check(L.zg_dev_dedup_lookup(idx.handle, d_h.data_ptr(), 50_000, d_out.data_ptr(), sp))
    torch.cuda.synchronize(dev)
    assert len(idx) == 25_000
    new = d_new.cpu().numpy().astype(bool)
    assert not new[:25_000].any() and new[25_000:].all()   # the occurrence with the smaller location is the one kept
    got = d_out.cpu().numpy().view(np.uint64)
    assert (got[:25_000] == locs[25_000:]).all() and (got[25_000:] == locs[25_000:]).all()
    assert (idx.lookup(hashes) == got).all()
    idx.close()


# ------------------------------------------------------------------------------------------------ CDC
@pytest.mark.parametrize("name", ["seed0_1MiB", "seed1_8MiB", "zeros_1MiB", "seed1_next_8191", "seed1_next_8192",
                                  "bf16_seed2_4MiB", "zipf_seed3_2MiB", "hello_world", "empty"])
def test_cdc_kats(ctx, oracle, vectors, name):
    rec = next(r for r in vectors["xet"] if r["name"] == name)
    data = golden_input(name)
    ends, hashes, fh = z.cdc_hash(data, ctx=ctx)
    assert list(ends) == list(oracle.cdc(data))
    assert z.hash_to_hex(fh) == rec["file_hash"]
    assert list(z.cdc(data, ctx=ctx)) == list(ends)
    if rec["xorb_hash"] is not None and rec["n_chunks"] == ends.size:
        assert list(ends) == rec["chunk_ends"]
        assert [z.hash_to_hex(h.tobytes()) for h in hashes] == rec["chunk_hashes"]
        lens = np.diff(np.concatenate([[0], ends])).astype(np.uint64)
        assert z.hash_to_hex(z.merkle_root(hashes, lens, ctx=ctx)) == rec["xorb_hash"]


@pytest.mark.parametrize("n", [1, 63, 64, 8191, 8192, 8193, 32767, 32768, 32769, 131072, 131073, 262144 + 8191, 1_000_003])
def test_cdc_sizes_vs_oracle(ctx, oracle, n):
    for gen in (oracle.synth_random, oracle.synth_zipf_text):
        data = gen(n, n)
        assert list(z.cdc(data, ctx=ctx)) == list(oracle.cdc(data))


def test_cdc_dense_candidates(ctx, oracle):
    """Adversarial input: a period-4 pattern whose windowed hash has its top 16 bits clear at one phase, so every
    fourth byte is a candidate (8,192 per 32 KiB tile) and tiles overflow their hit lists -- the 'dense, re-scan'
    path of the walker.  Cuts must still be exact."""
    import os
    import re
    from conftest import ROOT
    T = np.array([int(x, 16) for x in re.findall(r"0x([0-9a-f]{16})ULL", open(os.path.join(ROOT, "include", "zest_gear_table.h")).read())],
                 dtype=np.uint64)
    assert T.size == 256
    w = [np.uint64(sum(1 << k for k in range(j, 64, 4))) for j in range(4)]   # weight of the byte j steps back
    found = None
    with np.errstate(over="ignore"):
        for a in range(256):
            for b in range(0, 256, 5):
                H = T[a] * w[0] + T[b] * w[1] + (T[:, None] * w[2] + T[None, :] * w[3])
                hit = np.argwhere((H >> np.uint64(48)) == 0)
                if hit.size:
                    found = (a, b, int(hit[0][0]), int(hit[0][1]))
                    break
            if found:
                break
    assert found, "no period-4 pattern found"
    a, b, c, d = found                                    # newest byte a, then b, c, d going back
    pat = np.array([d, c, b, a], np.uint8)
    data = np.concatenate([oracle.synth_random(50_001, 1), np.tile(pat, 150_000), oracle.synth_random(70_000, 2),
                           np.tile(pat, 20_000), oracle.synth_random(9_000, 3)])
    want = oracle.cdc(data)
    lens = np.diff(np.concatenate([[0], want]))
    assert (lens[2:40] < 8200).sum() > 20                 # the pattern really does cut at the minimum size
    assert list(z.cdc(data, ctx=ctx)) == list(want)


# ------------------------------------------------------------------------------------------------ full-size properties
def test_config1_full_size(ctx, vectors):
    """BASELINE.json configs[0] at full size: 1 GiB random -> 16,741 chunks, hf_xet's file hash (golden)."""
    rec = next(r for r in vectors["xet"] if r["name"] == "seed0_1GiB")
    data = golden_input("seed0_1GiB")
    ends, hashes, fh = z.cdc_hash(data, ctx=ctx)
    assert ends.size == rec["n_chunks"] and int(ends[-1]) == data.size
    assert z.hash_to_hex(fh) == rec["file_hash"]
    # linearity of the batch API: hashing two halves of the chunk list equals hashing the whole
    off = np.concatenate([[0], ends]).astype(np.uint64)
    k = ends.size // 2
    both = np.concatenate([z.hash_chunks(data, off[:k + 1], ctx=ctx), z.hash_chunks(data, off[k:], ctx=ctx)])
    assert (both == hashes).all()


def test_config2_full_size_roundtrip(ctx, oracle):
    """BASELINE.json configs[1] at full size (~550 MB Zipf text, LZ4): build -> verify all ok -> reconstruct
    bit-exact -> flip one byte -> exactly that xorb fails."""
    data = oracle.synth_zipf_text(550_000_000, 2)
    xs, ends, hashes = oracle.make_xorbs(data, 1, threads=8)
    blobs = [x.blob for x in xs]
    res = z.verify_xorbs(blobs, [x.expected for x in xs], want_chunk_hashes=True, ctx=ctx)
    assert res.ok and res.bytes_verified == data.size and (res.chunk_hashes == hashes).all()
    pos = 0
    for x in xs[:2] + xs[-1:]:
        pass
    first = z.XorbReader(blobs[0], ctx).extract_chunk_range(0, xs[0].n_chunks)
    assert first == data[:xs[0].ulen_total].tobytes()
    bad = bytearray(blobs[3]); bad[len(bad) // 2] ^= 2; blobs[3] = bytes(bad)
    res2 = z.verify_xorbs(blobs, [x.expected for x in xs], ctx=ctx)
    assert [int(v == 1) for v in res2.verdict] == [int(i != 3) for i in range(len(xs))]


def test_bench_data_file_hash_on_gpu(ctx, oracle, vectors):
    """The bytes of the bench workloads through the GPU seed path (CDC + chunk hashes + file hash in one call) against
    hf_xet's file hash of the same bytes (tests/golden/vectors.json "bench"; the 64 MiB record is xorb 0 of the default set)."""
    for rec in vectors["bench"]:
        data = oracle.synth_bf16_fast(rec["n_bytes"], rec["seed"])
        ends, hashes, fh = z.cdc_hash(data, ctx=ctx)
        assert z.hash_to_hex(fh) == rec["file_hash"]
        assert np.array_equal(ends, oracle.cdc(data))


def test_config3_bg4_at_scale(ctx, oracle):
    """BASELINE.json configs[2] kind of input at a size that means something for scheme 2 (2 GiB of bf16, BG4-LZ4, 32
    xorbs, ~33 k chunks; bench.py repeats this check on the full 16 GB): every verdict, every root and every chunk hash
    against the oracle's, through the host-buffer call; one whole xorb reconstructed bit-exact; one damaged byte caught."""
    from concurrent.futures import ThreadPoolExecutor
    n_x, raw = 32, 64 << 20
    with ThreadPoolExecutor(16) as ex:
        built = list(ex.map(lambda i: oracle.synth_xorb("bf16", raw, 9000 + i, 2), range(n_x)))
    blobs = [b[0] for b in built]
    expected = np.frombuffer(b"".join(b[1] for b in built), np.uint8).reshape(-1, 32)
    gold = np.concatenate([b[2] for b in built])
    assert all(int(np.frombuffer(bl[:8], np.uint8)[4]) == 2 for bl in blobs)  # scheme byte of the first chunk header
    res = z.verify_xorbs(blobs, expected, want_chunk_hashes=True, ctx=ctx)
    assert res.ok and res.bytes_verified == n_x * raw
    assert np.array_equal(res.roots, expected) and np.array_equal(res.chunk_hashes, gold)
    k = 17
    want = oracle.synth_bf16_fast(raw, 9000 + k)
    got = z.XorbReader(blobs[k], ctx).extract_chunk_range(0, built[k][3].size)
    assert got == want.tobytes()
    bad = blobs[5].copy(); bad[bad.size // 3] ^= 0x10
    res2 = z.verify_xorbs([bad], expected[5:6], ctx=ctx)
    assert res2.verdict[0] != 1


def test_config4_cdc_every_boundary_at_16GB(ctx, oracle):
    """BASELINE.json configs[3]: Gearhash CDC of one 16 GB bf16 shard -- the size changes the kernel's segment geometry
    (1,024 segments of 512 tiles instead of 512 x 64 at 1 GiB) -- with EVERY boundary compared with the oracle's serial
    chunker (~10 s on the host), then every chunk hash on a 2 GB slice."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    if torch.cuda.mem_get_info(0)[0] < 40 << 30:
        pytest.skip("needs 40 GB of free device memory")
    n = 16_000_000_000
    piece = 256 << 20
    data = np.empty(n, np.uint8)
    with ThreadPoolExecutor(16) as ex:
        list(ex.map(lambda k: oracle.synth_bf16_fast(min(piece, n - k * piece), 333 + k, out=data[k * piece:k * piece + min(piece, n - k * piece)]),
                    range((n + piece - 1) // piece)))
    ends = z.cdc(data, ctx=ctx)
    gold = oracle.cdc(data)
    assert ends.size == gold.size and np.array_equal(ends, gold)
    m = int(np.searchsorted(gold, 2_000_000_000))
    off = np.concatenate([[0], gold[:m]]).astype(np.uint64)
    assert np.array_equal(z.hash_chunks(data, off, ctx=ctx), oracle.hash_chunks(data, off, threads=16))


def test_real_xet_core_xorbs_made_on_this_box(ctx, oracle):
    """hf_xet (xet-core, the production implementation of the protocol) is in this image: let it chunk, compress
    (lz4_flex, byte-grouped) and store 160 MiB of bf16 on the spot, then read the stored xorb files back as they lie on
    disk (chunk stream + XETBLOB footer): content address = file name, chunk hashes = the footer's, reconstructed
    bytes = the file.  tools/real_xorb_replay.py does the same at 2 GB with timing."""
    hf_xet = pytest.importorskip("hf_xet")
    import shutil
    import tempfile
    n = 160 << 20
    data = oracle.synth_bf16_fast(n, 77077)
    d = tempfile.mkdtemp(prefix="xetreplay")
    os.environ["HF_XET_XORB_COMPRESSION_POLICY"] = "bg4-lz4"
    try:
        s = hf_xet.XetSession()
        with s.new_upload_commit(endpoint="local://" + d) as c:
            h = c.start_upload_bytes(data.tobytes())
        file_hash = h.result().xet_info.hash
        xdir = os.path.join(d, "xet", "xorbs", "xorbs")
        files = sorted(os.listdir(xdir), key=lambda f: os.path.getmtime(os.path.join(xdir, f)))
        raws = [(f.split(".", 1)[1], np.fromfile(os.path.join(xdir, f), np.uint8)) for f in files]
    finally:
        shutil.rmtree(d, ignore_errors=True)
    assert len(raws) >= 2
    expected = [z.hex_to_hash(name) for name, _ in raws]
    res = z.verify_xorbs([r for _, r in raws], expected, want_chunk_hashes=True, ctx=ctx)
    assert res.ok and res.bytes_verified == n
    gold = np.concatenate([footer_chunk_hashes(r.tobytes())[0] for _, r in raws])
    assert np.array_equal(res.chunk_hashes, gold)
    out = b"".join(z.XorbReader(r, ctx).extract_chunk_range(0, z.xorb_index(r).n_chunks) for _, r in raws)
    assert out == data.tobytes()
    # and the whole file's Xet hash through the GPU seed path equals what xet-core reported for the upload
    assert z.hash_to_hex(z.cdc_hash(data, ctx=ctx)[2]) == file_hash


def test_device_hash_never_writes_past_its_scratch(ctx, oracle):
    """zg_dev_hash_chunks sizes its scratch from the data_bytes the caller states, while the chunk lengths are read on the
    device.  Lengths that add up to four times the stated bytes must not push the kernels past the scratch: the chunks
    that fit hash correctly, the others come out as 32 zero bytes, and a guard region behind the scratch stays intact."""
    import ctypes as C
    import torch
    from zest_b200 import _native
    from zest_b200._native import check
    L = _native.lib()
    dev = torch.device("cuda", 0)
    n, each = 100, 80_000
    data = oracle.synth_random(n * each, 31)
    off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(each))
    d_data = torch.from_numpy(np.concatenate([data, np.zeros(16, np.uint8)])).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    stated = n * each // 4                                   # the caller understates the data size
    need = L.zg_dev_hash_scratch_bytes(stated, n)
    guard = 1 << 20
    scratch = torch.full((need + guard,), 0xAB, dtype=torch.uint8, device=dev)
    out = torch.full((32 * n,), 0xCD, dtype=torch.uint8, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(L.zg_dev_hash_chunks(ctx.handle, d_data.data_ptr(), stated, d_off.data_ptr(), None, n, None, out.data_ptr(),
                               scratch.data_ptr(), need, stream))
    torch.cuda.synchronize()
    assert bool((scratch[need:] == 0xAB).all()), "the kernels wrote behind the scratch"
    got = out.cpu().numpy().reshape(n, 32)
    want = oracle.hash_chunks(data, off, threads=4)
    good = [i for i in range(n) if np.array_equal(got[i], want[i])]
    zero = [i for i in range(n) if not got[i].any()]
    assert len(good) + len(zero) == n and good and zero
    assert good == list(range(len(good)))                    # the chunks that fit are the leading ones
    # with the true size stated, every chunk hashes
    need2 = L.zg_dev_hash_scratch_bytes(n * each, n)
    scratch2 = torch.empty(need2, dtype=torch.uint8, device=dev)
    check(L.zg_dev_hash_chunks(ctx.handle, d_data.data_ptr(), n * each, d_off.data_ptr(), None, n, None, out.data_ptr(),
                               scratch2.data_ptr(), need2, stream))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().reshape(n, 32), want)
