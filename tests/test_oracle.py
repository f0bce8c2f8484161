"""Pins the CPU oracle against the golden vectors (hf_xet / xet-core and PyPI blake3 outputs).

The reference (praveer13/zest) holds no vectors for this path; these are the
substitutes named in SURVEY.md 8c and A.7.  Everything here runs on CPU.
"""
import hashlib

import numpy as np
import pytest

from conftest import footer_chunk_hashes, golden_input, read_golden_xorb


def test_blake3_kats(oracle, vectors):
    for v in vectors["blake3"]:
        d = bytes([v["fill"]]) * v["len"] if "fill" in v else bytes(i % 251 for i in range(v["len"]))
        assert oracle.blake3(d).hex() == v["plain"], v["len"]
        if "data_key" in v:
            assert oracle.chunk_hash(d).hex() == v["data_key"], v["len"]
            node_key = bytes([1, 126, 197, 199, 165, 71, 41, 150, 253, 148, 102, 102, 180, 138, 2, 230, 93, 221,
                              83, 111, 55, 199, 109, 210, 248, 99, 82, 230, 74, 83, 113, 63])
            assert oracle.blake3_keyed(node_key, d).hex() == v["node_key"], v["len"]


def test_compression_count(oracle):
    # sum over leaves of max(1, ceil(len/64)) + (leaves - 1)
    assert oracle.blake3_compressions(0) == 1
    assert oracle.blake3_compressions(64) == 1
    assert oracle.blake3_compressions(1024) == 16
    assert oracle.blake3_compressions(1025) == 18
    assert oracle.blake3_compressions(131072) == 128 * 16 + 127


def test_hex_form(oracle):
    h = bytes(range(32))
    s = oracle.hash_to_hex(h)
    assert s[:16] == "0706050403020100"          # each 8-byte group little-endian (A.1)
    assert oracle.hex_to_hash(s) == h


@pytest.mark.parametrize("name", ["hello_world", "empty", "seed0_1MiB", "seed1_8MiB", "zeros_1MiB",
                                  "seed1_next_8191", "seed1_next_8192", "bf16_seed2_4MiB", "zipf_seed3_2MiB"])
def test_xet_kats(oracle, vectors, name):
    rec = next(r for r in vectors["xet"] if r["name"] == name)
    data = golden_input(name)
    assert hashlib.sha256(data.tobytes()).hexdigest() == rec["sha256"], "input recipe drifted"
    ends = oracle.cdc(data)
    off = np.concatenate([np.zeros(1, np.uint64), ends])
    hashes = oracle.hash_chunks(data, off)
    lens = np.diff(off)
    assert oracle.hash_to_hex(oracle.file_hash(hashes, lens)) == rec["file_hash"]
    if rec["xorb_hash"] is None:
        return
    gold_hashes = [oracle.hex_to_hash(h) for h in rec["chunk_hashes"]]
    if rec["n_chunks"] == ends.size:
        assert list(ends) == rec["chunk_ends"]
        assert [hashes[i].tobytes() for i in range(ends.size)] == gold_hashes
        assert oracle.hash_to_hex(oracle.merkle_root(hashes, lens)) == rec["xorb_hash"]
    else:
        # xet-core dedups identical chunks inside one upload (the all-zero file keeps 1 of its 8 chunks)
        ours = {hashes[i].tobytes() for i in range(ends.size)}
        assert set(gold_hashes) == ours
        gl = np.diff(np.concatenate([[0], rec["chunk_ends"]])).astype(np.uint64)
        gh = np.frombuffer(b"".join(gold_hashes), np.uint8).reshape(-1, 32)
        assert oracle.hash_to_hex(oracle.merkle_root(gh, gl)) == rec["xorb_hash"]


def test_config1_full_size(oracle, vectors):
    """1 GiB random, seed 0: 16,741 chunks and hf_xet's file hash (BASELINE.json configs[0])."""
    rec = next(r for r in vectors["xet"] if r["name"] == "seed0_1GiB")
    data = golden_input("seed0_1GiB")
    ends = oracle.cdc(data)
    assert ends.size == rec["n_chunks"]
    off = np.concatenate([np.zeros(1, np.uint64), ends])
    lens = np.diff(off)
    assert lens.max() == 131072 and lens[:-1].min() >= 8192
    hashes = oracle.hash_chunks(data, off, threads=8)
    assert oracle.hash_to_hex(oracle.file_hash(hashes, lens)) == rec["file_hash"]


@pytest.mark.parametrize("name", ["xorb_bf16", "xorb_text", "xorb_rand", "xorb_mixed_auto", "xorb_tiny"])
def test_real_xorbs(oracle, vectors, name):
    """Real xet-core xorbs: walk, decode every chunk, hash, fold; compare with the XETBLOB footer."""
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == name)
    stream, raw = read_golden_xorb(rec)
    gold_hashes, unc_ends, xorb_hash = footer_chunk_hashes(raw)
    idx = oracle.xorb_index(stream)
    assert idx.n == rec["n_chunks"]
    assert {str(s): int((idx.scheme == s).sum()) for s in np.unique(idx.scheme)} == rec["scheme_hist"]
    assert list(np.cumsum(idx.ulen.astype(np.uint64))) == list(unc_ends)
    full = oracle.extract_chunk_range(stream, 0, idx.n)
    off = np.concatenate([np.zeros(1, np.uint64), unc_ends])
    hashes = oracle.hash_chunks(np.frombuffer(full, np.uint8), off)
    assert (hashes == gold_hashes).all()
    assert oracle.merkle_root(hashes, idx.ulen) == xorb_hash
    assert oracle.hash_to_hex(xorb_hash) == rec["xorb_hash"]
    verdict, roots, nbytes = oracle.verify_xorbs([stream], expected=[xorb_hash], threads=2)
    assert verdict[0] == 1 and roots[0].tobytes() == xorb_hash and nbytes == idx.total_ulen
    # partial ranges concatenate to the whole
    if idx.n >= 3:
        parts = oracle.extract_chunk_range(stream, 0, 1) + oracle.extract_chunk_range(stream, 1, idx.n - 1) \
            + oracle.extract_chunk_range(stream, idx.n - 1, idx.n)
        assert parts == full


def test_builder_roundtrip(oracle):
    """The synthetic xorb builder (liblz4 block encoder + our framing) decodes back bit-exactly."""
    for policy, gen in [(1, oracle.synth_zipf_text), (2, oracle.synth_bf16), (0, oracle.synth_random), (1, oracle.synth_random)]:
        data = gen(700_001, 5)
        for block_max in (0, 65536):
            xs, ends, hashes = oracle.make_xorbs(data, policy, block_max=block_max)
            assert len(xs) == 1
            out = oracle.extract_chunk_range(xs[0].blob, 0, xs[0].n_chunks)
            assert out == data.tobytes()
            v, roots, nb = oracle.verify_xorbs([xs[0].blob], expected=[xs[0].expected])
            assert v[0] == 1 and nb == data.size


def test_frame_variants(oracle):
    """Optional LZ4 frame features (content size, checksums, linked and small blocks) round-trip through the oracle."""
    for kind, policy in ((oracle.synth_zipf_text, 1), (oracle.synth_bf16, 2)):
        data = kind(400_001, 9)
        for ff in (1, 2, 4, 8, 15):
            for bm in (0, 4096):
                xs, _, _ = oracle.make_xorbs(data, policy, block_max=bm, fflags=ff)
                assert oracle.extract_chunk_range(xs[0].blob, 0, xs[0].n_chunks) == data.tobytes()


def test_corruption_detected(oracle, vectors):
    rec = next(r for r in vectors["real_xorbs"] if r["name"] == "xorb_text")
    stream, raw = read_golden_xorb(rec)
    _, _, xorb_hash = footer_chunk_hashes(raw)
    bad = bytearray(stream)
    bad[len(bad) // 2] ^= 0x10                       # payload bit flip: decode error or hash mismatch
    v, _, _ = oracle.verify_xorbs([bytes(bad)], expected=[xorb_hash])
    assert v[0] <= 0
    bad = bytearray(stream); bad[0] = 1              # header version
    assert oracle.verify_xorbs([bytes(bad)], expected=[xorb_hash])[0][0] == -4
    bad = bytearray(stream); bad[4] = 3              # scheme 3 (zig-xet's FullBitsliceLZ4) is rejected
    assert oracle.verify_xorbs([bytes(bad)], expected=[xorb_hash])[0][0] == -5
    assert oracle.verify_xorbs([stream[:-5]], expected=[xorb_hash])[0][0] == -3
    bad = bytearray(stream); bad[8] ^= 1             # LZ4 frame magic
    assert oracle.verify_xorbs([bytes(bad)], expected=[xorb_hash])[0][0] == -6


def test_crafted_streams(oracle):
    """Hand-rolled LZ4 blocks (length extensions, > 64 KiB matches, offsets 1 / 65535) decode exactly."""
    from conftest import chunk_record, craft_lz4_block, frame_single_block
    rng = np.random.default_rng(77)
    for style, n in [("tiny", 5000), ("mixed", 60000), ("long", 900000), ("long", 300000), ("tiny", 9), ("mixed", 1)]:
        blk, want = craft_lz4_block(rng, n, style)
        stream = chunk_record(frame_single_block(blk, len(want)), 1, len(want))
        assert oracle.extract_chunk_range(stream, 0, 1) == want


def test_range_hashes_against_hf_xet_shards(oracle, vectors):
    """Term verification hashes: for every golden file the chunk hashes of each term, keyed
    with xet-core's VERIFICATION_KEY, give the range hash hf_xet wrote into its shard -- including the two-term
    70 MiB file, whose first xorb closes at 1,021 chunks."""
    from conftest import golden_input
    for rec in vectors["terms"]:
        data = golden_input(rec["name"])
        ends = oracle.cdc(data)
        off = np.concatenate([np.zeros(1, np.uint64), ends])
        hashes = oracle.hash_chunks(data, off, threads=8)
        lens = np.diff(off)
        assert oracle.hash_to_hex(oracle.file_hash(hashes, lens)) == rec["file_hash"]
        base, bounds = 0, [0]
        for t in rec["terms"]:
            assert t["chunk_start"] == 0                      # each term starts a fresh xorb in these files
            n = t["chunk_end"] - t["chunk_start"]
            assert int(lens[base:base + n].sum()) == t["unpacked"]
            assert oracle.hash_to_hex(oracle.merkle_root(hashes[base:base + n], lens[base:base + n])) == t["xorb_hash"]
            base += n
            bounds.append(base)
        assert base == ends.size
        got = oracle.range_hashes(hashes, bounds)
        assert [g.tobytes().hex() for g in got] == [t["range_hash"] for t in rec["terms"]]


def test_bench_data_file_hash(oracle, vectors):
    """The bytes the bench workloads are made of (the oracle's fast bf16 generator; the 64 MiB record is xorb 0 of the
    default 16 GB set): the generator reproduces them bit for bit, and the oracle's CDC + chunk hashes + Merkle fold +
    file hash over them equal what hf_xet.hash_files reported when tests/golden/make_golden.py --only-bench ran."""
    import hashlib
    for rec in vectors["bench"]:
        data = oracle.synth_bf16_fast(rec["n_bytes"], rec["seed"])
        assert hashlib.sha256(data.tobytes()).hexdigest() == rec["sha256"]
        ends = oracle.cdc(data)
        off = np.concatenate([np.zeros(1, np.uint64), ends])
        hashes = oracle.hash_chunks(data, off, threads=8)
        assert oracle.hash_to_hex(oracle.file_hash(hashes, np.diff(off))) == rec["file_hash"]
