#!/usr/bin/env python
"""Regenerate tests/golden/*: known-answer vectors and real xorbs from HuggingFace's own implementation.

Run in the BUILD container only (needs hf_xet 1.5.0 = xet-core, and PyPI blake3); the outputs are small
and committed, so neither package is needed by the tests or on the GPU box.

Why these and not the reference: praveer13/zest has no golden vector for this path
and its arithmetic lives in the un-vendored zig-xet package; xet-core is the protocol's reference
implementation that zig-xet must interoperate with.

Outputs
  vectors.json        BLAKE3 KATs (PyPI blake3), Xet KATs of SURVEY.md A.7 (hf_xet.hash_files + local:// upload)
  xorb_*.bin          real xet-core xorbs (chunk stream + XETBLOB footer) for scheme 0 / 1 / 2 / mixed
Usage:  python tests/golden/make_golden.py            (writes next to this file)
"""
import hashlib
import json
import os
import shutil
import struct
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DATA_KEY = bytes([102, 151, 245, 119, 91, 149, 80, 222, 49, 53, 203, 172, 165, 151, 24, 28, 157, 228, 33, 16,
                  155, 235, 43, 88, 180, 208, 176, 75, 147, 173, 242, 41])
NODE_KEY = bytes([1, 126, 197, 199, 165, 71, 41, 150, 253, 148, 102, 102, 180, 138, 2, 230, 93, 221, 83, 111,
                  55, 199, 109, 210, 248, 99, 82, 230, 74, 83, 113, 63])


def check_gear_table():
    import glob
    import hf_xet
    so = glob.glob(os.path.join(os.path.dirname(hf_xet.__file__), "*.so"))[0]
    b = open(so, "rb").read()
    i = b.find(bytes.fromhex("59f540e8a9d388b0"))
    t = b[i:i + 2048]
    assert hashlib.sha256(t).hexdigest() == "e1d3936666d7ae7a977c958e9afcc75f90aaca758ce5fbe4ece61dffefe1912c"
    words = struct.unpack("<256Q", t)
    hdr = open(os.path.join(HERE, "..", "..", "include", "zest_gear_table.h")).read()
    for w in words:
        assert "0x%016xULL" % w in hdr
    return hashlib.sha256(t).hexdigest()


def blake3_kats():
    import blake3
    lens = [0, 1, 2, 3, 63, 64, 65, 127, 128, 129, 1023, 1024, 1025, 2048, 2049, 3072, 3073, 4096, 5000, 5120,
            6144, 7168, 8192, 8193, 14336, 16384, 31744, 65536, 66560, 100000, 103424, 131071, 131072]
    out = []
    for n in lens:
        d = bytes(i % 251 for i in range(n))
        out.append({"len": n,
                    "plain": blake3.blake3(d).hexdigest(),
                    "data_key": blake3.blake3(d, key=DATA_KEY).hexdigest(),
                    "node_key": blake3.blake3(d, key=NODE_KEY).hexdigest()})
    # bench.zig-equivalent input (src/bench.zig:207-223): 64 KiB of 0x42, unkeyed
    out.append({"len": 65536, "fill": 0x42, "plain": blake3.blake3(b"\x42" * 65536).hexdigest()})
    return out


def parse_footer(raw: bytes):
    """XETBLOB footer -> dict(stream_len, xorb_hash, chunk_hashes, comp_ends, unc_ends)."""
    info_len = struct.unpack("<I", raw[-4:])[0]
    foot = raw[-4 - info_len:-4]
    assert foot[:7] == b"XETBLOB" and foot[7] == 1
    p = 8
    xorb_hash = foot[p:p + 32]; p += 32
    assert foot[p:p + 7] == b"XBLBHSH" and foot[p + 7] == 0; p += 8
    n = struct.unpack("<I", foot[p:p + 4])[0]; p += 4
    hashes = [foot[p + 32 * i:p + 32 * i + 32] for i in range(n)]; p += 32 * n
    assert foot[p:p + 7] == b"XBLBBND" and foot[p + 7] == 1; p += 8
    n2 = struct.unpack("<I", foot[p:p + 4])[0]; p += 4
    assert n2 == n
    comp = list(struct.unpack("<%dI" % n, foot[p:p + 4 * n])); p += 4 * n
    unc = list(struct.unpack("<%dI" % n, foot[p:p + 4 * n])); p += 4 * n
    return {"stream_len": len(raw) - 4 - info_len, "xorb_hash": xorb_hash, "chunk_hashes": hashes,
            "comp_ends": comp, "unc_ends": unc}


def to_hex(h: bytes) -> str:
    return "".join("%016x" % int.from_bytes(h[8 * g:8 * g + 8], "little") for g in range(4))


def upload_local(data: bytes, policy: str):
    """-> (file_hash_hex, [raw xorb files in creation order])."""
    import hf_xet
    d = tempfile.mkdtemp(prefix="xetgold")
    os.environ["HF_XET_XORB_COMPRESSION_POLICY"] = policy
    try:
        s = hf_xet.XetSession()
        with s.new_upload_commit(endpoint="local://" + d) as c:
            h = c.start_upload_bytes(data)
        fh = h.result().xet_info.hash
        xdir = os.path.join(d, "xet", "xorbs", "xorbs")
        files = sorted(os.listdir(xdir), key=lambda f: os.path.getmtime(os.path.join(xdir, f)))
        raws = [(f.split(".", 1)[1], open(os.path.join(xdir, f), "rb").read()) for f in files]
        return fh, raws
    finally:
        shutil.rmtree(d, ignore_errors=True)


def hash_file(data: bytes) -> str:
    import hf_xet
    with tempfile.NamedTemporaryFile(delete=False) as f:
        f.write(data)
        name = f.name
    try:
        r = hf_xet.hash_files([name])
        return r[0].hash
    finally:
        os.unlink(name)


def bf16_a7(n_elems: int, seed: int) -> bytes:
    x = (np.random.default_rng(seed).standard_normal(n_elems).astype(np.float32).view(np.uint32) >> 16)
    return x.astype(np.uint16).view(np.uint8).tobytes()


def zipf_text(nbytes: int, seed: int) -> bytes:
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from oracle.xet_oracle import synth_zipf_text
    return synth_zipf_text(nbytes, seed).tobytes()


def xet_kats():
    rng1 = np.random.default_rng(1)
    s1_8m = rng1.integers(0, 256, 8 << 20, dtype=np.uint8).tobytes()
    s1_8191 = rng1.integers(0, 256, 8191, dtype=np.uint8).tobytes()
    s1_8192 = rng1.integers(0, 256, 8192, dtype=np.uint8).tobytes()
    cases = [
        ("hello_world", b"hello world", "auto"),
        ("empty", b"", "auto"),
        ("seed0_1MiB", np.random.default_rng(0).integers(0, 256, 1 << 20, dtype=np.uint8).tobytes(), "auto"),
        ("seed1_8MiB", s1_8m, "auto"),
        ("zeros_1MiB", bytes(1 << 20), "auto"),
        ("seed1_next_8191", s1_8191, "auto"),
        ("seed1_next_8192", s1_8192, "auto"),
        ("bf16_seed2_4MiB", bf16_a7(2 << 20, 2), "bg4-lz4"),
        ("zipf_seed3_2MiB", zipf_text(2 << 20, 3), "lz4"),
    ]
    out = []
    for name, data, policy in cases:
        fh = hash_file(data)
        rec = {"name": name, "n_bytes": len(data), "sha256": hashlib.sha256(data).hexdigest(), "file_hash": fh,
               "policy": policy}
        if data:
            fh2, raws = upload_local(data, policy)
            assert fh2 == fh, (name, fh, fh2)
            assert len(raws) == 1
            xh, raw = raws[0]
            ft = parse_footer(raw)
            assert to_hex(ft["xorb_hash"]) == xh
            rec.update({"xorb_hash": xh, "n_chunks": len(ft["chunk_hashes"]), "chunk_ends": ft["unc_ends"],
                        "chunk_hashes": [to_hex(h) for h in ft["chunk_hashes"]]})
        else:
            rec.update({"xorb_hash": None, "n_chunks": 0, "chunk_ends": [], "chunk_hashes": []})
        print(name, rec["n_chunks"], rec["file_hash"], rec.get("xorb_hash"))
        out.append(rec)
    # config 1 at full size: file hash only (SURVEY.md A.7 last-but-one row)
    big = np.random.default_rng(0).integers(0, 256, 1 << 30, dtype=np.uint8).tobytes()
    out.append({"name": "seed0_1GiB", "n_bytes": len(big), "sha256": hashlib.sha256(big).hexdigest(),
                "file_hash": hash_file(big), "policy": None, "xorb_hash": None, "n_chunks": 16741,
                "chunk_ends": None, "chunk_hashes": None})
    print("seed0_1GiB", out[-1]["file_hash"])
    return out


def real_xorbs():
    """Small real xet-core xorbs, one per scheme plus a mixed one, stored with their XETBLOB footer."""
    rng = np.random.default_rng(7)
    mixed = (zipf_text(200_000, 11) + rng.integers(0, 256, 150_000, dtype=np.uint8).tobytes()
             + bf16_a7(100_000, 12) + bytes(70_000) + zipf_text(3000, 13))
    cases = [
        ("xorb_bf16", bf16_a7(192 * 1024, 21), "bg4-lz4"),
        ("xorb_text", zipf_text(384 * 1024, 22), "lz4"),
        ("xorb_rand", rng.integers(0, 256, 200_000, dtype=np.uint8).tobytes(), "auto"),
        ("xorb_mixed_auto", mixed, "auto"),
        ("xorb_tiny", b"xet" * 11, "lz4"),
    ]
    meta = []
    for name, data, policy in cases:
        fh, raws = upload_local(data, policy)
        assert len(raws) == 1
        xh, raw = raws[0]
        ft = parse_footer(raw)
        open(os.path.join(HERE, name + ".bin"), "wb").write(raw)
        schemes = []
        p = 0
        while p < ft["stream_len"]:
            cl = int.from_bytes(raw[p + 1:p + 4], "little")
            schemes.append(raw[p + 4])
            p += 8 + cl
        rec = {"name": name, "file": name + ".bin", "policy": policy, "n_bytes": len(data),
               "sha256": hashlib.sha256(data).hexdigest(), "file_hash": fh, "xorb_hash": xh,
               "stream_len": ft["stream_len"], "n_chunks": len(ft["chunk_hashes"]),
               "scheme_hist": {str(s): schemes.count(s) for s in sorted(set(schemes))}}
        print(name, len(raw), rec["scheme_hist"], xh)
        meta.append(rec)
    return meta


def shard_terms(data: bytes, policy: str):
    """Upload `data` to a local:// endpoint and read the file's reconstruction terms back from the .mdb shard
    xet-core writes: per term the xorb hash, the chunk index range inside that xorb, the unpacked byte count and
    the FileVerificationEntry range hash (raw byte order).  Shard layout (xet spec, "MDB shard file format"):
    48-byte header, then per file a 48-byte header row {file hash, flags u32, n_entries u32, ..}, n_entries rows
    {xorb hash, cas_flags u32, unpacked_bytes u32, chunk_start u32, chunk_end u32}, then -- when flags has bit 31 --
    n_entries verification rows {range hash, 16 unused}, then -- bit 30 -- one metadata row; a row of 0xFF ends
    the section."""
    import glob
    import hf_xet
    d = tempfile.mkdtemp(prefix="xetgold")
    os.environ["HF_XET_XORB_COMPRESSION_POLICY"] = policy
    try:
        s = hf_xet.XetSession()
        with s.new_upload_commit(endpoint="local://" + d) as c:
            h = c.start_upload_bytes(data)
        fh = h.result().xet_info.hash
        terms = []
        for p in sorted(glob.glob(os.path.join(d, "xet", "xorbs", "shards", "*.mdb"))):
            b = open(p, "rb").read()
            pos = 48
            while b[pos:pos + 32] != b"\xff" * 32:
                fhash = b[pos:pos + 32]
                flags, n = struct.unpack("<II", b[pos + 32:pos + 40])
                pos += 48
                rows = [b[pos + 48 * i:pos + 48 * i + 48] for i in range(n)]
                pos += 48 * n
                assert flags & 0x80000000, "no verification entries"
                ver = [b[pos + 48 * i:pos + 48 * i + 32] for i in range(n)]
                pos += 48 * n
                if flags & 0x40000000:
                    pos += 48
                if to_hex(fhash) != fh:
                    continue
                for r, v in zip(rows, ver):
                    _, unpacked, cs, ce = struct.unpack("<IIII", r[32:48])
                    terms.append({"xorb_hash": to_hex(r[:32]), "chunk_start": cs, "chunk_end": ce, "unpacked": unpacked,
                                  "range_hash": v.hex()})
        return fh, terms
    finally:
        shutil.rmtree(d, ignore_errors=True)


def term_kats():
    """Reconstruction terms with their verification ("range") hashes, SURVEY.md 8f item 3: one-term files of 1 to
    ~130 chunks and a 70 MiB file that spans two xorbs (two terms)."""
    rng1 = np.random.default_rng(1)
    cases = [
        ("hello_world", b"hello world", "auto"),
        ("seed0_1MiB", np.random.default_rng(0).integers(0, 256, 1 << 20, dtype=np.uint8).tobytes(), "auto"),
        ("seed1_8MiB", rng1.integers(0, 256, 8 << 20, dtype=np.uint8).tobytes(), "auto"),
        ("zipf_seed3_2MiB", zipf_text(2 << 20, 3), "lz4"),
        ("seed4_70MiB", np.random.default_rng(4).integers(0, 256, 70 << 20, dtype=np.uint8).tobytes(), "auto"),
    ]
    out = []
    for name, data, policy in cases:
        fh, terms = shard_terms(data, policy)
        assert terms and sum(t["unpacked"] for t in terms) == len(data), (name, terms)
        print(name, fh, [(t["chunk_start"], t["chunk_end"]) for t in terms])
        out.append({"name": name, "n_bytes": len(data), "file_hash": fh, "terms": terms})
    return out


def bench_kats():
    """The data kind of the bench workloads (oracle's table-driven bf16 generator, oz_synth_bf16_q16) at 256 MiB: xet-core's
    file hash pins the oracle's CDC + chunk hashes + Merkle fold + file hash on exactly those bytes."""
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from oracle.xet_oracle import synth_bf16_fast
    out = []
    for seed, n in ((7, 256 << 20), (1_000_003, 64 << 20)):
        data = synth_bf16_fast(n, seed).tobytes()
        fh = hash_file(data)
        out.append({"name": f"bf16_q16_seed{seed}", "n_bytes": n, "seed": seed, "sha256": hashlib.sha256(data).hexdigest(), "file_hash": fh})
        print(out[-1]["name"], fh)
    return out


def main():
    path = os.path.join(HERE, "vectors.json")
    if "--only-terms" in sys.argv:      # add the term vectors to an existing file without redoing the rest
        res = json.load(open(path))
        res["terms"] = term_kats()
    elif "--only-bench" in sys.argv:    # the same for the bench-data vectors
        res = json.load(open(path))
        res["bench"] = bench_kats()
    else:
        res = {"generator": "tests/golden/make_golden.py", "hf_xet": "1.5.0", "blake3_pypi": "1.0.8",
               "gear_table_sha256": check_gear_table(),
               "blake3": blake3_kats(), "xet": xet_kats(), "real_xorbs": real_xorbs(), "terms": term_kats(),
               "bench": bench_kats()}
    json.dump(res, open(path, "w"), indent=0)
    print("wrote", path)


if __name__ == "__main__":
    main()
