"""Shared fixtures.  GPU tests are marked `gpu`; everything else must pass on a CPU-only box."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # the product library is built in-tree by __graft_entry__.build(); build it here if a fresh checkout has none
    # (nvcc cross-compiles without a GPU).  The tests never fall back to anything else.
    lib = os.path.join(ROOT, "zest_b200", "libzest_gpu.so")
    if not os.path.exists(lib):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "zest_b200", "csrc"), "-j8"], stdout=subprocess.DEVNULL)


@pytest.fixture(scope="session")
def vectors():
    return json.load(open(os.path.join(GOLDEN, "vectors.json")))


@pytest.fixture(scope="session")
def oracle():
    from oracle import xet_oracle
    xet_oracle.lib()
    return xet_oracle


def golden_input(name: str) -> np.ndarray:
    """Regenerate the input bytes of a vectors.json 'xet' record (recipes of tests/golden/make_golden.py)."""
    from oracle.xet_oracle import synth_zipf_text
    rng = np.random.default_rng
    if name == "hello_world":
        return np.frombuffer(b"hello world", np.uint8)
    if name == "empty":
        return np.zeros(0, np.uint8)
    if name == "seed0_1MiB":
        return rng(0).integers(0, 256, 1 << 20, dtype=np.uint8)
    if name == "seed0_1GiB":
        return rng(0).integers(0, 256, 1 << 30, dtype=np.uint8)
    if name in ("seed1_8MiB", "seed1_next_8191", "seed1_next_8192"):
        g = rng(1)
        a = g.integers(0, 256, 8 << 20, dtype=np.uint8)
        b = g.integers(0, 256, 8191, dtype=np.uint8)
        c = g.integers(0, 256, 8192, dtype=np.uint8)
        return {"seed1_8MiB": a, "seed1_next_8191": b, "seed1_next_8192": c}[name]
    if name == "seed4_70MiB":
        return rng(4).integers(0, 256, 70 << 20, dtype=np.uint8)
    if name == "zeros_1MiB":
        return np.zeros(1 << 20, np.uint8)
    if name == "bf16_seed2_4MiB":
        x = rng(2).standard_normal(2 << 20).astype(np.float32).view(np.uint32) >> 16
        return x.astype(np.uint16).view(np.uint8)
    if name == "zipf_seed3_2MiB":
        return synth_zipf_text(2 << 20, 3)
    raise KeyError(name)


def read_golden_xorb(rec):
    raw = open(os.path.join(GOLDEN, rec["file"]), "rb").read()
    return raw[:rec["stream_len"]], raw


def footer_chunk_hashes(raw: bytes):
    """Chunk hashes and uncompressed end offsets from a XETBLOB footer."""
    import struct
    info_len = struct.unpack("<I", raw[-4:])[0]
    foot = raw[-4 - info_len:-4]
    assert foot[:7] == b"XETBLOB"
    p = 8 + 32 + 8
    n = struct.unpack("<I", foot[p:p + 4])[0]
    p += 4
    hashes = np.frombuffer(foot[p:p + 32 * n], np.uint8).reshape(n, 32).copy()
    p += 32 * n + 8 + 4 + 4 * n
    unc = np.array(struct.unpack("<%dI" % n, foot[p:p + 4 * n]), np.uint64)
    return hashes, unc, foot[8:40]


# ---------------------------------------------------------------------------------------------------------------
# Hand-rolled LZ4 streams: sequences liblz4 / lz4_flex would never emit but the format allows (multi-byte length
# extensions, matches longer than 64 KiB, offsets of 1 and 65535, empty literal runs, many tiny sequences).
# ---------------------------------------------------------------------------------------------------------------
def craft_lz4_block(rng, target_len: int, style: str):
    """-> (block bytes, decoded bytes).  A valid LZ4 block of exactly target_len decoded bytes."""
    out = bytearray()
    blk = bytearray()

    def emit(lit: bytes, mlen: int, off: int, last: bool):
        tok_l = min(len(lit), 15)
        tok_m = 0 if last else min(mlen - 4, 15)
        blk.append((tok_l << 4) | tok_m)
        if tok_l == 15:
            r = len(lit) - 15
            while r >= 255:
                blk.append(255); r -= 255
            blk.append(r)
        blk.extend(lit)
        out.extend(lit)
        if last:
            return
        blk.extend(int(off).to_bytes(2, "little"))
        if tok_m == 15:
            r = mlen - 4 - 15
            while r >= 255:
                blk.append(255); r -= 255
            blk.append(r)
        for _ in range(mlen):
            out.append(out[-off])

    while True:
        remaining = target_len - len(out)
        if style == "tiny":
            lit_n = int(rng.integers(0, 4))
            mlen = int(rng.integers(4, 9))
        elif style == "edge":
            # around what the decoder's hop table holds: a sequence of up to 255 stream bytes (token, one length byte
            # each, literals, offset) -- literal runs of 225-260 bytes, match lengths of 255-300, and short ones between
            lit_n = int(rng.choice([0, 3, 14, 15, 16, int(rng.integers(225, 261)), int(rng.integers(225, 261))]))
            mlen = int(rng.choice([4, 18, 19, int(rng.integers(255, 301)), int(rng.integers(255, 301))]))
        elif style == "long":
            lit_n = int(rng.choice([0, 1, 15, 16, 17, 270, 271, 600, 70000]))
            mlen = int(rng.choice([4, 18, 19, 20, 21, 274, 275, 70000, 140000]))
        else:
            lit_n = int(rng.integers(0, 40))
            mlen = int(rng.integers(4, 60))
        if lit_n + mlen + 5 >= remaining or len(out) + lit_n == 0:
            if remaining > 0 and len(out) + remaining > 0:
                emit(bytes(rng.integers(0, 256, remaining, dtype=np.uint8)), 0, 0, True)
            else:
                emit(b"", 0, 0, True)
            break
        lit = bytes(rng.integers(0, 256, lit_n, dtype=np.uint8))
        avail = len(out) + lit_n
        if avail == 0:
            continue
        choice = int(rng.integers(0, 5))
        off = [1, 2, min(avail, 65535), int(rng.integers(1, min(avail, 65535) + 1)), min(avail, 3)][choice]
        emit(lit, mlen, off, False)
    return bytes(blk), bytes(out)


def frame_single_block(block: bytes, ulen: int) -> bytes:
    """LZ4 frame around one compressed block, the way xet-core frames chunks."""
    from oracle.xet_oracle import lib
    import ctypes as C
    bd = 0x70  # 4 MiB blocks: crafted blocks may be large
    desc = bytes([0x60, bd])
    hdr = np.frombuffer(desc, np.uint8).copy()
    hc = (lib().oz_xxh32(hdr.ctypes.data, 2, 0) >> 8) & 0xFF
    return b"\x04\x22\x4d\x18" + desc + bytes([hc]) + len(block).to_bytes(4, "little") + block + b"\x00\x00\x00\x00"


def chunk_record(payload: bytes, scheme: int, ulen: int) -> bytes:
    return bytes([0]) + len(payload).to_bytes(3, "little") + bytes([scheme]) + ulen.to_bytes(3, "little") + payload
