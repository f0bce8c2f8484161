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
