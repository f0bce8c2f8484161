"""Two-GPU checks (skipped on a single-GPU box): the dedup index split across ranks with NCCL all-to-all routing."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch
    import zest_b200 as z
    from zest_b200 import distributed as zd
    from zest_b200.dedup import ShardedDedupIndex
    r, w, local = zd.init_from_env("nccl")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    ctx = z.Context(local)
    idx = ShardedDedupIndex(400_000, ctx=ctx)
    pool = np.random.default_rng(100).integers(0, 256, (300_000, 32), dtype=np.uint8)
    mine = np.random.default_rng(r).integers(0, pool.shape[0], 200_000)
    hashes = torch.from_numpy(pool[mine]).to(dev)
    locs = torch.from_numpy((mine.astype(np.int64) << 8) | r).to(dev)
    new = idx.insert(hashes, locs)
    torch.cuda.synchronize()
    zd.barrier()
    found = idx.lookup(torch.from_numpy(pool).to(dev))
    # timing of a lookup round trip (device events, max over ranks is taken by the parent)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    zd.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        idx.lookup(hashes)
    e1.record(); torch.cuda.synchronize()
    q.put((r, mine.tolist(), new.cpu().tolist(), found.cpu().tolist(), len(idx.local), e0.elapsed_time(e1) / 5))
    zd.shutdown()


def test_sharded_dedup_index_nccl():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=300) for _ in ps)
    for p in ps:
        p.join(timeout=120)
        assert p.exitcode == 0
    (_, m0, n0, f0, c0, t0), (_, m1, n1, f1, c1, t1) = res
    want = {}
    for r, m in ((0, m0), (1, m1)):
        for i in m:
            want[i] = min(want.get(i, 1 << 62), (i << 8) | r)
    assert f0 == f1 == [want.get(i, -1) for i in range(300_000)]
    assert c0 + c1 == len(want) and abs(c0 - c1) < 0.05 * len(want)
    kept = {}
    for r, m, n in ((0, m0, n0), (1, m1, n1)):
        for i, flag in zip(m, n):
            if flag:
                kept.setdefault(i, set()).add(r)
    assert set(kept) == set(want) and all(v == {want[i] & 0xff} for i, v in kept.items())
    print(f"sharded lookup of 200k hashes per rank: {max(t0, t1):.3f} ms")


def _group_case(oz, n_x=9):
    blobs, expected, gold = [], [], []
    for i in range(n_x):
        data = oz.synth_zipf_text(400_000 + 90_000 * i, 40 + i) if i % 3 else oz.synth_bf16(500_000 + 64 * i, 40 + i)
        xs, _, hs = oz.make_xorbs(data, 1 if i % 3 else 2)
        blobs.append(xs[0].blob); expected.append(xs[0].expected); gold.append(hs[:xs[0].n_chunks])
    return blobs, expected, gold


@pytest.mark.gpu
@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 0, 0], "all"])
def test_group_verify_matches_oracle_in_caller_order(devices):
    """One process, several contexts (zg_group): the batch is LPT-partitioned, every part verified through its own
    context, verdicts / roots / chunk hashes come back in the caller's order and equal the oracle's.  [0, 0] and
    [0, 0, 0] put two and three contexts on one GPU, so the split-and-reassemble logic is exercised on a one-GPU box
    too; "all" takes every visible device."""
    import torch
    import zest_b200 as z
    from oracle import xet_oracle as oz
    blobs, expected, gold = _group_case(oz)
    bad = bytearray(blobs[4]); bad[len(bad) // 2] ^= 0x21; blobs[4] = bytes(bad)      # payload damage
    blobs[7] = blobs[7][:len(blobs[7]) - 11]                                          # truncated stream
    with z.Group(None if devices == "all" else devices) as g:
        assert g.size == (torch.cuda.device_count() if devices == "all" else len(devices))
        r = g.verify_xorbs(blobs, expected, want_chunk_hashes=True)
        r2 = g.verify_xorbs(blobs, expected)                      # without chunk hashes (no counting pass)
        assert g.context(0).stats()["calls"] > 0
    v, roots, nb, ch = oz.verify_xorbs(blobs, expected, threads=4, want_chunk_hashes=True)
    assert np.array_equal(r.verdict, v) and np.array_equal(r2.verdict, v)
    assert r.verdict[4] != 1 and r.verdict[7] == -3 and (np.delete(r.verdict, [4, 7]) == 1).all()
    ok = [i for i in range(len(blobs)) if v[i] >= 0]
    assert np.array_equal(r.roots[ok], roots[ok]) and np.array_equal(r2.roots[ok], roots[ok])
    want_bytes = sum(int(oz.xorb_index(blobs[i]).total_ulen) for i in range(len(blobs)) if v[i] >= 0)  # xorbs that decoded
    assert r.bytes_verified == want_bytes == r2.bytes_verified
    # chunk hashes: back to back in xorb order; the truncated blob (walk error) contributes none
    pos = 0
    for i, gh in enumerate(gold):
        if i == 7:
            continue
        k = gh.shape[0]
        if i != 4:
            assert np.array_equal(r.chunk_hashes[pos:pos + k], gh), i
        pos += k
    assert pos == r.chunk_hashes.shape[0]


@pytest.mark.gpu
def test_group_on_two_gpus_uses_both():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import zest_b200 as z
    from oracle import xet_oracle as oz
    blobs, expected, _ = _group_case(oz, 8)
    with z.Group([0, 1]) as g:
        r = g.verify_xorbs(blobs, expected)
        s0, s1 = g.context(0).stats(), g.context(1).stats()
    assert r.ok and s0["h2d_bytes"] > 0 and s1["h2d_bytes"] > 0
    parts = z.partition_native([len(b) for b in blobs], 2)
    assert abs(s0["h2d_bytes"] - sum(len(blobs[i]) for i in parts[0])) < 200_000


@pytest.mark.gpu
def test_calls_leave_the_callers_device_alone():
    """A context on device 1 used from a thread whose current device is 0 (torch on cuda:0 next to a context elsewhere,
    a zg_group worker next to its caller): every entry point works on its context's device and puts the thread's
    device back -- checked through the CUDA runtime torch itself uses."""
    import ctypes as C
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import zest_b200 as z
    from oracle import xet_oracle as oz
    rt = C.CDLL("libcudart.so.12")

    def current():
        d = C.c_int(-1)
        assert rt.cudaGetDevice(C.byref(d)) == 0
        return d.value
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda:0")
    assert current() == 0
    ctx = z.Context(1)
    assert current() == 0
    data = oz.synth_zipf_text(600_000, 3)
    xs, ends, hashes = oz.make_xorbs(data, 1)
    r = z.verify_xorbs([xs[0].blob], [xs[0].expected], ctx=ctx)
    assert r.ok and current() == 0
    assert z.XorbReader(xs[0].blob, ctx).extract_chunk_range(0, xs[0].n_chunks) == data.tobytes() and current() == 0
    assert np.array_equal(z.cdc(data, ctx=ctx), oz.cdc(data)) and current() == 0
    p = ctx.pinned_empty(4096)
    ctx.pinned_free(p)
    assert ctx.stats()["h2d_bytes"] > 0 and current() == 0
    ctx.close()
    assert current() == 0
