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
