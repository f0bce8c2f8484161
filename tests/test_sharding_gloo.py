"""N>1 path on CPU: two gloo ranks shard a xorb list (LPT by compressed size), 'verify' their shard with the
oracle, and assemble verdicts and whole-job throughput exactly as bench.py does on GPUs.  No collective touches
the data path."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from oracle import xet_oracle as oz
    from zest_b200 import distributed as zd
    from zest_b200.verify import partition_xorbs
    r, w, _ = zd.init_from_env("gloo")
    blobs, expected = [], []
    for i in range(7):  # every rank builds the same global list; each verifies only its shard
        data = oz.synth_zipf_text(300_000 + 50_000 * i, i)
        x = oz.make_xorbs(data, 1)[0][0]
        blobs.append(x.blob); expected.append(x.expected)
    bad = bytearray(blobs[4]); bad[1000] ^= 8; blobs[4] = bytes(bad)
    parts = partition_xorbs([len(b) for b in blobs], w)
    mine = parts[r]
    v, roots, nbytes = oz.verify_xorbs([blobs[i] for i in mine], [expected[i] for i in mine])
    zd.barrier()
    allv = zd.gather_verdicts(mine, v, len(blobs))
    total = zd.reduce_sum(nbytes)
    tmax = zd.reduce_max(0.5 + r)
    q.put((r, [int(x) for x in allv], total, tmax, mine))


def test_two_rank_sharding():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    (r0, v0, t0, m0, p0), (r1, v1, t1, m1, p1) = res
    assert v0 == v1 and sorted(p0 + p1) == list(range(7)) and not set(p0) & set(p1)
    assert [int(x == 1) for x in v0] == [1, 1, 1, 1, 0, 1, 1]
    assert t0 == t1 and m0 == m1 == 1.5


class _DictShard:
    """CPU stand-in for the device shard of ShardedDedupIndex (test infrastructure: same semantics as dedup.cu)."""

    def __init__(self):
        self.table = {}

    def insert_t(self, hashes_t, locs_t):
        import torch
        keys = [bytes(h.tolist()) for h in hashes_t]
        fresh = {k for k in keys if k not in self.table}
        for k, l in zip(keys, locs_t.tolist()):
            self.table[k] = min(self.table.get(k, 1 << 62), int(l))
        return torch.tensor([int(k in fresh and self.table[k] == int(l)) for k, l in zip(keys, locs_t.tolist())], dtype=torch.uint8)

    def lookup_t(self, hashes_t):
        import torch
        return torch.tensor([self.table.get(bytes(h.tolist()), -1) for h in hashes_t], dtype=torch.int64)


def _dedup_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch
    from zest_b200 import distributed as zd
    from zest_b200.dedup import ShardedDedupIndex, shard_of
    r, w, _ = zd.init_from_env("gloo")
    shard = _DictShard()
    idx = ShardedDedupIndex(0, local=shard)
    rng = np.random.default_rng(100)                       # same pool on both ranks
    pool = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    mine = np.random.default_rng(r).integers(0, 4000, 3000)  # overlapping picks, different per rank
    hashes = torch.from_numpy(pool[mine])
    locs = torch.from_numpy((mine.astype(np.int64) << 8) | r)   # location encodes (pool index, inserting rank)
    new = idx.insert(hashes, locs)
    zd.barrier()
    found = idx.lookup(torch.from_numpy(pool))
    owned_ok = all(int(shard_of(np.frombuffer(k, np.uint8), w)[0]) == r for k in shard.table)
    q.put((r, mine.tolist(), new.tolist(), found.tolist(), len(shard.table), owned_ok))
    zd.shutdown()


def test_two_rank_sharded_dedup_index():
    """Routing of the split dedup index over two gloo ranks: every hash lands on its owner, answers come back in
    input order, both ranks see the same global answer, and it equals a single dict fed with both ranks' batches."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_dedup_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=180) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, m0, n0, f0, c0, ok0), (_, m1, n1, f1, c1, ok1) = res
    assert ok0 and ok1 and f0 == f1
    want = {}
    for r, m in ((0, m0), (1, m1)):
        for i in m:
            want[i] = min(want.get(i, 1 << 62), (i << 8) | r)
    assert c0 + c1 == len(want)
    assert f0 == [want.get(i, -1) for i in range(4000)]
    # was_new: exactly one occurrence per distinct hash overall claims it (the one whose location was kept);
    # duplicates of a hash inside one rank's batch carry the same location and all report it
    kept = {}
    for r, m, n in ((0, m0, n0), (1, m1, n1)):
        for i, flag in zip(m, n):
            if flag:
                kept.setdefault(i, set()).add(r)
    assert set(kept) == set(want) and all(len(v) == 1 for v in kept.values())
    assert all(((want[i] & 0xff) in v) for i, v in kept.items())


def _set_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    import bench
    from oracle import xet_oracle as oz
    from zest_b200 import distributed as zd
    r, w, _ = zd.init_from_env("gloo")
    bench.XORB_RAW = 1 << 20          # 1 MiB xorbs so the set is small

    def all_gather_max(a):
        t = torch.from_numpy(a.copy())
        if w > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.numpy()
    xs, n_all, comp_all, part_bytes = bench.build_xorb_set("llama8b-bg4", 7 * (1 << 20) + 12345, 3, r, w, 2,
                                                           lambda b: np.zeros(b, np.uint8), all_gather_max, lambda m: None)
    v, roots, nb = oz.verify_xorbs(xs.views, xs.expected, threads=2, want_chunk_hashes=False)[:3]
    ok = bool((v == 1).all()) and nb == xs.total and np.array_equal(roots, xs.expected)
    q.put((r, xs.index, xs.sizes, [bytes(b[:64]) + bytes(b[-64:]) for b in xs.views], n_all, comp_all, part_bytes, ok,
           int(xs.first[-1]), xs.hashes.shape[0], xs.total))
    zd.shutdown()


def test_bench_set_is_one_global_list_partitioned_over_ranks():
    """bench.build_xorb_set on two gloo ranks against one rank: the same global xorb list (same bytes, whoever built
    them), split by partition_xorbs on compressed bytes, every rank's part verifies under the oracle."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_set_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=180) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    q1 = ctx.Queue()
    p = ctx.Process(target=_set_worker, args=(0, 1, port + 1 if port < 65000 else port - 1, q1))
    p.start()
    one = q1.get(timeout=180)
    p.join(timeout=60)
    assert p.exitcode == 0 and one[7] and one[1] == list(range(8))
    (_, i0, s0, e0, n0, c0, pb0, ok0, nc0, nh0, t0), (_, i1, s1, e1, n1, c1, pb1, ok1, nc1, nh1, t1) = res
    assert ok0 and ok1 and n0 == n1 == 8 and c0 == c1 == one[5] and pb0 == pb1
    assert sorted(i0 + i1) == list(range(8)) and not set(i0) & set(i1)
    from zest_b200.verify import partition_xorbs
    assert [i0, i1] == partition_xorbs(one[2], 2)                      # the split is LPT over the true compressed sizes
    assert pb0 == [sum(one[2][i] for i in i0), sum(one[2][i] for i in i1)]
    for idx, sizes, edges in ((i0, s0, e0), (i1, s1, e1)):
        for i, sz, ed in zip(idx, sizes, edges):
            assert sz == one[2][i] and ed == one[3][i]                   # same blob whoever built it
    assert nc0 == nh0 and nc1 == nh1 and t0 + t1 == one[10] == 7 * (1 << 20) + 12345
