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
