#!/usr/bin/env python
"""Developer tool: throughput of the reconstruct call (zg_extract_chunk_range: host blob in, decoded bytes out to
pinned host memory) with 1..N concurrent callers, the way zest's task group would use it (one call per term).
  python tools/extract_bench.py [workload] [bytes] [callers,...]"""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import zest_b200 as z  # noqa: E402
from zest_b200 import _native  # noqa: E402
from zest_b200._native import check  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "gpt2-lz4"
    nbytes = int(sys.argv[2]) if len(sys.argv) > 2 else 550_000_000
    callers = [int(x) for x in (sys.argv[3] if len(sys.argv) > 3 else "1,2,4,8").split(",")]
    blobs, expected, gold, total, nchunks = bench.make_workload(wl, nbytes, 1, os.cpu_count())
    ctx = z.Context(0)
    L = _native.lib()
    jobs = []
    for b in blobs:
        ix = z.xorb_index(b)
        src = ctx.pinned_empty(len(b) + 64)
        src[:len(b)] = np.frombuffer(b, np.uint8)
        dst = ctx.pinned_empty(int(ix.ulen.sum()) + 64)
        jobs.append((src, len(b), ix.n_chunks, dst, int(ix.ulen.sum())))
    print(f"{wl}: {total/1e6:.0f} MB in {len(jobs)} xorbs, {nchunks} chunks")

    def run(nthreads, reps):
        nxt, lock = [0], threading.Lock()
        order = list(range(len(jobs))) * reps

        def worker():
            got = C.c_uint64()
            while True:
                with lock:
                    if nxt[0] >= len(order):
                        return
                    k = order[nxt[0]]
                    nxt[0] += 1
                src, n, nch, dst, ul = jobs[k]
                check(L.zg_extract_chunk_range(ctx.handle, src.ctypes.data, n, 0, nch, dst.ctypes.data, ul, C.byref(got)))
        ts = [threading.Thread(target=worker) for _ in range(nthreads)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return (time.perf_counter() - t0) / reps

    run(2, 1)
    for c in callers:
        dt = run(c, 3)
        print(f"callers={c:2d}: {dt*1e3:8.2f} ms per pass  {total/dt/1e9:7.1f} GB/s decoded bytes delivered to the host")


if __name__ == "__main__":
    main()
