#!/usr/bin/env python
"""Developer tool: insert / lookup rates of the device-resident dedup index on device-resident hashes.
  python tools/dedup_bench.py [n_hashes]      (default 2.2 M: the chunk count of BASELINE.json configs[4], 140 GB)
Reports M hashes/s and the HBM traffic that implies at ~3 sectors (96 B) per probe plus the 32-byte hash read."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import zest_b200 as z  # noqa: E402
from zest_b200 import _native  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_200_000
    ctx = z.Context(0)
    L = _native.lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(1)
    d_h = torch.randint(0, 256, (n, 32), dtype=torch.uint8, device=dev, generator=g)
    d_miss = torch.randint(0, 256, (n, 32), dtype=torch.uint8, device=dev, generator=g)
    d_l = torch.arange(n, dtype=torch.int64, device=dev)
    d_new = torch.zeros(n, dtype=torch.uint8, device=dev)
    d_out = torch.zeros(n, dtype=torch.int64, device=dev)
    sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def timed(fn, reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    best = 1e9
    for _ in range(3):                      # a fresh table each time: inserts of new keys
        idx = z.DedupIndex(n, ctx)
        ms = timed(lambda: _native.check(L.zg_dev_dedup_insert(idx.handle, d_h.data_ptr(), d_l.data_ptr(), n, d_new.data_ptr(), sp)), 1)
        best = min(best, ms)
        assert len(idx) == n
        if _ < 2:
            idx.close()
    print(f"insert (new keys)  : {best:8.3f} ms  {n/best/1e3:8.1f} M hashes/s")
    ms = timed(lambda: _native.check(L.zg_dev_dedup_insert(idx.handle, d_h.data_ptr(), d_l.data_ptr(), n, d_new.data_ptr(), sp)), 5)
    print(f"insert (all known) : {ms:8.3f} ms  {n/ms/1e3:8.1f} M hashes/s")
    ms = timed(lambda: _native.check(L.zg_dev_dedup_lookup(idx.handle, d_h.data_ptr(), n, d_out.data_ptr(), sp)), 10)
    print(f"lookup (hits)      : {ms:8.3f} ms  {n/ms/1e3:8.1f} M hashes/s  ~{n*(96+32+8)/ms/1e6:7.1f} GB/s of sectors")
    assert bool((d_out == d_l).all())
    ms = timed(lambda: _native.check(L.zg_dev_dedup_lookup(idx.handle, d_miss.data_ptr(), n, d_out.data_ptr(), sp)), 10)
    print(f"lookup (misses)    : {ms:8.3f} ms  {n/ms/1e3:8.1f} M hashes/s")
    assert bool((d_out == -1).all())


if __name__ == "__main__":
    main()
