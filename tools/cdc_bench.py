#!/usr/bin/env python
"""Developer tool: Gearhash CDC (+ chunk hash) on a large device-resident buffer (BASELINE.json configs[3] shape).
  python tools/cdc_bench.py 8589934592
The buffer is bf16-like synthetic data generated on the device in pieces (torch), boundaries are spot-checked
against the oracle on the first 64 MiB."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import zest_b200 as z  # noqa: E402
from zest_b200 import _native  # noqa: E402
from zest_b200._native import check  # noqa: E402
from oracle import xet_oracle as oz  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 30
dev = torch.device("cuda", 0)
ctx = z.Context(0)
L = _native.lib()
g = torch.Generator(device=dev).manual_seed(1)
raw = torch.empty(n + 16, dtype=torch.uint8, device=dev)
step = 1 << 28
for o in range(0, n, step):
    m = min(step, n - o)
    x = (torch.randn(m // 2 + 1, generator=g, device=dev) * 0.02).view(torch.int32) >> 16
    raw[o:o + m] = x.to(torch.int16).view(torch.uint8)[:m]
cap = n // 8192 + 2
d_ends = torch.zeros(cap + 1, dtype=torch.int64, device=dev)
d_n = torch.zeros(4, dtype=torch.int64, device=dev)
s_cdc = torch.empty(L.zg_dev_cdc_scratch_bytes(n), dtype=torch.uint8, device=dev)
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def cdc():
    check(L.zg_dev_cdc(ctx.handle, raw.data_ptr(), n, d_ends.data_ptr() + 8, cap, d_n.data_ptr(), s_cdc.data_ptr(), s_cdc.numel(), sp))


def timed(fn, k=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k


ms = timed(cdc)
cnt = int(d_n[0])
print(f"cdc      : {n/1e9:.2f} GB  {ms:8.3f} ms  {n/ms/1e6:8.1f} GB/s  {cnt} chunks  overflow={int(d_n[1])}")
# spot check against the oracle on a prefix
m = min(n, 64 << 20)
want = oz.cdc(raw[:m].cpu().numpy())
got = d_ends[1:1 + want.size].cpu().numpy().astype(np.uint64)
ok = (got[:-1] == want[:-1]).all()
print("prefix boundaries match oracle:", bool(ok))
# chunk hashes of all chunks (offsets = [0, ends...])
d_hash = torch.empty(32 * cnt, dtype=torch.uint8, device=dev)
s_hash = torch.empty(L.zg_dev_hash_scratch_bytes(n, cnt), dtype=torch.uint8, device=dev)
hs = lambda: check(L.zg_dev_hash_chunks(ctx.handle, raw.data_ptr(), n, d_ends.data_ptr(), None, cnt, None, d_hash.data_ptr(),
                                         s_hash.data_ptr(), s_hash.numel(), sp))
ms2 = timed(hs)
print(f"hash     : {ms2:8.3f} ms  {n/ms2/1e6:8.1f} GB/s")
off = np.concatenate([[0], want[:8]]).astype(np.uint64)
assert (d_hash[:32 * 8].cpu().numpy().reshape(8, 32) == oz.hash_chunks(raw[:int(want[7])].cpu().numpy(), off)).all()
print(f"cdc+hash : {n/(ms+ms2)/1e6:8.1f} GB/s")
ctx.close()
