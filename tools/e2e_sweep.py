#!/usr/bin/env python
"""Developer tool: time zg_verify_xorbs (host buffers in, verdicts out) for several sub-batch sizes / lane counts.
Each setting runs in a fresh process because the knobs are read once."""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, time, os, numpy as np
sys.path.insert(0, %r)
import bench, zest_b200 as z
blobs, expected, gold, total, nchunks = bench.make_workload(sys.argv[1], int(sys.argv[2]), 1, os.cpu_count())
ctx = z.Context(0)
host = ctx.pinned_empty(int(sum(len(b) for b in blobs)) + 64)
views, p = [], 0
for b in blobs:
    host[p:p+len(b)] = np.frombuffer(b, np.uint8); views.append(host[p:p+len(b)]); p += len(b)
for _ in range(3): r = z.verify_xorbs(views, expected, ctx=ctx)
assert r.ok
t0 = time.perf_counter(); K = 8
for _ in range(K): z.verify_xorbs(views, expected, ctx=ctx)
dt = (time.perf_counter() - t0) / K
print(f"{dt*1e3:8.2f} ms  {total/dt/1e9:7.1f} GB/s uncompressed  {p/dt/1e9:6.1f} GB/s H2D")
''' % ROOT
wl, nbytes = sys.argv[1], sys.argv[2]
for mb in [int(x) for x in (sys.argv[3] if len(sys.argv) > 3 else '64,128,256,1024').split(',')]:
    for lanes in [int(x) for x in (sys.argv[4] if len(sys.argv) > 4 else '1,2,4').split(',')]:
        env = dict(os.environ, ZEST_VERIFY_SUBBATCH_MB=str(mb), ZEST_VERIFY_LANES=str(lanes))
        out = subprocess.run([sys.executable, "-c", CODE, wl, nbytes], env=env, capture_output=True, text=True)
        print(f"subbatch={mb:5d} MB lanes={lanes}: {out.stdout.strip() or out.stderr.strip()[-300:]}", flush=True)
