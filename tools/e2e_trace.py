import sys, time, os, numpy as np
sys.path.insert(0, '/root/repo')
import bench, zest_b200 as z
blobs, expected, gold, total, nchunks = bench.make_workload('gpt2-lz4', 550000000, 1, os.cpu_count())
ctx = z.Context(0)
host = ctx.pinned_empty(sum(len(b) for b in blobs) + 64)
views, p = [], 0
for b in blobs:
    host[p:p+len(b)] = np.frombuffer(b, np.uint8); views.append(host[p:p+len(b)]); p += len(b)
for _ in range(3): r = z.verify_xorbs(views, expected, ctx=ctx)
print("---- traced call", file=sys.stderr, flush=True)
t0=time.perf_counter(); z.verify_xorbs(views, expected, ctx=ctx); print("wall %.2f ms" % ((time.perf_counter()-t0)*1e3), file=sys.stderr)
