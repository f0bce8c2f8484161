#!/usr/bin/env python
"""Single-process, several-GPU verify through zg_group_verify_xorbs (what a one-process caller like zest would use):
end-to-end GB/s with pinned host buffers, beside bench.py's torchrun figure.  Not part of bench.py's contract.
  python tools/group_bench.py --gpus 8 --bytes 16000000000 --steps 5
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import zest_b200 as z  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0, help="0: every visible device")
    ap.add_argument("--workload", default="llama8b-bg4")
    ap.add_argument("--bytes", type=int, default=16_000_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    import torch
    n_dev = args.gpus or torch.cuda.device_count()
    g = z.Group(list(range(n_dev)))
    ctx0 = g.context(0)
    cores = os.cpu_count() or 1
    xs, n_all, comp_all, _ = bench.build_xorb_set(args.workload, args.bytes, 1, 0, 1, cores, ctx0.pinned_empty, lambda a: a, bench.log)
    views, exp = xs.views, xs.expected
    for _ in range(args.warmup):
        r = g.verify_xorbs(views, exp)
        assert r.ok and np.array_equal(r.roots, exp) and r.bytes_verified == xs.total
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = g.verify_xorbs(views, exp)
    dt = (time.perf_counter() - t0) / args.steps
    assert r.ok
    per_dev = [g.context(i).stats() for i in range(n_dev)]
    parts = z.partition_native(xs.sizes, n_dev)
    print(json.dumps({"what": "zg_group_verify_xorbs, one process, pinned host buffers", "n_gpus": n_dev, "workload": args.workload,
                      "uncompressed_bytes": xs.total, "compressed_bytes": comp_all, "n_xorbs": n_all, "steps": args.steps,
                      "ms_per_step": round(dt * 1e3, 2), "e2e_GBps": round(xs.total / dt / 1e9, 2),
                      "h2d_GBps_aggregate": round(comp_all / dt / 1e9, 2),
                      "xorbs_per_device": [len(p) for p in parts],
                      "h2d_bytes_per_device_per_step": [int(s["h2d_bytes"] // (args.steps + args.warmup)) for s in per_dev]}))
    g.close()


if __name__ == "__main__":
    main()
