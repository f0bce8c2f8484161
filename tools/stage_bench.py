#!/usr/bin/env python
"""Developer tool: time the decode / hash / merkle / cdc kernels alone on one workload, optionally sweeping the
decode kernel's grid size (ZEST_LZ4_WARPS).  Not part of the product or of bench.py's contract.
  python tools/stage_bench.py --workload gpt2-lz4 --bytes 550000000 --warps 1480,2220,2960
With a library built with -DZG_LZ4_PROFILE, ZG_PROF=1 prints the decoder's phase counters.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import zest_b200 as z  # noqa: E402
from zest_b200 import _native  # noqa: E402
from zest_b200._native import ChunkDesc, check  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="gpt2-lz4")
    ap.add_argument("--bytes", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warps", default="")
    ap.add_argument("--cdc", action="store_true")
    ap.add_argument("--regroup", action="store_true", help="time the decode of the reconstruct path (byte-grouped chunks regrouped "
                    "into the output) and the hash over plain bytes, instead of the verify path's halves")
    args = ap.parse_args()
    nbytes = args.bytes or bench.WORKLOADS[args.workload]["nbytes"]
    dev = torch.device("cuda", 0)
    ctx = z.Context(0)
    L = _native.lib()
    t0 = time.time()
    xs, n_all, comp_bytes, _ = bench.build_xorb_set(args.workload, nbytes, 1, 0, 1, os.cpu_count(), lambda b: np.zeros(b, np.uint8),
                                                    lambda a: a, lambda m: None)
    total, gold, expected = xs.total, xs.hashes, xs.expected
    nx = len(xs.index)
    descs, out_bytes = bench.build_descs(xs, L)
    nchunks = descs.size
    print(f"workload {args.workload}: {total/1e6:.0f} MB, {nx} xorbs, {nchunks} chunks, ratio {comp_bytes/total:.3f}, built in {time.time()-t0:.1f}s", flush=True)
    cpos = xs.offsets[-1] + (xs.sizes[-1] + 15) // 16 * 16
    d_comp = torch.from_numpy(xs.buf[:cpos + 64]).to(dev)
    d_desc = torch.from_numpy(descs.view(np.uint8).reshape(-1)).to(dev)
    d_first = torch.from_numpy(xs.first.view(np.uint8).copy()).to(dev)
    d_exp = torch.from_numpy(expected.reshape(-1).copy()).to(dev)
    off = descs["dst_off"].copy()
    lens = descs["ulen"].copy()
    d_off, d_len = torch.from_numpy(off.view(np.uint8).copy()).to(dev), torch.from_numpy(lens.view(np.uint8).copy()).to(dev)
    d_out = torch.zeros(out_bytes + 16, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(nchunks, dtype=torch.int32, device=dev)
    d_hashes = torch.zeros(32 * nchunks, dtype=torch.uint8, device=dev)
    d_roots = torch.zeros(32 * nx, dtype=torch.uint8, device=dev)
    d_verdict = torch.zeros(nx, dtype=torch.int8, device=dev)
    s_dec = torch.empty(L.zg_dev_decode_scratch_bytes(nchunks, out_bytes), dtype=torch.uint8, device=dev)
    s_hash = torch.empty(max(L.zg_dev_hash_scratch_bytes(out_bytes, nchunks), L.zg_dev_hash_decoded_scratch_bytes(out_bytes, nchunks)),
                         dtype=torch.uint8, device=dev)
    s_mer = torch.empty(L.zg_dev_merkle_scratch_bytes(nchunks, nx), dtype=torch.uint8, device=dev)
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def k_decode():
        f = L.zg_dev_decode_chunks if args.regroup else L.zg_dev_decode_for_hash
        check(f(ctx.handle, d_comp.data_ptr(), cpos, d_desc.data_ptr(), nchunks, d_out.data_ptr(), out_bytes,
                d_status.data_ptr(), s_dec.data_ptr(), s_dec.numel(), sp))

    def k_hash():
        if args.regroup:
            check(L.zg_dev_hash_chunks(ctx.handle, d_out.data_ptr(), out_bytes, d_off.data_ptr(), d_len.data_ptr(), nchunks, None,
                                       d_hashes.data_ptr(), s_hash.data_ptr(), s_hash.numel(), sp))
        else:
            check(L.zg_dev_hash_decoded(ctx.handle, d_out.data_ptr(), out_bytes, d_desc.data_ptr(), nchunks, s_dec.data_ptr(),
                                        d_hashes.data_ptr(), s_hash.data_ptr(), s_hash.numel(), sp))

    def k_merkle():
        check(L.zg_dev_merkle(ctx.handle, d_hashes.data_ptr(), d_len.data_ptr(), d_first.data_ptr(), nx, nchunks, d_status.data_ptr(),
                              d_exp.data_ptr(), d_roots.data_ptr(), d_verdict.data_ptr(), s_mer.data_ptr(), s_mer.numel(), sp))

    def timed(fn, steps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    def check_all():
        k_decode(); k_hash(); k_merkle()
        torch.cuda.synchronize()
        ok = int(d_status.abs().sum()) == 0 and bool((d_verdict == 1).all()) and np.array_equal(d_hashes.cpu().numpy().reshape(-1, 32), gold)
        return ok

    res = {}
    sweeps = [""]
    warps = [w for w in args.warps.split(",") if w] or [""]
    overs = [""]
    for w in [(a, b) for a in warps for b in overs]:
        w, ov = w
        for s in sweeps:
            os.environ.pop("ZEST_LZ4_WARPS", None)
            if w:
                os.environ["ZEST_LZ4_WARPS"] = w
            d_out.zero_()
            ok = check_all()
            ms = timed(k_decode, args.steps)
            if os.environ.get("ZG_PROF"):
                g8 = (C.c_ulonglong * 8)()
                C.CDLL(_native.LIB_PATH).zg_lz4_profile_read(g8, 1)
                k_decode(); torch.cuda.synchronize()
                C.CDLL(_native.LIB_PATH).zg_lz4_profile_read(g8, 0)
                g = list(g8)
                print(f"      execute steps {g[3]}: cycles/step pass1 {g[0]/max(g[3],1):.0f} long copies {g[1]/max(g[3],1):.0f} match rounds {g[2]/max(g[3],1):.0f}; rounds/step {g[4]/max(g[3],1):.2f} pending/round {g[5]/max(g[4],1):.1f} long lits/step {g[6]/max(g[3],1):.2f} long old matches/step {g[7]/max(g[3],1):.2f}")
                pr = s_dec[16:88].cpu().numpy().view(np.uint64)
                tot = float(pr[:4].sum()) or 1.0
                print(f"      phase cycles: parse {pr[0]/tot:.3f} resolve {pr[1]/tot:.3f} execute {pr[2]/tot:.3f} serial {pr[3]/tot:.3f}  spec rounds {pr[4]} serial rounds {pr[5]}  true records {pr[6]} parsed records {pr[7]} pending matches {pr[8]}  total warp-cycles {tot:.3e}")
            print(f"decode warps={w or 'auto':>5}: {ms:8.3f} ms  {total/ms/1e6:8.1f} GB/s uncompressed  {(total+comp_bytes)/ms/1e6:8.1f} GB/s alg  parity={'ok' if ok else 'FAIL'}", flush=True)
            res[f"decode_s{s}_w{w}"] = ms
    os.environ.pop("ZEST_LZ4_WARPS", None)
    ms = timed(k_hash, args.steps)
    print(f"hash   : {ms:8.3f} ms  {total/ms/1e6:8.1f} GB/s")
    ms = timed(k_merkle, args.steps)
    print(f"merkle : {ms:8.3f} ms")
    if args.cdc:
        raw = torch.from_numpy(np.frombuffer(getattr(__import__('oracle.xet_oracle', fromlist=['x']), bench.WORKLOADS[args.workload]['gen'])(min(total, 1 << 30), 5), np.uint8).copy()).to(dev)
        n = raw.numel()
        cap = n // 8192 + 2
        d_ends = torch.zeros(cap, dtype=torch.int64, device=dev)
        d_n = torch.zeros(4, dtype=torch.int64, device=dev)
        s_cdc = torch.empty(L.zg_dev_cdc_scratch_bytes(n), dtype=torch.uint8, device=dev)
        f = lambda: check(L.zg_dev_cdc(ctx.handle, raw.data_ptr(), n, d_ends.data_ptr(), cap, d_n.data_ptr(), s_cdc.data_ptr(), s_cdc.numel(), sp))
        ms = timed(f, args.steps)
        print(f"cdc    : {ms:8.3f} ms  {n/ms/1e6:8.1f} GB/s  ({int(d_n[0])} chunks)")
    ctx.close()


if __name__ == "__main__":
    main()
