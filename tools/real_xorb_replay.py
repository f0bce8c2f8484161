#!/usr/bin/env python
"""Replay of REAL xet-core xorbs through the library (the part that needs no Zig): hf_xet -- HuggingFace's
production Rust implementation, the one zest's peers and HF's CAS speak -- chunks, compresses (lz4_flex, byte-grouped) and
stores a synthetic bf16 tensor file as xorb files in a local store; every stored file (chunk stream + XETBLOB footer) is
then read back as it lies on disk and goes through zg_verify_xorbs (content address = the file name, chunk hashes = the
footer's) and zg_extract_chunk_range (bytes = the original file).  Reports parity and the device-resident / host-buffer
throughput on these compressor-made streams next to the liblz4-made ones bench.py uses.  Not part of bench.py's contract.
  python tools/real_xorb_replay.py --bytes 2000000000
"""
import argparse
import ctypes as C
import json
import os
import shutil
import struct
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import zest_b200 as z  # noqa: E402
from zest_b200 import _native  # noqa: E402
from zest_b200._native import check  # noqa: E402


def footer(raw):
    info_len = struct.unpack("<I", raw[-4:])[0]
    foot = raw[-4 - info_len:-4]
    assert foot[:7] == b"XETBLOB"
    p = 8 + 32 + 8
    n = struct.unpack("<I", foot[p:p + 4])[0]
    hashes = np.frombuffer(foot[p + 4:p + 4 + 32 * n], np.uint8).reshape(n, 32).copy()
    return len(raw) - 4 - info_len, bytes(foot[8:40]), hashes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=2_000_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--kind", default="bf16", choices=["bf16", "text"], help="bf16 -> policy bg4-lz4, text -> policy lz4")
    args = ap.parse_args()
    import hf_xet
    import torch
    from oracle import xet_oracle as oz
    dev = torch.device("cuda", 0)
    ctx = z.Context(0)
    L = _native.lib()
    n = args.bytes // 2 * 2
    piece = 256 << 20
    data = np.empty(n, np.uint8)
    for k in range((n + piece - 1) // piece):
        m = min(piece, n - k * piece)
        if args.kind == "bf16":
            oz.synth_bf16_fast(m, 424242 + k, out=data[k * piece:k * piece + m])
        else:
            data[k * piece:k * piece + m] = oz.synth_zipf_text(m, 424242 + k)
    d = tempfile.mkdtemp(prefix="xetreplay", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    os.environ["HF_XET_XORB_COMPRESSION_POLICY"] = "bg4-lz4" if args.kind == "bf16" else "lz4"
    try:
        t0 = time.time()
        s = hf_xet.XetSession()
        with s.new_upload_commit(endpoint="local://" + d) as c:
            h = c.start_upload_bytes(data.tobytes())
        file_hash = h.result().xet_info.hash
        t_up = time.time() - t0
        xdir = os.path.join(d, "xet", "xorbs", "xorbs")
        files = sorted(os.listdir(xdir), key=lambda f: os.path.getmtime(os.path.join(xdir, f)))
        raws = [(f.split(".", 1)[1], np.fromfile(os.path.join(xdir, f), np.uint8)) for f in files]
    finally:
        shutil.rmtree(d, ignore_errors=True)
    nx = len(raws)
    stored = sum(r.size for _, r in raws)
    # ---- parity: stored files as they are (footer included) through verify; names are the content addresses
    expected = np.stack([np.frombuffer(z.hex_to_hash(name), np.uint8) for name, _ in raws])
    pinned = []
    for _, r in raws:
        p = ctx.pinned_empty(r.size)
        p[:] = r
        pinned.append(p)
    res = z.verify_xorbs(pinned, expected, want_chunk_hashes=True, ctx=ctx)
    assert res.ok, res.verdict
    gold = np.concatenate([footer(r.tobytes())[2] for _, r in raws])
    assert np.array_equal(res.chunk_hashes, gold), "chunk hashes differ from the footers xet-core wrote"
    assert res.bytes_verified == n
    schemes = np.concatenate([z.xorb_index(p).scheme for p in pinned])
    # reconstruction: xorbs are consecutive pieces of the file here (one file, no dedup hits across its own chunks expected)
    pos = 0
    for p in pinned[:3] + pinned[-1:]:
        ix = z.xorb_index(p)
        out = z.XorbReader(p, ctx).extract_chunk_range(0, ix.n_chunks)
        if p is pinned[-1]:
            assert out == data[n - len(out):].tobytes()
        else:
            assert out == data[pos:pos + len(out)].tobytes()
            pos += len(out)
    # the oracle on the same files (its walk does not know the footer: hand it the stream part)
    streams = [r[:footer(r.tobytes())[0]] for _, r in raws[:4]]
    v, roots, _ = oz.verify_xorbs(streams, expected[:4], threads=os.cpu_count())
    assert (v == 1).all()
    # ---- timing
    t0 = time.perf_counter()
    for _ in range(args.steps):
        z.verify_xorbs(pinned, expected, ctx=ctx)
    t_e2e = (time.perf_counter() - t0) / args.steps
    # device-resident: same layout bench.py uses
    xs = bench.XorbSet()
    xs.index = list(range(nx))
    slen = [footer(r.tobytes())[0] for _, r in raws]
    xs.sizes = slen
    off, p0 = [], 0
    for sz in slen:
        off.append(p0)
        p0 += (sz + 15) // 16 * 16
    xs.offsets = off
    xs.buf = ctx.pinned_empty(p0 + 64)
    for o, sz, (_, r) in zip(off, slen, raws):
        xs.buf[o:o + sz] = r[:sz]
    counts = [z.xorb_index(v).n_chunks for v in xs.views]
    xs.first = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
    xs.ulens = np.concatenate([z.xorb_index(v).ulen for v in xs.views]).astype(np.uint32)
    descs, out_bytes = bench.build_descs(xs, L)
    nchunks = descs.size
    cpos = p0
    d_comp = torch.from_numpy(xs.buf[:cpos + 64]).to(dev)
    d_desc = torch.from_numpy(descs.view(np.uint8).reshape(-1)).to(dev)
    d_first = torch.from_numpy(xs.first.view(np.uint8).copy()).to(dev)
    d_exp = torch.from_numpy(expected.reshape(-1).copy()).to(dev)
    d_roots = torch.zeros(32 * nx, dtype=torch.uint8, device=dev)
    d_verdict = torch.zeros(nx, dtype=torch.int8, device=dev)
    d_hashes = torch.zeros(32 * nchunks, dtype=torch.uint8, device=dev)
    sb = L.zg_dev_verify_scratch_bytes(nchunks, nx, out_bytes)
    d_scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    d_out = torch.empty(out_bytes + 16, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(nchunks, dtype=torch.int32, device=dev)
    s_dec = torch.empty(L.zg_dev_decode_scratch_bytes(nchunks, out_bytes), dtype=torch.uint8, device=dev)
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def step():
        check(L.zg_dev_verify_xorbs(ctx.handle, d_comp.data_ptr(), cpos, d_desc.data_ptr(), nchunks, d_first.data_ptr(), nx, d_exp.data_ptr(),
                                    out_bytes, d_roots.data_ptr(), d_verdict.data_ptr(), d_hashes.data_ptr(), d_scratch.data_ptr(), sb, sp))

    def dec():
        check(L.zg_dev_decode_chunks(ctx.handle, d_comp.data_ptr(), cpos, d_desc.data_ptr(), nchunks, d_out.data_ptr(), out_bytes,
                                     d_status.data_ptr(), s_dec.data_ptr(), s_dec.numel(), sp))

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps / 1e3
    t_dev = timed(step)
    assert bool((d_verdict == 1).all()) and np.array_equal(d_hashes.cpu().numpy().reshape(-1, 32), gold)
    t_dec = timed(dec)
    print(json.dumps({
        "what": f"real xet-core xorbs (hf_xet local store, lz4_flex, {args.kind} data, policy {os.environ['HF_XET_XORB_COMPRESSION_POLICY']}) replayed through libzest_gpu",
        "hf_xet_upload_seconds": round(t_up, 1), "file_hash": file_hash, "bytes": n, "xorbs": nx, "chunks": int(nchunks),
        "stored_bytes": int(stored), "ratio": round(sum(slen) / n, 4),
        "scheme_histogram": {str(k): int((schemes == k).sum()) for k in (0, 1, 2)},
        "parity": "verdicts, roots (= file names) and every chunk hash (= footers) equal; reconstruction of 4 xorbs byte-exact; oracle agrees on 4",
        "device_resident_GBps": round(n / t_dev / 1e9, 2), "decode_alone_GBps": round(n / t_dec / 1e9, 2),
        "host_buffer_GBps": round(n / t_e2e / 1e9, 2)}))
    ctx.close()


if __name__ == "__main__":
    main()
