#!/usr/bin/env python
"""bench.py -- xorb verify GB/s (decompress + BLAKE3 + Merkle) on N B200s, next to the CPU path.

Contract (driver): python bench.py --gpus N --steps K --warmup W [--impl reference]; under torchrun one rank per
GPU.  Rank 0 prints ONE JSON line.

  step        one pass of the hot path over the workload's synthetic xorb set.  The set is ONE global list of xorbs;
              zest_b200.partition_xorbs (greedy LPT by compressed bytes) gives every rank a disjoint
              part, nothing but 32-byte roots and verdicts leaves a GPU -> no collective, "scaling": "strong".
  value       device-resident GB/s: the rank's compressed xorbs + chunk descriptors already in HBM, one
              zg_dev_verify_xorbs per step on torch's current stream, CUDA events, max over ranks.
              GB = 1e9 uncompressed bytes, whole job (all ranks).
  e2e         same metric through the reference-facing call (zg_verify_xorbs) with pinned HOST buffers: host walk
              of the chunk stream, H2D of the compressed bytes, kernels, D2H of roots + verdicts, all inside the
              timed region.  `e2e.two_callers`: the same calls issued by two host threads (how zest's task group
              would); `e2e.pageable`: the same call on ordinary malloc'ed blobs (what an unmodified zest allocator
              hands over).  The streaming workload replaces the single call by producer threads -> pinned ring ->
              consumer threads (see stream_e2e).
  roofline    the dominant kernel of the step (timed alone with CUDA events, same inputs, as the verify call runs it:
              zg_dev_decode_for_hash / zg_dev_hash_decoded / zg_dev_merkle), algorithmic bytes / time against the
              measured HBM copy bandwidth in MEASURED_PEAKS.json; `stages` lists every kernel; `blake3_alu` reports the
              hash kernel against the integer-ALU roof.
  cpu_baseline  the CPU oracle (oracle/, a port -- the reference's zig-xet code cannot be built here) on all host
              cores over a bounded sample of the same xorbs; rows C1-C6 of BASELINE.md section 3 where they apply.
  --impl reference   times only that CPU arm and prints it in the same shape.

Workloads (--workload), BASELINE.json configs[]:
  llama8b-bg4  (default) configs[2]: 16 GB of bf16 Gaussians, BG4-LZ4 chunks, ~239 xorbs of 64 MiB
  gpt2-lz4     configs[1]: ~550 MB Zipf text, LZ4 chunks
  llama70b-bg4 configs[4]: 140 GB as configs[2] over 8 GPUs (17.5 GB per rank), e2e streamed through a pinned ring
  hash-1gib    configs[0]: keyed chunk hashes of 1 GiB random data in its 16,741 CDC chunks (+ the bench.zig row)
  cdc-16g      configs[3]: Gearhash CDC + chunk hashes of one 16 GB bf16 shard, every boundary checked
  random-none  incompressible bytes, stored chunks
Every verify workload asserts, before timing, all verdicts and EVERY chunk hash of the device-resident pass
against the hashes the oracle computed when the xorbs were built, so a bench run is also a full-size parity test.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import queue
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

XORB_RAW = 64 * 1024 * 1024  # raw bytes per synthetic xorb (xet's xorb capacity)
WORKLOADS = {
    "llama8b-bg4": dict(mode="verify", kind="bf16", policy=2, nbytes=16_000_000_000,
                        desc="BASELINE.json configs[2]: Llama-3.1-8B-sized synthetic xorbs (16 GB bf16 Gaussians, BG4-LZ4 chunks)"),
    "gpt2-lz4": dict(mode="verify", kind="text", policy=1, nbytes=550_000_000,
                     desc="BASELINE.json configs[1]: gpt2-sized synthetic xorb set (~550 MB Zipf text, LZ4 chunks)"),
    "llama70b-bg4": dict(mode="verify", kind="bf16", policy=2, nbytes=140_000_000_000, per_rank_cap=17_500_000_000, stream=True,
                         desc="BASELINE.json configs[4]: Llama-3.1-70B-sized synthetic xorbs (140 GB bf16, BG4-LZ4) over 8 GPUs, "
                              "e2e streamed through a pinned ring"),
    "random-none": dict(mode="verify", kind="random", policy=0, nbytes=550_000_000, desc="incompressible bytes, stored chunks"),
    "hash-1gib": dict(mode="hash", nbytes=1 << 30,
                      desc="BASELINE.json configs[0]: keyed Xet chunk hashes of 1 GiB random data (numpy seed 0) in its 16,741 CDC chunks"),
    "cdc-16g": dict(mode="cdc", nbytes=16_000_000_000,
                    desc="BASELINE.json configs[3]: Gearhash CDC + chunk hashes of one 16 GB bf16 shard, bit-exact boundaries"),
}
INT_OPS_PER_COMPRESSION = 792  # 7 rounds x 8 G x 14 + 8
METRIC = "xorb verify GB/s (decompress+BLAKE3+Merkle)"
DESC_DTYPE = np.dtype([("src_off", "<u8"), ("dst_off", "<u8"), ("clen", "<u4"), ("ulen", "<u4"), ("scheme", "<u4"), ("order", "<u4")])


# ------------------------------------------------------------------------------------------------ workload
def plan_xorbs(nbytes: int):
    """Raw size of every xorb of the global set (the last one may be short)."""
    n = (nbytes + XORB_RAW - 1) // XORB_RAW
    return [min(XORB_RAW, nbytes - i * XORB_RAW) for i in range(n)]


class XorbSet:
    """The xorbs one rank owns: blobs back to back (16-byte aligned) in one pinned buffer + the oracle's answers."""

    def __init__(self):
        self.index = []       # global xorb numbers
        self.sizes = []       # compressed bytes
        self.offsets = []     # start inside `buf`
        self.expected = None  # (n, 32) u8
        self.hashes = None    # (n_chunks, 32) u8, oracle chunk hashes in order
        self.ulens = None     # u32 per chunk
        self.first = None     # u32[n+1] chunk index of each xorb's first chunk
        self.buf = None       # pinned uint8
        self.total = 0        # uncompressed bytes
        self.gen_seconds = 0.0

    @property
    def views(self):
        return [self.buf[o:o + s] for o, s in zip(self.offsets, self.sizes)]

    @property
    def comp_bytes(self):
        return int(sum(self.sizes))


def build_xorb_set(name: str, nbytes: int, seed: int, rank: int, world: int, threads: int, alloc, all_gather, log):
    """Generate the global set cooperatively and keep this rank's part (partition_xorbs by compressed bytes).
    Phase 1: rank r builds xorbs i = r (mod world) -- which gives every rank the compressed size of every xorb after
    one all-gather of sizes.  Phase 2: LPT partition; a rank rebuilds the xorbs it was assigned but did not build
    (each xorb is a pure function of (seed, i), so no blob crosses ranks)."""
    from oracle import xet_oracle as oz
    from zest_b200.verify import partition_xorbs
    w = WORKLOADS[name]
    raw = plan_xorbs(nbytes)
    n = len(raw)
    t0 = time.perf_counter()

    def make(i, out=None):
        return oz.synth_xorb(w["kind"], raw[i], seed * 1_000_003 + i, w["policy"], out)

    mine1 = list(range(rank, n, world))
    with ThreadPoolExecutor(max(1, threads)) as ex:
        held = dict(zip(mine1, ex.map(make, mine1)))
    sizes = np.zeros(n, np.int64)
    for i, r in held.items():
        sizes[i] = r[0].size
    sizes = all_gather(sizes)  # element-wise max over ranks == every rank's sizes
    parts = partition_xorbs(sizes.tolist(), world)
    mine = parts[rank]
    s = XorbSet()
    s.index = mine
    s.sizes = [int(sizes[i]) for i in mine]
    off, pos = [], 0
    for sz in s.sizes:
        off.append(pos)
        pos += (sz + 15) // 16 * 16
    s.offsets = off
    s.buf = alloc(pos + 64)
    need = [i for i in mine if i not in held]
    slot = {i: (o, sz) for i, o, sz in zip(mine, off, s.sizes)}
    for i in mine:
        if i in held:
            o, sz = slot[i]
            C.memmove(s.buf.ctypes.data + o, held[i][0].ctypes.data, sz)
    keep = {i: held[i][1:] for i in mine if i in held}
    del held

    def remake(i):
        o, sz = slot[i]
        r = make(i, s.buf[o:o + sz])
        assert r[0] == sz, "a xorb must rebuild to the same bytes"
        return r[1:]
    with ThreadPoolExecutor(max(1, threads)) as ex:
        keep.update(zip(need, ex.map(remake, need)))
    s.expected = np.frombuffer(b"".join(keep[i][0] for i in mine), np.uint8).reshape(-1, 32).copy() if mine else np.zeros((0, 32), np.uint8)
    s.hashes = np.concatenate([keep[i][1] for i in mine]) if mine else np.zeros((0, 32), np.uint8)
    s.ulens = np.concatenate([keep[i][2] for i in mine]) if mine else np.zeros(0, np.uint32)
    s.first = np.concatenate([[0], np.cumsum([keep[i][2].size for i in mine])]).astype(np.uint32)
    s.total = int(sum(raw[i] for i in mine))
    s.gen_seconds = time.perf_counter() - t0
    log(f"rank {rank}: {len(mine)} of {n} xorbs, {s.total/1e9:.2f} GB raw, {s.comp_bytes/1e9:.2f} GB compressed, "
        f"{len(need)} rebuilt after the partition, {s.gen_seconds:.1f} s on {threads} threads")
    return s, n, int(sizes.sum()), [int(sum(sizes[i] for i in p)) for p in parts]


def make_workload(name: str, nbytes: int, seed: int, threads: int):
    """The whole set on one rank as plain arrays -> (blobs, expected (n,32) u8, chunk hashes, uncompressed bytes, chunks):
    what the developer tools under tools/ take."""
    xs, _, _, _ = build_xorb_set(name, nbytes, seed, 0, 1, threads, lambda b: np.zeros(b, np.uint8), lambda a: a, lambda m: None)
    return xs.views, xs.expected, xs.hashes, xs.total, int(xs.first[-1])


def build_descs(xs: XorbSet, L):
    """Chunk descriptors of the rank's set from the host walk (zg_xorb_index), vectorised; decoded chunks 16-byte aligned."""
    import zest_b200 as z
    nchunks = int(xs.first[-1])
    d = np.zeros(nchunks, DESC_DTYPE)
    k = 0
    for o, v in zip(xs.offsets, xs.views):
        ix = z.xorb_index(v)
        m = ix.n_chunks
        d["src_off"][k:k + m] = o + ix.hdr_off.astype(np.uint64) + 8
        d["clen"][k:k + m] = ix.clen
        d["ulen"][k:k + m] = ix.ulen
        d["scheme"][k:k + m] = ix.scheme
        k += m
    assert k == nchunks and np.array_equal(d["ulen"], xs.ulens)
    al = (d["ulen"].astype(np.uint64) + 15) // 16 * 16
    d["dst_off"] = np.concatenate([[0], np.cumsum(al)[:-1]]).astype(np.uint64) if nchunks else 0
    out_bytes = int(al.sum())
    if nchunks:
        L.zg_chunk_order(d.ctypes.data, nchunks)
    return d, out_bytes


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # "under load": the samples of the busiest half (the region also holds host-side set-up between legs)
        load = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arms
def time_best(fn, min_seconds=2.0, max_reps=10):
    best, reps, t_all = None, 0, time.perf_counter()
    while reps < 2 or (time.perf_counter() - t_all < min_seconds and reps < max_reps):
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
        reps += 1
    return best, reps


def cpu_verify_rows(views, expected, ulen_per_xorb, cores, budget_bytes=3_000_000_000):
    """BASELINE.md row C2 on a bounded sample: full verify (walk + decode + hash + fold + compare) with the oracle
    port, one thread and all cores, plain decoder and liblz4's block decoder."""
    from oracle import xet_oracle as oz
    n, tot = 0, 0
    while n < len(views) and (n == 0 or tot + ulen_per_xorb[n] <= budget_bytes):
        tot += ulen_per_xorb[n]
        n += 1
    bl, ex = views[:n], expected[:n]
    rows = {}

    def run(threads, m):
        v, _, nb = oz.verify_xorbs(bl[:m], ex[:m], threads=threads)
        assert (v == 1).all()
        return nb
    for lz in (False, True):
        used = oz.set_lz4_decoder(lz)
        if lz and not used:
            continue
        tag = "port+liblz4" if lz else "port"
        m1 = max(1, min(n, 4))  # one thread: a few xorbs are enough
        b1 = sum(ulen_per_xorb[:m1])
        t, _ = time_best(lambda: run(1, m1), 1.0, 3)
        rows[f"C2 verify, {tag}, 1 thread"] = round(b1 / t / 1e9, 4)
        t, reps = time_best(lambda: run(cores, n), 2.0, 8)
        rows[f"C2 verify, {tag}, {cores} threads"] = round(tot / t / 1e9, 4)
    oz.set_lz4_decoder(False)
    best_key = max((k for k in rows if k.endswith(f"{cores} threads")), key=lambda k: rows[k])
    sample = f"first {n} xorbs of rank 0 ({tot/1e6:.0f} MB uncompressed), best of repeated passes"
    return rows[best_key], best_key, rows, sample


def _claim_stdout():
    """Keep stdout for the single JSON line: anything libraries print to fd 1 (NCCL's version banner, for one) goes
    to stderr instead.  Returns a writer for the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def log(msg):
    print(f"[bench] {msg}", file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------ main
def main():
    out_stream = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="llama8b-bg4", choices=sorted(WORKLOADS))
    ap.add_argument("--bytes", type=int, default=0, help="override the workload's total uncompressed size")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        if rank != 0:
            return 0
        return reference_arm(args, wl, out_stream)
    if wl["mode"] == "hash":
        return hash_bench(args, wl, out_stream)
    if wl["mode"] == "cdc":
        return cdc_bench(args, wl, out_stream)
    return verify_bench(args, wl, out_stream)


def workload_bytes(args, wl, world):
    nbytes = args.bytes or wl["nbytes"]
    if not args.bytes and wl.get("per_rank_cap"):
        nbytes = min(nbytes, wl["per_rank_cap"] * world)  # the 8-GPU set; fewer GPUs keep the per-GPU share
    return int(nbytes)


def base_config(args, wl, world, nbytes):
    return {"workload": wl["desc"], "name": args.workload, "uncompressed_bytes_total": nbytes,
            "l2_policy": "inputs larger than L2 (compressed + decoded working set >> 126 MB), no flush needed",
            "parallelism": f"one global xorb list, LPT-partitioned by compressed bytes over {world} rank(s), no collective",
            "chunk_index": "host walk of the 8-byte chunk headers (zg_xorb_index): inside the timed region of e2e; "
                           "value starts from blobs + 32-byte chunk descriptors in HBM"}


def reference_arm(args, wl, out_stream):
    """--impl reference: the CPU implementation of the path (the oracle port; zig-xet cannot be built here) on all host
    cores, whole workload per step, same config / metric / unit as our arm."""
    from oracle import xet_oracle as oz
    K, W = args.steps, args.warmup
    cores = os.cpu_count() or 1
    world = max(1, args.gpus)
    nbytes = workload_bytes(args, wl, world)
    config = base_config(args, wl, world, nbytes)
    if wl["mode"] == "verify":
        xs, n_all, comp_all, _ = build_xorb_set(args.workload, nbytes, 1, 0, 1, cores, lambda b: np.empty(b, np.uint8), lambda a: a, log)
        views, exp = xs.views, xs.expected
        lz = oz.set_lz4_decoder(True)
        # a step = one verify pass over a bounded sample (the whole set when it is small enough to keep the run short)
        ul = np.add.reduceat(xs.ulens.astype(np.int64), xs.first[:-1].astype(np.int64)).tolist()
        budget, m, tot = 16_000_000_000, 0, 0
        while m < len(views) and (m == 0 or tot + ul[m] <= budget):
            tot += ul[m]; m += 1

        def step():
            v, _, nb = oz.verify_xorbs(views[:m], exp[:m], threads=cores)
            assert (v == 1).all() and nb == tot
        sample = f"{m} of {n_all} xorbs ({tot/1e6:.0f} MB uncompressed) per step, {'liblz4 block decoder' if lz else 'plain decoder'}"
        work = tot
        config = dict(config, n_xorbs=n_all, compressed_bytes_total=comp_all, compression_ratio=round(comp_all / nbytes, 4))
    elif wl["mode"] == "hash":
        data = oz.synth_random(nbytes, 0)
        off = np.concatenate([np.zeros(1, np.uint64), oz.cdc(data)])

        def step():
            oz.hash_chunks(data, off, cores)
        sample, work = f"all {off.size-1} chunks ({nbytes/1e6:.0f} MB) per step", nbytes
    else:
        m = min(nbytes, 2_000_000_000)
        data = oz.synth_bf16_fast(m, 11)

        def step():
            e = oz.cdc(data)
            oz.hash_chunks(data, np.concatenate([np.zeros(1, np.uint64), e]), cores)
        sample, work = f"first {m/1e6:.0f} MB of the shard per step (CDC serial, hashing on all cores)", m
    for _ in range(W):
        step()
    t0 = time.perf_counter()
    for _ in range(K):
        step()
    dt = time.perf_counter() - t0
    val = work * K / dt / 1e9
    metric = METRIC if wl["mode"] == "verify" else ("chunk hash GB/s (BLAKE3 keyed)" if wl["mode"] == "hash" else "CDC + chunk hash GB/s")
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": round(val, 4), "unit": "GB/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
        "ms_per_step": round(dt / K * 1e3, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic", "config": config,
        "cpu_baseline": {"value": round(val, 4), "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample,
                         "note": "oracle/xet_oracle.c (CPU restatement; zest's zig-xet path needs a Zig toolchain, absent here)"},
        "e2e": {"value": round(val, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), file=out_stream, flush=True)
    return 0


def gpu_setup():
    import torch
    import zest_b200 as z
    from zest_b200 import _native, distributed as zd
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: zest_b200 has no CPU fallback")
    rank, world, local = zd.init_from_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ctx = z.Context(local)
    return torch, z, _native, zd, rank, world, local, dev, ctx


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return {}


def load_traffic(workload):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(workload, {})
    except (OSError, ValueError):
        return {}


def verify_bench(args, wl, out_stream):
    torch, z, _native, zd, rank, world, local, dev, ctx = gpu_setup()
    import torch.distributed as dist
    from zest_b200._native import check
    L = _native.lib()
    K, W = args.steps, max(args.warmup, 3)
    cores = os.cpu_count() or 1
    host_threads = max(1, cores // max(world, 1))
    nbytes = workload_bytes(args, wl, world)
    config = base_config(args, wl, world, nbytes)

    def all_gather_max(a):
        if world == 1:
            return a
        t = torch.from_numpy(a).to(dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().numpy()

    xs, n_all, comp_all, part_bytes = build_xorb_set(args.workload, nbytes, 1, rank, world, host_threads, ctx.pinned_empty, all_gather_max, log)
    nx, total, comp_bytes = len(xs.index), xs.total, xs.comp_bytes
    views, expected = xs.views, xs.expected
    descs, out_bytes = build_descs(xs, L)
    nchunks = descs.size
    cpos = (xs.offsets[-1] + (xs.sizes[-1] + 15) // 16 * 16) if nx else 0

    # ---- device-resident inputs
    d_comp = torch.empty(cpos + 64, dtype=torch.uint8, device=dev)
    pin_t = torch.from_numpy(xs.buf[:cpos])
    d_comp[:cpos].copy_(pin_t, non_blocking=True)
    d_desc = torch.from_numpy(descs.view(np.uint8).reshape(-1)).to(dev)
    d_first = torch.from_numpy(xs.first.view(np.uint8).copy()).to(dev)
    d_exp = torch.from_numpy(expected.reshape(-1).copy()).to(dev)
    d_roots = torch.zeros(32 * max(nx, 1), dtype=torch.uint8, device=dev)
    d_verdict = torch.zeros(max(nx, 1), dtype=torch.int8, device=dev)
    d_hashes = torch.zeros(32 * max(nchunks, 1), dtype=torch.uint8, device=dev)
    sbytes = L.zg_dev_verify_scratch_bytes(nchunks, nx, out_bytes)
    d_scratch = torch.empty(sbytes, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)

    def step_dev():
        check(L.zg_dev_verify_xorbs(ctx.handle, d_comp.data_ptr(), cpos, d_desc.data_ptr(), nchunks, d_first.data_ptr(), nx,
                                    d_exp.data_ptr(), out_bytes, d_roots.data_ptr(), d_verdict.data_ptr(), d_hashes.data_ptr(),
                                    d_scratch.data_ptr(), sbytes, sp))

    def timed(fn, steps):
        """CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks -> seconds."""
        torch.cuda.synchronize(); zd.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(); zd.barrier()
        return zd.reduce_max(e0.elapsed_time(e1) / 1e3)

    for _ in range(W):
        step_dev()
    torch.cuda.synchronize()
    assert bool((d_verdict[:nx] == 1).all()), "device-resident verify: verdicts"
    assert np.array_equal(d_hashes.cpu().numpy().reshape(-1, 32)[:nchunks], xs.hashes), "device-resident verify: chunk hashes"
    assert np.array_equal(d_roots.cpu().numpy().reshape(-1, 32)[:nx], expected), "device-resident verify: roots"

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    k0 = ctx.stats()["kernels_launched"]
    t_dev = timed(step_dev, K)
    launches = ctx.stats()["kernels_launched"] - k0
    all_bytes = zd.reduce_sum(total)
    value = all_bytes * K / t_dev / 1e9

    # ---- per-kernel timing for the roofline (rank 0's numbers are reported; same inputs, each stage alone).  The
    # stage buffers replace the one-call scratch, so the two never sit in HBM together.
    del d_scratch
    torch.cuda.empty_cache()
    d_off = torch.from_numpy(descs["dst_off"].copy().view(np.uint8)).to(dev)
    d_len = torch.from_numpy(descs["ulen"].copy().view(np.uint8)).to(dev)
    d_out = torch.empty(out_bytes + 16, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(max(nchunks, 1), dtype=torch.int32, device=dev)
    s_dec = torch.empty(L.zg_dev_decode_scratch_bytes(nchunks, out_bytes), dtype=torch.uint8, device=dev)
    s_hash = torch.empty(L.zg_dev_hash_decoded_scratch_bytes(out_bytes, nchunks), dtype=torch.uint8, device=dev)
    s_mer = torch.empty(L.zg_dev_merkle_scratch_bytes(nchunks, nx), dtype=torch.uint8, device=dev)

    # the stages as zg_dev_verify_xorbs runs them: byte-grouped chunks stay as planes, the hash regroups them as it reads
    def k_decode():
        check(L.zg_dev_decode_for_hash(ctx.handle, d_comp.data_ptr(), cpos, d_desc.data_ptr(), nchunks, d_out.data_ptr(), out_bytes,
                                       d_status.data_ptr(), s_dec.data_ptr(), s_dec.numel(), sp))

    def k_hash():
        check(L.zg_dev_hash_decoded(ctx.handle, d_out.data_ptr(), out_bytes, d_desc.data_ptr(), nchunks, s_dec.data_ptr(),
                                    d_hashes.data_ptr(), s_hash.data_ptr(), s_hash.numel(), sp))

    def k_merkle():
        check(L.zg_dev_merkle(ctx.handle, d_hashes.data_ptr(), d_len.data_ptr(), d_first.data_ptr(), nx, nchunks, d_status.data_ptr(),
                              d_exp.data_ptr(), d_roots.data_ptr(), d_verdict.data_ptr(), s_mer.data_ptr(), s_mer.numel(), sp))

    stages = {}
    ks = max(3, min(K, 5))
    d_hashes.zero_()
    for name, fn in (("lz4_decode", k_decode), ("blake3_hash", k_hash), ("merkle_fold", k_merkle)):
        for _ in range(2):
            fn()
        stages[name] = timed(fn, ks) / ks
    assert int(d_status.abs().sum()) == 0 and bool((d_verdict[:nx] == 1).all())
    assert np.array_equal(d_hashes.cpu().numpy().reshape(-1, 32)[:nchunks], xs.hashes), "stage-wise verify: chunk hashes"
    # what the link alone does: the same compressed bytes, pinned host -> device, nothing else (bounds e2e from above)
    for _ in range(2):
        d_comp[:cpos].copy_(pin_t, non_blocking=True)
    t_link = timed(lambda: d_comp[:cpos].copy_(pin_t, non_blocking=True), 3) / 3
    link_gbps = zd.reduce_sum(cpos) / t_link / 1e9  # aggregate over ranks
    del d_out, s_dec, s_hash, s_mer
    torch.cuda.empty_cache()

    # ---- e2e: reference-facing call with pinned host buffers
    s0 = ctx.stats()
    extra = {}
    if wl.get("stream"):
        t_e2e, extra, s0, s_end = stream_e2e(z, ctx, xs, K, W, zd, torch, host_threads)
    else:
        def step_e2e():
            return z.verify_xorbs(views, expected, ctx=ctx)
        for _ in range(W):
            r0 = step_e2e()
        assert r0.ok and r0.bytes_verified == total
        s0 = ctx.stats()
        torch.cuda.synchronize(); zd.barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            step_e2e()
        torch.cuda.synchronize(); zd.barrier()
        t_e2e = zd.reduce_max(time.perf_counter() - t0)
    s1 = s_end if wl.get("stream") else ctx.stats()
    e2e_val = all_bytes * K / t_e2e / 1e9
    if not wl.get("stream"):
        # The same K calls issued by two host threads (zest runs up to 128 download tasks at once,
        # src/parallel_download.zig:117): the tail of one call then overlaps the copies of the next.
        def two_caller_pass(steps):
            nxt, lock = [0], threading.Lock()

            def caller():
                while True:
                    with lock:
                        if nxt[0] >= steps:
                            return
                        nxt[0] += 1
                    step_e2e()
            torch.cuda.synchronize(); zd.barrier()
            t0 = time.perf_counter()
            ths = [threading.Thread(target=caller) for _ in range(2)]
            for th in ths:
                th.start()
            for th in ths:
                th.join()
            torch.cuda.synchronize(); zd.barrier()
            return zd.reduce_max(time.perf_counter() - t0)
        k2 = max(2, min(K, 6))
        t2 = sorted(two_caller_pass(k2) for _ in range(3))[1]  # median of three: host-thread scheduling makes it noisy
        extra["two_callers"] = {"value": round(all_bytes * k2 / t2 / 1e9, 3), "ms_per_step": round(t2 / k2 * 1e3, 3)}
        # Pageable blobs: what zest's allocator-owned []u8 (src/xet_bridge.zig:45-52) gives the call when the caller does
        # not adopt zg_pinned_alloc.  A bounded sample of the set (ordinary numpy memory), same call.
        m = max(1, min(nx, 24))
        pg = [np.array(v) for v in views[:m]]
        pg_bytes = int(sum(xs.ulens[int(xs.first[0]):int(xs.first[m])].astype(np.int64)))
        z.verify_xorbs(pg, expected[:m], ctx=ctx)
        torch.cuda.synchronize(); zd.barrier()
        t0 = time.perf_counter()
        for _ in range(2):
            rp = z.verify_xorbs(pg, expected[:m], ctx=ctx)
        torch.cuda.synchronize(); zd.barrier()
        t_pg = zd.reduce_max(time.perf_counter() - t0) / 2
        assert rp.ok
        extra["pageable"] = {"value": round(zd.reduce_sum(pg_bytes) / t_pg / 1e9, 3),
                             "sample": f"first {m} xorbs of every rank in ordinary (pageable) host memory, 2 passes"}
        del pg
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        ctx.close()
        zd.shutdown()
        return 0

    # ---- roofline of the dominant kernel (rank 0's launch: its own bytes and times)
    peaks = load_peaks()
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    alg = {"lz4_decode": comp_bytes + total, "blake3_hash": total + 32 * nchunks, "merkle_fold": 40 * nchunks + 33 * nx}
    dom = max(stages, key=lambda s: stages[s])
    achieved = alg[dom] / stages[dom] / 1e9
    tr = load_traffic(args.workload)
    traffic = None
    if tr.get(dom) and tr.get("uncompressed_bytes"):
        traffic = int(tr[dom] * (total / tr["uncompressed_bytes"]))  # ncu capture of a smaller launch, scaled by bytes
    elif tr.get(dom):
        traffic = tr[dom]
    sm_mhz = (clocks or {}).get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
    lz4_issue = None
    if tr.get("lz4_decode_warp_instructions"):
        n_inst = tr["lz4_decode_warp_instructions"] * (total / tr["uncompressed_bytes"] if tr.get("uncompressed_bytes") else 1.0)
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        floor_ms = n_inst / (4.0 * n_sm * sm_mhz * 1e6) * 1e3
        lz4_issue = {"warp_instructions": int(n_inst), "issue_floor_ms": round(floor_ms, 4),
                     "frac_of_issue_floor": round(floor_ms / (stages["lz4_decode"] * 1e3), 4),
                     "note": "smsp__inst_executed.sum of one launch (ncu, profiles/) scaled to this launch's bytes / (4 schedulers x SMs x SM clock)"}
    from oracle import xet_oracle as oz
    lens_u, cnt_u = np.unique(xs.ulens, return_counts=True)
    compressions = int(sum(oz.blake3_compressions(int(n)) * int(c) for n, c in zip(lens_u, cnt_u)))
    int_ops = compressions * INT_OPS_PER_COMPRESSION
    sink = torch.zeros(4, dtype=torch.uint8, device=dev)
    g_alu, g_fma, g_both = C.c_double(), C.c_double(), C.c_double()
    check(L.zg_dev_int_pipe_peak(sink.data_ptr(), C.byref(g_alu), C.byref(g_fma), C.byref(g_both), sp))
    t_floor = compressions * max(448 / (g_alu.value * 1e9), 336 / (g_fma.value * 1e9), 784 / (g_both.value * 1e9))
    t_step_rank0 = sum(stages.values())
    roofline = {"kernel": dom, "bound": "hbm", "achieved": round(achieved, 2), "peak": hbm_peak, "unit": "GB/s",
                "frac": round(achieved / hbm_peak, 4), "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(alg[dom]), "avg_launch_ms": round(stages[dom] * 1e3, 4),
                "stages": {s: {"ms": round(t * 1e3, 4), "alg_GBps": round(alg[s] / t / 1e9, 2),
                               "uncompressed_GBps": round(total / t / 1e9, 2)} for s, t in stages.items()},
                "lz4_issue": lz4_issue,
                "blake3_alu": {"compressions": compressions, "int_ops": int(int_ops),
                               "achieved_Tops": round(int_ops / stages["blake3_hash"] / 1e12, 3),
                               "whole_step_Tops": round(int_ops / t_step_rank0 / 1e12, 3),
                               "measured_pipe_peaks_Tops": {"alu": round(g_alu.value / 1e3, 2), "fma_imad": round(g_fma.value / 1e3, 2),
                                                            "both": round(g_both.value / 1e3, 2)},
                               "frac_of_alu_pipe": round(int_ops / stages["blake3_hash"] / 1e9 / g_alu.value, 4),
                               "whole_step_frac_of_alu_pipe": round(int_ops / t_step_rank0 / 1e9 / g_alu.value, 4),
                               "frac_of_issue_floor": round(t_floor / stages["blake3_hash"], 4),
                               "note": "int_ops = 792 per compression (SURVEY.md 8d); the kernel issues 448 ALU-pipe + 336 "
                                       "FMA-pipe (IMAD adds) thread-instructions per compression; issue floor = "
                                       "max(448/alu, 336/fma, 784/both) from the measured pipe rates",
                               "sm_mhz": sm_mhz}}
    cpu = None
    if not args.no_cpu:
        ul = np.add.reduceat(xs.ulens.astype(np.int64), xs.first[:-1].astype(np.int64)).tolist() if nx else []
        v, which, rows, sample = cpu_verify_rows(views, expected, ul, cores)
        cpu = {"value": v, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample, "which": which, "rows": rows}
    out = {
        "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(t_dev / K * 1e3, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(config, n_xorbs=n_all, compressed_bytes_total=comp_all, compression_ratio=round(comp_all / nbytes, 4)),
        "run": {"rank0": {"xorbs": nx, "chunks": int(nchunks), "uncompressed_bytes": total, "compressed_bytes": comp_bytes},
                "partition_compressed_bytes": part_bytes, "generate_seconds": round(xs.gen_seconds, 1)},
        "e2e": dict({"value": round(e2e_val, 3), "unit": "GB/s", "ms_per_step": round(t_e2e / K * 1e3, 3),
                     "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) // K),
                     "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) // K),
                     "h2d_GBps_rank0": round((s1["h2d_bytes"] - s0["h2d_bytes"]) / t_e2e / 1e9, 2),
                     "h2d_link_GBps": round(link_gbps, 2),
                     "link_bound_GBps_uncompressed": round(link_gbps * nbytes / comp_all, 2),
                     "frac_of_link_bound": round(e2e_val / (link_gbps * nbytes / comp_all), 4)}, **extra),
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(out), file=out_stream, flush=True)
    ctx.close()
    zd.shutdown()
    return 0


def stream_e2e(z, ctx, xs, K, W, zd, torch, host_threads):
    """configs[4]: the rank's xorbs go through a bounded ring of pinned slots instead of one call over the whole set.
    Producer threads stand where zest's network / disk readers stand (src/parallel_download.zig:117,174-182): each takes
    the next xorb, waits for a free slot, makes the blob appear in it and queues it; two consumer threads take batches
    of ready slots through zg_verify_xorbs and hand the slots back, so at most `ring_slots` xorbs are in flight.
    Two kinds of producer:
      in place  the bytes land in the pinned slot without the CPU touching them, as a NIC's or an NVMe's DMA would
                deliver them (the slot IS the xorb's place in the pinned set) -- the timed figure: ring, batching,
                consumer calls, H2D link;
      copy      a CPU memcpy into a slot of a separate pinned ring (zg_pinned_alloc) stands in for the network -- reported
                beside it with the copy rates: on a box with few cores per GPU the memcpy, not the GPU side, is the limit.
    Checked every pass: all verdicts 1 and every root equal to the oracle's; once: a planted corruption is caught.
    -> (seconds for K timed in-place passes, max over ranks; extras; stats before / after the timed region)."""
    nx = len(xs.index)
    slot_bytes = max(xs.sizes) + 64
    n_slots = min(nx, 48)
    batch = 6
    ring = ctx.pinned_empty(slot_bytes * n_slots)
    views = xs.views
    n_prod = max(2, min(8, host_threads - 2))
    stats = {"wait_slot_s": 0.0, "wait_ready_s": 0.0, "copy_s": 0.0}

    def one_pass(copy, corrupt=None):
        free_q, ready_q = queue.Queue(), queue.Queue()
        for s in range(n_slots):
            free_q.put(s)
        nxt, lock = [0], threading.Lock()
        verdict = np.zeros(nx, np.int8)
        roots = np.zeros((nx, 32), np.uint8)
        err = []

        def producer():
            while True:
                with lock:
                    i = nxt[0]
                    nxt[0] += 1
                if i >= nx:
                    return
                t0 = time.perf_counter()
                s = free_q.get()
                t1 = time.perf_counter()
                if copy:
                    C.memmove(ring.ctypes.data + s * slot_bytes, views[i].ctypes.data, xs.sizes[i])
                    blob = ring[s * slot_bytes:s * slot_bytes + xs.sizes[i]]
                    if corrupt is not None and i == corrupt:
                        blob[xs.sizes[i] // 2] ^= 0x5A
                else:
                    blob = views[i]
                t2 = time.perf_counter()
                stats["wait_slot_s"] += t1 - t0
                stats["copy_s"] += t2 - t1
                ready_q.put((i, s, blob))

        def consumer():
            try:
                while True:
                    t0 = time.perf_counter()
                    item = ready_q.get()
                    stats["wait_ready_s"] += time.perf_counter() - t0
                    if item is None:
                        ready_q.put(None)
                        return
                    items = [item]
                    while len(items) < batch:
                        try:
                            it = ready_q.get_nowait()
                        except queue.Empty:
                            break
                        if it is None:
                            ready_q.put(None)
                            break
                        items.append(it)
                    ids = [i for i, _, _ in items]
                    r = z.verify_xorbs([b for _, _, b in items], xs.expected[ids], ctx=ctx)
                    verdict[ids] = r.verdict
                    roots[ids] = r.roots
                    for _, s, _ in items:
                        free_q.put(s)
            except Exception as e:  # surfaced by the caller
                err.append(e)
                ready_q.put(None)
        prods = [threading.Thread(target=producer) for _ in range(n_prod)]
        cons = [threading.Thread(target=consumer) for _ in range(2)]
        for t in prods + cons:
            t.start()
        for t in prods:
            t.join()
        ready_q.put(None)
        for t in cons:
            t.join()
        if err:
            raise err[0]
        if corrupt is None:
            assert (verdict == 1).all(), "streamed verify: verdicts"
            assert np.array_equal(roots, xs.expected), "streamed verify: roots"
        return verdict

    def timed_passes(copy, n):
        for k in stats:
            stats[k] = 0.0
        sa = ctx.stats()
        torch.cuda.synchronize(); zd.barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            one_pass(copy)
        torch.cuda.synchronize(); zd.barrier()
        t = zd.reduce_max(time.perf_counter() - t0)
        return t, sa, ctx.stats(), dict(stats)

    for _ in range(max(1, W - 1)):
        one_pass(False)
    bad = nx // 2
    v = one_pass(True, corrupt=bad)
    assert v[bad] != 1 and (np.delete(v, bad) == 1).all(), "a planted corruption must fail exactly its xorb"
    n_copy = max(1, min(K, 2))
    t_copy, _, _, st_c = timed_passes(True, n_copy)
    t, s0, s1, st = timed_passes(False, K)
    comp_all = zd.reduce_sum(float(sum(xs.sizes)))
    unc_all = zd.reduce_sum(float(xs.total))
    extra = {"stream": {"ring_slots": n_slots, "slot_bytes": int(slot_bytes), "batch_xorbs": batch, "producers": n_prod, "consumers": 2,
                        "timed": "in-place producers (bytes land in the pinned slot by DMA, the CPU only queues the slot)",
                        "consumers_waiting_for_data_frac": round(st["wait_ready_s"] / (2 * t), 3),
                        "producers_waiting_for_slot_frac": round(st["wait_slot_s"] / (n_prod * t), 3),
                        "copy_producers": {
                            "value": round(unc_all * n_copy / t_copy / 1e9, 3), "unit": "GB/s", "passes": n_copy,
                            "producer_copy_GBps_per_thread_rank0": round(float(sum(xs.sizes)) * n_copy / max(st_c["copy_s"], 1e-9) * 1e-9, 2),
                            "producer_threads_busy_frac": round(st_c["copy_s"] / (n_prod * t_copy), 3),
                            "consumers_waiting_for_data_frac": round(st_c["wait_ready_s"] / (2 * t_copy), 3),
                            "host_memcpy_GBps_all_ranks": round(comp_all * n_copy / t_copy / 1e9, 2),
                            "note": "a CPU memcpy per blob stands in for the network: with few host cores per GPU it, not the "
                                    "GPU side, bounds this figure"},
                        "planted_corruption": f"xorb {bad} of every rank: verdict {int(v[bad])}, all others 1"}}
    ctx.pinned_free(ring)
    return t, extra, s0, s1


def hash_bench(args, wl, out_stream):
    """configs[0]: what `zest bench --synthetic` would time if it hashed Xet chunks: keyed BLAKE3 of every CDC chunk of
    1 GiB of random bytes.  value = device-resident zg_dev_hash_chunks; e2e = zg_hash_chunks from pinned host memory.
    Side row: the bench.zig input itself (plain BLAKE3 of 64 KiB of 0x42, 1000 of them) next to the published figure."""
    torch, z, _native, zd, rank, world, local, dev, ctx = gpu_setup()
    from zest_b200._native import check
    from oracle import xet_oracle as oz
    L = _native.lib()
    K, W = args.steps, max(args.warmup, 3)
    cores = os.cpu_count() or 1
    nbytes = args.bytes or wl["nbytes"]
    data = oz.synth_random(nbytes, 0 + rank)
    ends = oz.cdc(data)
    off = np.concatenate([np.zeros(1, np.uint64), ends])
    n = ends.size
    gold = oz.hash_chunks(data, off, cores)
    pin = ctx.pinned_empty(nbytes + 16)
    pin[:nbytes] = data
    d_data = torch.from_numpy(pin).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_out = torch.zeros(32 * n, dtype=torch.uint8, device=dev)
    scratch = torch.empty(L.zg_dev_hash_scratch_bytes(nbytes, n), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)

    def step():
        check(L.zg_dev_hash_chunks(ctx.handle, d_data.data_ptr(), nbytes, d_off.data_ptr(), None, n, None, d_out.data_ptr(),
                                   scratch.data_ptr(), scratch.numel(), sp))

    def timed(fn, steps):
        torch.cuda.synchronize(); zd.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(); zd.barrier()
        return zd.reduce_max(e0.elapsed_time(e1) / 1e3)
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy().reshape(-1, 32), gold), "chunk hashes"
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start(); time.sleep(0.3)
    k0 = ctx.stats()["kernels_launched"]
    t_dev = timed(step, K)
    launches = ctx.stats()["kernels_launched"] - k0
    all_bytes = zd.reduce_sum(nbytes)
    value = all_bytes * K / t_dev / 1e9

    def step_e2e():
        return z.hash_chunks(pin[:nbytes], off, ctx=ctx)
    for _ in range(W):
        h = step_e2e()
    assert np.array_equal(np.asarray(h).reshape(-1, 32), gold)
    s0 = ctx.stats()
    torch.cuda.synchronize(); zd.barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        step_e2e()
    torch.cuda.synchronize(); zd.barrier()
    t_e2e = zd.reduce_max(time.perf_counter() - t0)
    s1 = ctx.stats()
    # side row: bench.zig's input, 1000 x 64 KiB of 0x42, plain BLAKE3 (src/bench.zig:207-223, published 3,517 "MB/s" = MiB/s)
    zig_n, zig_len = 1000, 65536
    zd_data = torch.full((zig_n * zig_len + 16,), 0x42, dtype=torch.uint8, device=dev)
    zd_off = torch.arange(0, (zig_n + 1) * zig_len, zig_len, dtype=torch.int64, device=dev)
    zd_out = torch.zeros(32 * zig_n, dtype=torch.uint8, device=dev)
    zs = torch.empty(L.zg_dev_hash_scratch_bytes(zig_n * zig_len, zig_n), dtype=torch.uint8, device=dev)
    plain = C.c_void_p(1)

    def step_zig():
        check(L.zg_dev_hash_chunks(ctx.handle, zd_data.data_ptr(), zig_n * zig_len, zd_off.data_ptr(), None, zig_n, plain,
                                   zd_out.data_ptr(), zs.data_ptr(), zs.numel(), sp))
    for _ in range(3):
        step_zig()
    t_zig = timed(step_zig, 20) / 20
    want = oz.blake3(b"\x42" * zig_len)
    assert bytes(zd_out[:32].cpu().numpy()) == want and bytes(zd_out[-32:].cpu().numpy()) == want
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        ctx.close(); zd.shutdown()
        return 0
    peaks = load_peaks()
    lens = np.diff(off).astype(np.int64)
    lens_u, cnt_u = np.unique(lens, return_counts=True)
    compressions = int(sum(oz.blake3_compressions(int(a)) * int(c) for a, c in zip(lens_u, cnt_u)))
    int_ops = compressions * INT_OPS_PER_COMPRESSION
    sink = torch.zeros(4, dtype=torch.uint8, device=dev)
    g_alu, g_fma, g_both = C.c_double(), C.c_double(), C.c_double()
    check(L.zg_dev_int_pipe_peak(sink.data_ptr(), C.byref(g_alu), C.byref(g_fma), C.byref(g_both), sp))
    t_one = t_dev / K
    t_floor = compressions * max(448 / (g_alu.value * 1e9), 336 / (g_fma.value * 1e9), 784 / (g_both.value * 1e9))
    roofline = {"kernel": "blake3_hash", "bound": "alu", "achieved": round(int_ops / t_one / 1e12, 3), "peak": round(g_alu.value / 1e3, 3),
                "unit": "Tops/s", "frac": round(int_ops / t_one / 1e9 / g_alu.value, 4), "traffic": load_traffic(args.workload).get("blake3_hash"),
                "peak_source": "measured on the spot (zg_dev_int_pipe_peak: ALU-pipe thread-ops/s; FMA pipe and both pipes listed beside it)",
                "measured_pipe_peaks_Tops": {"alu": round(g_alu.value / 1e3, 2), "fma_imad": round(g_fma.value / 1e3, 2), "both": round(g_both.value / 1e3, 2)},
                "frac_of_issue_floor": round(t_floor / t_one, 4), "compressions": compressions, "int_ops_per_compression": INT_OPS_PER_COMPRESSION,
                "hbm": {"achieved_GBps": round((nbytes + 32 * n) / t_one / 1e9, 2), "peak_GBps": float(peaks.get("hbm_gbs", 6650.0))},
                "avg_launch_ms": round(t_one * 1e3, 4)}
    cpu = None
    if not args.no_cpu:
        rows = {}
        t, _ = time_best(lambda: oz.hash_chunks(data, off, 1), 1.0, 3)
        rows["C1 keyed chunk hash, port, 1 thread"] = round(nbytes / t / 1e9, 4)
        t, _ = time_best(lambda: oz.hash_chunks(data, off, cores), 1.0, 5)
        rows[f"C1 keyed chunk hash, port, {cores} threads"] = round(nbytes / t / 1e9, 4)
        best = rows[f"C1 keyed chunk hash, port, {cores} threads"]
        try:
            import blake3 as pyb3
            key = bytes([102, 151, 245, 119, 91, 149, 80, 222, 49, 53, 203, 172, 165, 151, 24, 28, 157, 228, 33, 16, 155, 235, 43, 88,
                         180, 208, 176, 75, 147, 173, 242, 41])
            mv = memoryview(data[:256 << 20])

            def c5():
                for i in range(0, len(mv), 65536):
                    pyb3.blake3(mv[i:i + 65536], key=key).digest()
            t, _ = time_best(c5, 1.0, 3)
            rows["C5 PyPI blake3 keyed, 64 KiB slices of 256 MiB, 1 thread"] = round(len(mv) / t / 1e9, 4)
            buf = b"\x42" * 65536

            def c6():
                for _ in range(1000):
                    pyb3.blake3(buf).digest()
            t, _ = time_best(c6, 0.5, 5)
            rows["C6 bench.zig-equivalent, PyPI blake3 plain 64 KiB x1000, 1 thread (MiB/s)"] = round(1000 * 65536 / t / 2 ** 20, 1)
        except ImportError:
            rows["C5/C6"] = "PyPI blake3 not importable on this box"
        buf = np.full(65536, 0x42, np.uint8)
        t, _ = time_best(lambda: [oz.blake3(buf) for _ in range(1000)], 0.5, 5)
        rows["C6 bench.zig-equivalent, oracle port plain 64 KiB x1000, 1 thread (MiB/s)"] = round(1000 * 65536 / t / 2 ** 20, 1)
        cpu = {"value": best, "unit": "GB/s", "cores": cores, "kind": "port", "sample": f"all {n} chunks ({nbytes/1e6:.0f} MB), best of repeated passes",
               "rows": rows}
    out = {"metric": "chunk hash GB/s (BLAKE3 keyed, Xet data key)", "value": round(value, 3), "unit": "GB/s", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": round(t_dev / K * 1e3, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
           "data": "synthetic",
           "config": {"workload": wl["desc"], "name": args.workload, "bytes_per_rank": nbytes, "chunks_per_rank": int(n),
                      "l2_policy": "input (1 GiB) larger than L2, no flush needed"},
           "e2e": {"value": round(all_bytes * K / t_e2e / 1e9, 3), "unit": "GB/s", "ms_per_step": round(t_e2e / K * 1e3, 3),
                   "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) // K), "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) // K)},
           "bench_zig_row": {"what": "plain BLAKE3, 64 KiB of 0x42, 1000 buffers in one launch (src/bench.zig:207-223)",
                             "gpu_MiBps": round(zig_n * zig_len / t_zig / 2 ** 20, 1), "published_cpu_MiBps": 3517,
                             "published_source": "README.md:309-313 (CPU model not stated)", "digest_checked": True},
           "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
    print(json.dumps(out), file=out_stream, flush=True)
    ctx.close(); zd.shutdown()
    return 0


def cdc_bench(args, wl, out_stream):
    """configs[3]: Gearhash CDC + chunk hashes of one large bf16 shard (seed / dedup path).  value = device-resident
    zg_dev_cdc + zg_dev_hash_chunks; e2e = zg_cdc_hash on the pinned host buffer.  EVERY boundary and every chunk hash is
    compared with the oracle (serial CDC on the host, ~20 s for 16 GB) before timing."""
    torch, z, _native, zd, rank, world, local, dev, ctx = gpu_setup()
    from zest_b200._native import check
    from oracle import xet_oracle as oz
    L = _native.lib()
    K, W = args.steps, max(args.warmup, 3)
    cores = os.cpu_count() or 1
    nbytes = args.bytes or wl["nbytes"]
    pin = ctx.pinned_empty(nbytes + 16)
    piece = 256 << 20
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max(1, cores // world)) as ex:
        list(ex.map(lambda k: oz.synth_bf16_fast(min(piece, nbytes - k * piece), 7_000_000 * (rank + 1) + k, out=pin[k * piece:k * piece + min(piece, nbytes - k * piece)]),
                    range((nbytes + piece - 1) // piece)))
    log(f"generated {nbytes/1e9:.1f} GB in {time.perf_counter()-t0:.1f} s")
    d_data = torch.from_numpy(pin).to(dev)
    cap = nbytes // 8192 + 2
    d_ends = torch.zeros(cap, dtype=torch.int64, device=dev)
    d_n = torch.zeros(4, dtype=torch.int64, device=dev)
    s_cdc = torch.empty(L.zg_dev_cdc_scratch_bytes(nbytes), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)

    def k_cdc():
        check(L.zg_dev_cdc(ctx.handle, d_data.data_ptr(), nbytes, d_ends.data_ptr(), cap, d_n.data_ptr(), s_cdc.data_ptr(), s_cdc.numel(), sp))
    k_cdc()
    torch.cuda.synchronize()
    n = int(d_n[0])
    t0 = time.perf_counter()
    gold_ends = oz.cdc(pin[:nbytes])
    t_cpu_cdc = time.perf_counter() - t0
    got = d_ends[:n].cpu().numpy().astype(np.uint64)
    assert n == gold_ends.size and np.array_equal(got, gold_ends), "CDC boundaries differ from the oracle"
    d_off = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), d_ends[:n]])
    d_out = torch.zeros(32 * n, dtype=torch.uint8, device=dev)
    s_hash = torch.empty(L.zg_dev_hash_scratch_bytes(nbytes, n), dtype=torch.uint8, device=dev)

    def k_hash():
        check(L.zg_dev_hash_chunks(ctx.handle, d_data.data_ptr(), nbytes, d_off.data_ptr(), None, n, None, d_out.data_ptr(),
                                   s_hash.data_ptr(), s_hash.numel(), sp))

    def step():
        k_cdc(); k_hash()

    def timed(fn, steps):
        torch.cuda.synchronize(); zd.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(); zd.barrier()
        return zd.reduce_max(e0.elapsed_time(e1) / 1e3)
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    off_h = np.concatenate([np.zeros(1, np.uint64), gold_ends])
    t0 = time.perf_counter()
    gold_h = oz.hash_chunks(pin[:nbytes], off_h, cores)
    t_cpu_hash = time.perf_counter() - t0
    assert np.array_equal(d_out.cpu().numpy().reshape(-1, 32), gold_h), "chunk hashes differ from the oracle"
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start(); time.sleep(0.3)
    k0 = ctx.stats()["kernels_launched"]
    t_dev = timed(step, K)
    launches = ctx.stats()["kernels_launched"] - k0
    t_cdc = timed(k_cdc, K) / K
    t_hash = timed(k_hash, K) / K
    all_bytes = zd.reduce_sum(nbytes)
    # e2e: one zg_cdc_hash call on the pinned host buffer (H2D of the shard, CDC, hashes, file hash, D2H of ends + hashes)
    del d_data, s_cdc, s_hash, d_out, d_off
    torch.cuda.empty_cache()
    ke = max(2, min(K, 3))
    r = z.cdc_hash(pin[:nbytes], ctx=ctx)
    assert np.array_equal(np.asarray(r[0], np.uint64), gold_ends) and np.array_equal(np.asarray(r[1]).reshape(-1, 32), gold_h)
    s0 = ctx.stats()
    torch.cuda.synchronize(); zd.barrier()
    t0 = time.perf_counter()
    for _ in range(ke):
        z.cdc_hash(pin[:nbytes], ctx=ctx)
    torch.cuda.synchronize(); zd.barrier()
    t_e2e = zd.reduce_max(time.perf_counter() - t0)
    s1 = ctx.stats()
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        ctx.close(); zd.shutdown()
        return 0
    peaks = load_peaks()
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    tr = load_traffic(args.workload)
    roofline = {"kernel": "cdc_scan", "bound": "hbm", "achieved": round((nbytes + 8 * n) / t_cdc / 1e9, 2), "peak": hbm_peak, "unit": "GB/s",
                "frac": round((nbytes + 8 * n) / t_cdc / 1e9 / hbm_peak, 4), "traffic": tr.get("cdc"),
                "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)",
                "algorithmic_bytes_per_launch": int(nbytes + 8 * n), "avg_launch_ms": round(t_cdc * 1e3, 4),
                "stages": {"cdc": {"ms": round(t_cdc * 1e3, 4), "GBps": round(nbytes / t_cdc / 1e9, 2)},
                           "blake3_hash": {"ms": round(t_hash * 1e3, 4), "GBps": round(nbytes / t_hash / 1e9, 2)}}}
    cpu = None
    if not args.no_cpu:
        rows = {"C3 CDC (serial) + chunk hash, port: CDC 1 thread": round(nbytes / t_cpu_cdc / 1e9, 4),
                f"C3 chunk hash of the result, port, {cores} threads": round(nbytes / t_cpu_hash / 1e9, 4),
                f"C3 CDC + hash, port (CDC serial, hash on {cores} threads)": round(nbytes / (t_cpu_cdc + t_cpu_hash) / 1e9, 4)}
        try:
            import hf_xet
            import tempfile
            m = min(nbytes, 4 << 30)
            with tempfile.NamedTemporaryFile(dir="/dev/shm" if os.path.isdir("/dev/shm") else None, suffix=".bin") as f:
                f.write(memoryview(pin[:m])); f.flush()
                t0 = time.perf_counter()
                hf_xet.hash_files([f.name])
                rows["C4 hf_xet.hash_files (xet-core Rust: CDC + hash + Merkle), first 4 GiB from tmpfs, library threads"] = round(m / (time.perf_counter() - t0) / 1e9, 4)
        except Exception as e:  # not importable / API differs: the row is a sanity figure, not part of the metric
            rows["C4 hf_xet.hash_files"] = f"unavailable: {type(e).__name__}"
        cpu = {"value": rows[f"C3 CDC + hash, port (CDC serial, hash on {cores} threads)"], "unit": "GB/s", "cores": cores, "kind": "port",
               "sample": f"the whole shard ({nbytes/1e9:.1f} GB), one pass (it is also the parity check)", "rows": rows}
    out = {"metric": "CDC + chunk hash GB/s (Gearhash + BLAKE3 keyed)", "value": round(all_bytes * K / t_dev / 1e9, 3), "unit": "GB/s",
           "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(t_dev / K * 1e3, 4), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8", "data": "synthetic",
           "config": {"workload": wl["desc"], "name": args.workload, "bytes_per_rank": nbytes, "chunks_per_rank": int(n),
                      "parity": "every boundary and every chunk hash equal to the oracle's", "l2_policy": "input larger than L2, no flush needed"},
           "e2e": {"value": round(all_bytes * ke / t_e2e / 1e9, 3), "unit": "GB/s", "ms_per_step": round(t_e2e / ke * 1e3, 3), "steps": ke,
                   "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) // ke), "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) // ke)},
           "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
    print(json.dumps(out), file=out_stream, flush=True)
    ctx.close(); zd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
