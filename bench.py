#!/usr/bin/env python
"""bench.py -- xorb verify GB/s (decompress + BLAKE3 + Merkle) on N B200s, next to the CPU path.

Contract (driver): python bench.py --gpus N --steps K --warmup W [--impl reference]; under torchrun one rank per
GPU.  Rank 0 prints ONE JSON line.

  step        one pass of the hot path over one batch of synthetic xorbs (every rank owns its own disjoint set:
              xorbs are independent, so the path shards with no collective -> "scaling": "weak").
  value       device-resident GB/s: compressed xorbs + chunk descriptors already in HBM, one zg_dev_verify_xorbs
              per step on torch's current stream, CUDA events, max over ranks.  GB = 1e9 uncompressed bytes.
  e2e         same metric through the reference-facing call (zg_verify_xorbs) with pinned HOST buffers: host walk
              of the chunk stream, H2D of the compressed bytes, kernels, D2H of roots + verdicts, all inside the
              timed region.  `e2e.two_callers` is the same K calls issued by two host threads (how zest would call
              it); the one-caller figure stays `e2e.value`.
  roofline    the dominant kernel of the step (timed alone with CUDA events, same inputs), algorithmic bytes /
              time against the measured HBM copy bandwidth in MEASURED_PEAKS.json; `stages` lists every kernel;
              `blake3_alu` reports the hash kernel against the integer-ALU roof.
  cpu_baseline  the CPU oracle (oracle/, a port -- the reference's zig-xet code cannot be built here) on all host
              cores over the same xorbs.
  --impl reference   times only that CPU arm and prints it in the same shape.

Workloads (--workload): "gpt2-lz4" = BASELINE.json configs[1] (default; ~550 MB Zipf text, scheme 1);
"llama8b-bg4" = configs[2] scaled by --bytes (bf16 Gaussians, scheme 2); "random-none" (scheme 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "gpt2-lz4": dict(gen="synth_zipf_text", policy=1, nbytes=550_000_000,
                     desc="BASELINE.json configs[1]: gpt2-sized synthetic xorb set (~550 MB Zipf text, LZ4 chunks)"),
    "llama8b-bg4": dict(gen="synth_bf16", policy=2, nbytes=2_000_000_000,
                        desc="BASELINE.json configs[2] (scaled): Llama-3.1-8B-like bf16 Gaussians, BG4-LZ4 chunks"),
    "random-none": dict(gen="synth_random", policy=0, nbytes=550_000_000, desc="incompressible bytes, stored chunks"),
}
INT_OPS_PER_COMPRESSION = 792  # 7 rounds x 8 G x 14 + 8


def make_workload(name: str, nbytes: int, seed: int, threads: int):
    """-> (blobs, expected (n,32) u8, chunk_hashes, total uncompressed bytes, n_chunks).  Built with the oracle's
    generators and xorb builder (input generation only)."""
    from oracle import xet_oracle as oz
    w = WORKLOADS[name]
    blobs, expected, hashes, total, nchunks = [], [], [], 0, 0
    piece = 256 * 1024 * 1024  # generate in pieces so the host never holds more than one at a time
    done = 0
    k = 0
    while done < nbytes:
        n = min(piece, nbytes - done)
        data = getattr(oz, w["gen"])(n, seed * 1000 + k)
        xs, ends, hs = oz.make_xorbs(data, w["policy"], threads=threads)
        for x in xs:
            blobs.append(x.blob)
            expected.append(x.expected)
            nchunks += x.n_chunks
        hashes.append(hs)
        total += n
        done += n
        k += 1
    return blobs, np.frombuffer(b"".join(expected), np.uint8).reshape(-1, 32).copy(), np.concatenate(hashes), total, nchunks


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
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_arm(blobs, expected, total_bytes, threads, min_seconds=3.0, max_reps=20):
    """Times oracle verify (walk + decode + hash + fold + compare) on `threads` host threads; best of the reps."""
    from oracle import xet_oracle as oz
    best, reps, t_all = None, 0, time.perf_counter()
    while reps < 2 or (time.perf_counter() - t_all < min_seconds and reps < max_reps):
        t0 = time.perf_counter()
        v, _, nb = oz.verify_xorbs(blobs, expected, threads=threads)
        dt = time.perf_counter() - t0
        assert (v == 1).all() and nb == total_bytes
        best = dt if best is None else min(best, dt)
        reps += 1
    return total_bytes / best / 1e9, best, reps


def _claim_stdout():
    """Keep stdout for the single JSON line: anything libraries print to fd 1 (NCCL's version banner, for one) goes
    to stderr instead.  Returns a writer for the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def main():
    out_stream = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="gpt2-lz4", choices=sorted(WORKLOADS))
    ap.add_argument("--bytes", type=int, default=0, help="override the workload's uncompressed size per rank")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    K, W = args.steps, max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    wl = WORKLOADS[args.workload]
    nbytes = args.bytes or wl["nbytes"]
    cores = os.cpu_count() or 1
    host_threads = max(1, cores // max(world, 1)) if args.impl == "ours" else cores
    config = {"workload": wl["desc"], "name": args.workload, "uncompressed_bytes_per_rank": nbytes,
              "l2_policy": "inputs larger than L2 (compressed + decoded working set >> 126 MB), no flush needed",
              "parallelism": f"xorb-sharded x{world}, no collective",
              "chunk_index": "host walk of the 8-byte chunk headers (zg_xorb_index): inside the timed region of e2e; "
                             "value starts from blobs + 32-byte chunk descriptors in HBM"}

    # ------------------------------------------------------------------ reference arm: CPU only, rank 0 only
    if args.impl == "reference":
        if rank != 0:
            return 0
        blobs, expected, _, total, nchunks = make_workload(args.workload, nbytes, 1, cores)
        from oracle import xet_oracle as oz
        for _ in range(W):
            oz.verify_xorbs(blobs, expected, threads=cores)
        t0 = time.perf_counter()
        for _ in range(K):
            v, _, nb = oz.verify_xorbs(blobs, expected, threads=cores)
            assert (v == 1).all() and nb == total
        dt = time.perf_counter() - t0
        val = total * K / dt / 1e9
        sample = f"all {len(blobs)} xorbs ({total/1e6:.0f} MB uncompressed) per step"
        print(file=out_stream, flush=True, *[json.dumps({
            "impl": "reference", "metric": "xorb verify GB/s (decompress+BLAKE3+Merkle)", "value": round(val, 4), "unit": "GB/s",
            "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": round(dt / K * 1e3, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": round(val, 4), "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample,
                             "note": "oracle/xet_oracle.c (CPU restatement; zest's zig-xet path needs a Zig toolchain, absent here)"},
            "e2e": {"value": round(val, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0})])
        return 0

    # ------------------------------------------------------------------ our arm
    import torch
    import zest_b200 as z
    from zest_b200 import _native, distributed as zd
    from zest_b200._native import ChunkDesc, check
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: zest_b200 has no CPU fallback")
    rank, world, local = zd.init_from_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ctx = z.Context(local)
    L = _native.lib()

    blobs, expected, gold_hashes, total, nchunks = make_workload(args.workload, nbytes, 1 + rank, host_threads)
    nx = len(blobs)
    comp_bytes = sum(len(b) for b in blobs)

    # ---- device-resident inputs: concatenated blobs (16-byte aligned), descriptors from the host walk
    descs = (ChunkDesc * nchunks)()
    first = np.zeros(nx + 1, np.uint32)
    comp_host = ctx.pinned_empty(comp_bytes + 16 * nx + 16)
    cpos = dpos = k = 0
    views = []
    for i, b in enumerate(blobs):
        ix = z.xorb_index(b)
        first[i] = k
        comp_host[cpos:cpos + len(b)] = np.frombuffer(b, np.uint8)
        views.append(comp_host[cpos:cpos + len(b)])
        for j in range(ix.n_chunks):
            d = descs[k]
            d.src_off, d.dst_off, d.clen, d.ulen, d.scheme = cpos + int(ix.hdr_off[j]) + 8, dpos, int(ix.clen[j]), int(ix.ulen[j]), int(ix.scheme[j])
            dpos += (int(ix.ulen[j]) + 15) // 16 * 16
            k += 1
        cpos += (len(b) + 15) // 16 * 16
    first[nx] = k
    L.zg_chunk_order(C.byref(descs), nchunks)
    out_bytes = dpos
    d_comp = torch.from_numpy(comp_host[:cpos + 16]).to(dev)
    d_desc = torch.from_numpy(np.frombuffer(descs, np.uint8).copy()).to(dev)
    d_first = torch.from_numpy(first.view(np.uint8).copy()).to(dev)
    d_exp = torch.from_numpy(expected.reshape(-1)).to(dev)
    d_roots = torch.zeros(32 * nx, dtype=torch.uint8, device=dev)
    d_verdict = torch.zeros(nx, dtype=torch.int8, device=dev)
    d_hashes = torch.zeros(32 * nchunks, dtype=torch.uint8, device=dev)
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
    assert bool((d_verdict == 1).all()), "device-resident verify: verdicts"
    assert np.array_equal(d_hashes.cpu().numpy().reshape(-1, 32), gold_hashes), "device-resident verify: chunk hashes"

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    k0 = ctx.stats()["kernels_launched"]
    t_dev = timed(step_dev, K)
    launches = ctx.stats()["kernels_launched"] - k0
    all_bytes = zd.reduce_sum(total)
    value = all_bytes * K / t_dev / 1e9

    # ---- e2e: reference-facing call with pinned host buffers
    def step_e2e():
        res = z.verify_xorbs(views, expected, ctx=ctx)
        return res
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
    s1 = ctx.stats()
    e2e_val = all_bytes * K / t_e2e / 1e9
    # The same K calls issued by two host threads (zest runs up to 128 download tasks at once,
    # src/parallel_download.zig:117): the tail of one call -- its last chunks are a serial chain of ~2 ms after the
    # last byte has arrived -- then overlaps the copies of the next.  Reported beside the one-caller figure, which
    # stays the headline.
    import threading

    def two_caller_pass():
        nxt, lock = [0], threading.Lock()

        def caller():
            while True:
                with lock:
                    if nxt[0] >= K:
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
    t_e2e2 = sorted(two_caller_pass() for _ in range(3))[1]  # median of three: host-thread scheduling makes it noisy
    clocks = sampler.stop() if rank == 0 else None
    # what the link alone does: the same compressed bytes, pinned host -> device, nothing else (bounds e2e from above)
    pin_t = torch.from_numpy(comp_host[:cpos])
    for _ in range(2):
        d_comp[:cpos].copy_(pin_t, non_blocking=True)
    t_link = timed(lambda: d_comp[:cpos].copy_(pin_t, non_blocking=True), 5) / 5
    link_gbps = cpos / t_link / 1e9

    # ---- per-kernel timing for the roofline (rank 0's numbers; same inputs, each stage alone)
    off = np.array([d.dst_off for d in descs], np.uint64)
    lens = np.array([d.ulen for d in descs], np.uint32)
    d_off, d_len = torch.from_numpy(off.view(np.uint8).copy()).to(dev), torch.from_numpy(lens.view(np.uint8).copy()).to(dev)
    d_out = torch.empty(out_bytes + 16, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(nchunks, dtype=torch.int32, device=dev)
    s_dec = torch.empty(L.zg_dev_decode_scratch_bytes(nchunks, out_bytes), dtype=torch.uint8, device=dev)
    s_hash = torch.empty(L.zg_dev_hash_scratch_bytes(out_bytes, nchunks), dtype=torch.uint8, device=dev)
    s_mer = torch.empty(L.zg_dev_merkle_scratch_bytes(nchunks, nx), dtype=torch.uint8, device=dev)

    def k_decode():
        check(L.zg_dev_decode_chunks(ctx.handle, d_comp.data_ptr(), cpos, d_desc.data_ptr(), nchunks, d_out.data_ptr(), out_bytes,
                                     d_status.data_ptr(), s_dec.data_ptr(), s_dec.numel(), sp))

    def k_hash():
        check(L.zg_dev_hash_chunks(ctx.handle, d_out.data_ptr(), out_bytes, d_off.data_ptr(), d_len.data_ptr(), nchunks, None,
                                   d_hashes.data_ptr(), s_hash.data_ptr(), s_hash.numel(), sp))

    def k_merkle():
        check(L.zg_dev_merkle(ctx.handle, d_hashes.data_ptr(), d_len.data_ptr(), d_first.data_ptr(), nx, nchunks, d_status.data_ptr(),
                              d_exp.data_ptr(), d_roots.data_ptr(), d_verdict.data_ptr(), s_mer.data_ptr(), s_mer.numel(), sp))

    stages = {}
    for name, fn in (("lz4_decode", k_decode), ("blake3_hash", k_hash), ("merkle_fold", k_merkle)):
        for _ in range(3):
            fn()
        stages[name] = timed(fn, K) / K
    assert int(d_status.abs().sum()) == 0 and bool((d_verdict == 1).all())

    if rank != 0:
        ctx.close()
        zd.shutdown()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    alg = {"lz4_decode": comp_bytes + total, "blake3_hash": total + 32 * nchunks, "merkle_fold": 40 * nchunks + 33 * nx}
    dom = max(stages, key=lambda s: stages[s])
    achieved = alg[dom] / stages[dom] / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {}).get(dom)
    except (OSError, ValueError):
        pass
    # The decoder is bound by instruction issue, not by HBM: put its launch time against the issue floor of the warp
    # instructions ncu counted for this workload (profiles/traffic.json), 4 schedulers per SM at the observed clock.
    lz4_issue = None
    try:
        n_inst = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {}).get("lz4_decode_warp_instructions")
        if n_inst:
            n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
            mhz = (clocks or {}).get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
            floor_ms = n_inst / (4.0 * n_sm * mhz * 1e6) * 1e3
            lz4_issue = {"warp_instructions": int(n_inst), "issue_floor_ms": round(floor_ms, 4),
                         "frac_of_issue_floor": round(floor_ms / (stages["lz4_decode"] * 1e3), 4),
                         "note": "smsp__inst_executed.sum of one launch (ncu, profiles/r1_ncu_decode_spec.txt) / (4 schedulers x SMs x SM clock)"}
    except (OSError, ValueError):
        pass
    from oracle import xet_oracle as oz
    compressions = sum(oz.blake3_compressions(int(n)) for n in lens)
    int_ops = compressions * INT_OPS_PER_COMPRESSION
    sm_mhz = (clocks or {}).get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
    # integer-pipe calibration (zg_dev_int_pipe_peak): measured thread-op rates of the ALU pipe (LOP3/SHF/IADD3), the
    # FMA pipe (IMAD) and both together.  The hash kernel issues, per 64-byte compression, 448 ALU-pipe and 336
    # FMA-pipe instructions per thread (adds as IMAD), so its floor is max(448/alu, 336/fma, 784/both) per compression.
    sink = torch.zeros(4, dtype=torch.uint8, device=dev)
    g_alu, g_fma, g_both = C.c_double(), C.c_double(), C.c_double()
    check(L.zg_dev_int_pipe_peak(sink.data_ptr(), C.byref(g_alu), C.byref(g_fma), C.byref(g_both), sp))
    t_floor = compressions * max(448 / (g_alu.value * 1e9), 336 / (g_fma.value * 1e9), 784 / (g_both.value * 1e9))
    roofline = {"kernel": dom, "bound": "hbm", "achieved": round(achieved, 2), "peak": hbm_peak, "unit": "GB/s",
                "frac": round(achieved / hbm_peak, 4), "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(alg[dom]), "avg_launch_ms": round(stages[dom] * 1e3, 4),
                "stages": {s: {"ms": round(t * 1e3, 4), "alg_GBps": round(alg[s] / t / 1e9, 2),
                               "uncompressed_GBps": round(total / t / 1e9, 2)} for s, t in stages.items()},
                "lz4_issue": lz4_issue,
                "blake3_alu": {"compressions": int(compressions), "int_ops": int(int_ops),
                               "achieved_Tops": round(int_ops / stages["blake3_hash"] / 1e12, 3),
                               "measured_pipe_peaks_Tops": {"alu": round(g_alu.value / 1e3, 2), "fma_imad": round(g_fma.value / 1e3, 2),
                                                            "both": round(g_both.value / 1e3, 2)},
                               "frac_of_alu_pipe": round(int_ops / stages["blake3_hash"] / 1e9 / g_alu.value, 4),
                               "frac_of_issue_floor": round(t_floor / stages["blake3_hash"], 4),
                               "note": "int_ops = 792 per compression (SURVEY.md 8d); the kernel issues 448 ALU-pipe + 336 "
                                       "FMA-pipe (IMAD adds) thread-instructions per compression; issue floor = "
                                       "max(448/alu, 336/fma, 784/both) from the measured pipe rates",
                               "sm_mhz": sm_mhz}}
    cpu = None
    if not args.no_cpu:
        v, best, reps = cpu_arm(blobs, expected, total, cores)
        cpu = {"value": round(v, 4), "unit": "GB/s", "cores": cores, "kind": "port",
               "sample": f"all {nx} xorbs of rank 0 ({total/1e6:.0f} MB uncompressed), best of {reps} passes"}
    out = {
        "metric": "xorb verify GB/s (decompress+BLAKE3+Merkle)", "value": round(value, 3), "unit": "GB/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(t_dev / K * 1e3, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(config, n_xorbs_per_rank=nx, n_chunks_per_rank=nchunks, compressed_bytes_per_rank=comp_bytes,
                       compression_ratio=round(comp_bytes / total, 4)),
        "e2e": {"value": round(e2e_val, 3), "unit": "GB/s", "ms_per_step": round(t_e2e / K * 1e3, 3),
                "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) // K),
                "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) // K),
                "h2d_GBps": round((s1["h2d_bytes"] - s0["h2d_bytes"]) / t_e2e / 1e9, 2),
                "two_callers": {"value": round(all_bytes * K / t_e2e2 / 1e9, 3), "ms_per_step": round(t_e2e2 / K * 1e3, 3)},
                "h2d_link_GBps": round(link_gbps, 2),
                "link_bound_GBps_uncompressed": round(link_gbps * total / comp_bytes, 2)},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(out), file=out_stream, flush=True)
    ctx.close()
    zd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
