#!/usr/bin/env python
"""Developer tool: share of warp-stall samples and of executed instructions per code region of one kernel, from an
.ncu-rep captured with `ncu --set full --import-source on` of a library built with -lineinfo.
  python tools/ncu_regions.py gpurun_out/x.ncu-rep [--top 25] > profiles/xxx_by_region.txt
Regions are cut at the marker lines below (found by text, so they follow the source as it moves); --top prints the
hottest source lines as well.
"""
import csv
import io
import re
import subprocess
import sys

# (region name, regex that matches the region's FIRST source line); order = order in lz4.cu
MARKERS = [
    ("rd8/xxh/load_window", r"^__device__ __forceinline__ uint32_t rd8"),
    ("load16 (global)", r"^template <bool READ_ONLY, int K>"),
    ("load16_win (smem)", r"^template <int K>\s*$|^// The same from the staged stream window"),
    ("store16", r"^template <int I>"),
    ("warp_copy", r"^// Whole-warp copy of non-overlapping ranges"),
    ("warp_match", r"^// Whole-warp LZ4 match copy"),
    ("regroup", r"^// Four bytes at any alignment"),
    ("cp_async/stage", r"^__device__ __forceinline__ void cp_async16"),
    ("tables (hop + exit)", r"^// Hop table \+ exit table"),
    ("seq_at", r"^// The sequence at window byte w"),
    ("hop fast", r"^struct Hop"),
    ("hop general", r"^__device__ __noinline__ Hop spec_hop_general\(.*\{"),
    ("match_rounds", r"^// The matches of up to 32 sequences in stream order"),
    ("exec_group", r"^// One execute step"),
    ("output ring", r"^// Output ring\."),
    ("scan", r"^__device__ __forceinline__ uint32_t warp_incl_scan"),
    ("kernel head+refill", r"^__global__ void __launch_bounds__\(32"),
    ("round: stage", r"// =+ (speculative|table) round"),
    ("spec: stage+run-up+parse loop", r"// ---- parse: lane k walks"),
    ("join", r"// ---- join:"),
    ("resolve", r"// ---- resolve:"),
    ("compose", r"// ---- compose:"),
    ("chain walk + list", r"// ---- chain walk:"),
    ("exec loop", r"// ---- execute:"),
    ("serial", r"// =+ serial round"),
    ("finish+regroup call", r"// ---- chunk finished this round"),
    ("(after kernel)", r"^}  // namespace$"),
]


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 0
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[hi]
    c_samp, c_inst = hdr.index("# Samples"), hdr.index("Instructions Executed")
    lines = []  # (line no, source, samples, instructions) of the FIRST file section (the kernel's own .cu file)
    other = {}  # later sections: inlined code from CUDA headers (shuffles, votes, funnel shifts), totals per file
    section = None
    for r in rows:
        if r and r[0] == "File Path":
            section = r[1]
            continue
        if section is not None and section != rows[0][1]:
            if r and r[0].strip().isdigit():
                try:
                    o = other.setdefault(section.split("/")[-1], [0.0, 0.0])
                    o[0] += float(r[c_samp]); o[1] += float(r[c_inst])
                except (ValueError, IndexError):
                    pass
            continue
        if r and r[0].strip().isdigit():
            def num(v):
                try:
                    return float(v)
                except ValueError:
                    return 0.0
            lines.append((int(r[0]), r[1], num(r[c_samp]) if len(r) > c_samp else 0.0, num(r[c_inst]) if len(r) > c_inst else 0.0))
    # the full source listing as captured in the report (lines without instructions are absent from the view above)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda"],
                         capture_output=True, text=True).stdout
    listing = [(int(r[0]), r[1]) for r in csv.reader(io.StringIO(src)) if len(r) >= 2 and r[0].strip().isdigit()]
    starts = []
    for name, pat in MARKERS:
        rx = re.compile(pat)
        ln = next((l for l, s in listing if rx.search(s)), None)
        if ln is not None:
            starts.append((ln, name))
    starts.sort()
    tot_s = (sum(x[2] for x in lines) + sum(o[0] for o in other.values())) or 1.0
    tot_i = (sum(x[3] for x in lines) + sum(o[1] for o in other.values())) or 1.0
    print(f"{rows[1][1][:120] if len(rows) > 1 and len(rows[1]) > 1 else ''}")
    print(f"total: {tot_s:.0f} stall samples, {tot_i:.4g} warp instructions")
    for k, (ln, name) in enumerate(starts):
        end = starts[k + 1][0] if k + 1 < len(starts) else 10 ** 9
        s = sum(x[2] for x in lines if ln <= x[0] < end)
        i = sum(x[3] for x in lines if ln <= x[0] < end)
        print(f"{name:32s} lines {ln}-{min(end, listing[-1][0] + 1) - 1}: stall samples {s / tot_s:.3f}  instructions {i / tot_i:.3f}")
    for f, o in sorted(other.items(), key=lambda kv: -kv[1][1]):
        print(f"{'(inlined from ' + f + ')':60s} stall samples {o[0] / tot_s:.3f}  instructions {o[1] / tot_i:.3f}")
    if top:
        print(f"\nhottest {top} source lines by instructions:")
        for l, s, sm, it in sorted(lines, key=lambda x: -x[3])[:top]:
            print(f"  {l:5d} inst {it / tot_i:.3f} samples {sm / tot_s:.3f}  {s.strip()[:110]}")
        print(f"\nhottest {top} source lines by stall samples:")
        for l, s, sm, it in sorted(lines, key=lambda x: -x[2])[:top]:
            print(f"  {l:5d} samples {sm / tot_s:.3f} inst {it / tot_i:.3f}  {s.strip()[:110]}")


if __name__ == "__main__":
    main()
