#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU): headline raw metrics per kernel and, from the
source page, shared-memory wavefronts / global tag requests per 32 records by instruction group.
  python tools/ncu_summary.py gpurun_out/x.ncu-rep --records 200000000"""
import argparse, collections, csv, io, re, subprocess, sys

ap = argparse.ArgumentParser()
ap.add_argument("rep")
ap.add_argument("--records", type=float, default=200e6)
ap.add_argument("--source", action="store_true")
args = ap.parse_args()
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
raw = subprocess.run(["ncu", "-i", args.rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
per32 = args.records / 32
for r in rows[2:]:
    print("----")
    print(r[hdr.index("Kernel Name")])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            extra = ""
            if w.endswith(".sum") and "wavefront" in w or "conflict" in w or "inst_executed" in w:
                try:
                    extra = f"   ({float(r[i]) / per32:.2f} per 32 records)"
                except ValueError:
                    pass
            print(f"  {w:72s} {r[i]} {units[i]}{extra}")
    stalls = []
    for i, h in enumerate(hdr):
        m = re.match(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active.ratio", h)
        if m:
            try:
                stalls.append((float(r[i]), m.group(1)))
            except ValueError:
                pass
    print("  stalls per issue:", ", ".join(f"{n}={v:.2f}" for v, n in sorted(stalls, reverse=True)[:8]))
if args.source:
    src = subprocess.run(["ncu", "-i", args.rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    for si, st in enumerate(starts):
        end = starts[si + 1] if si + 1 < len(starts) else len(rows)
        print("==== source:", rows[st][1][:100])
        hdr = rows[st + 1]
        ix = {h: i for i, h in enumerate(hdr)}
        def I(x):
            try:
                return int(x)
            except ValueError:
                return 0
        agg = collections.OrderedDict()
        for r in rows[st + 2:end]:
            if len(r) < len(hdr):
                continue
            w = I(r[ix["L1 Wavefronts Shared"]]); wi = I(r[ix["L1 Wavefronts Shared Ideal"]])
            tg = I(r[ix["L1 Tag Requests Global"]]); smp = I(r[ix["# Samples"]])
            s = re.sub(r"^@!?U?P\d\s+", "", r[ix["Source"]].strip())
            op = s.split()[0]
            a = agg.setdefault(op, [0, 0, 0, 0, 0])
            a[0] += I(r[ix["Instructions Executed"]]); a[1] += w; a[2] += wi; a[3] += tg; a[4] += smp
        tot = sum(a[4] for a in agg.values()) or 1
        for op, a in sorted(agg.items(), key=lambda kv: -kv[1][4]):
            if a[1] or a[3] or a[4] * 50 > tot:
                print(f"  {op:28s} inst/32={a[0] / per32:7.2f} smem_wf/32={a[1] / per32:6.2f} (ideal {a[2] / per32:5.2f})"
                      f" gtag/32={a[3] / per32:5.2f} samples={100 * a[4] / tot:5.1f}%")
