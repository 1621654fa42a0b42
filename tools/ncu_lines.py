#!/usr/bin/env python3
"""Per-source-line hot spots of one kernel in an .ncu-rep (read here, no GPU): warp instructions
executed, shared-memory wavefronts and stall samples per CUDA source line, scaled per 32 records.
  python tools/ncu_lines.py gpurun_out/x.ncu-rep --kernel local_count --records 1e9 [--top 40]"""
import argparse, csv, io, subprocess

ap = argparse.ArgumentParser()
ap.add_argument("rep")
ap.add_argument("--kernel", required=True)
ap.add_argument("--records", type=float, default=1e9)
ap.add_argument("--top", type=int, default=40)
ap.add_argument("--by", default="inst", choices=["inst", "wf", "samp"])
args = ap.parse_args()
raw = subprocess.run(["ncu", "-i", args.rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                      "--kernel-name", "regex:" + args.kernel], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
per32 = args.records / 32
fname, hdr, out = "", None, []
seen_fn = 0
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r and r[0] == "Function Name":
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if not hdr or not r or r[0] == "":
        continue
    def g(name):
        try:
            return float(r[hdr.index(name)])
        except (ValueError, IndexError):
            return 0.0
    out.append({"file": fname, "line": r[0], "src": r[1].strip()[:100], "inst": g("Instructions Executed"),
                "wf": g("L1 Wavefronts Shared"), "wfi": g("L1 Wavefronts Shared Ideal"), "samp": g("# Samples"),
                "gl": g("L1 Tag Requests Global")})
tot = {k: sum(o[k] for o in out) for k in ("inst", "wf", "wfi", "samp", "gl")}
print(f"total per 32 records: inst {tot['inst']/per32:.1f}  smem wavefronts {tot['wf']/per32:.2f} (ideal {tot['wfi']/per32:.2f})  "
      f"global tag requests {tot['gl']/per32:.2f}  samples {tot['samp']:.0f}")
for o in sorted(out, key=lambda o: -o[args.by])[: args.top]:
    print(f"{o['inst']/per32:7.2f} inst  wf {o['wf']/per32:5.2f}/{o['wfi']/per32:5.2f}  samp {100*o['samp']/max(tot['samp'],1):5.1f}%  "
          f"{o['file']}:{o['line']}: {o['src']}")
