"""per-CUDA-source-line executed instructions and stall samples of one ncu report taken with --import-source on
    python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur = None; hdr = None; agg = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No":
        hdr = r; iss = hdr.index("Warp Stall Sampling (All Samples)"); iex = hdr.index("Instructions Executed"); continue
    if hdr and len(r) > iex and r[0].isdigit() and r[iss].isdigit():
        a = agg.setdefault((cur, int(r[0])), [r[1], 0, 0]); a[1] += int(r[iss]); a[2] += int(r[iex] or 0)
tot_s = sum(a[1] for a in agg.values()); tot_e = sum(a[2] for a in agg.values())
print("stall samples", tot_s, "warp instructions", tot_e)
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print(f"{k[0]:16s} {k[1]:5d} exec {a[2]:11d} ({100*a[2]/tot_e:4.1f}%) smp {a[1]:6d}  {a[0].strip()[:110]}")
