"""Summarise one `ncu --set full` report (.ncu-rep) into a small text file for profiles/.

    python tools/ncu_summary.py gpurun_out/prof_r1_q1.ncu-rep profiles/r1_q1_full.txt [--json profiles/traffic.json --key q1]

Writes the headline metrics (duration, DRAM bytes, instruction counts, IPC, stall reasons, occupancy) and the
hottest SASS regions by executed instructions and by stall samples. With --json the per-launch DRAM traffic is merged
into a JSON file that bench.py reads for `roofline.traffic`.
"""
import argparse
import csv
import io
import json
import subprocess


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__cycles_elapsed.avg.per_second",
]


def to_bytes(val, unit):
    v = float(val.replace(",", ""))
    u = unit.lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("out")
    ap.add_argument("--json")
    ap.add_argument("--key")
    ap.add_argument("--rows", type=int, default=0, help="rows the profiled launch processed (for per-row figures)")
    a = ap.parse_args()
    raw = ncu_csv(a.rep, "raw")
    hdr, units, vals = raw[0], raw[1], raw[2]
    m = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
    lines = [f"report: {a.rep}", f"kernel: {m['Kernel Name'][0]}", ""]
    for w in WANT:
        if w in m:
            lines.append(f"{w:70s} {m[w][0]} {m[w][1]}")
    stalls = sorted(((float(v[0]), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
                     for h, v in m.items() if "issue_stalled" in h and "per_issue_active" in h), reverse=True)
    lines += ["", "warp stall reasons (stalled warps per issue-active cycle):"] + [f"  {n:28s} {x:.3f}" for x, n in stalls[:10]]
    rd = to_bytes(*m["dram__bytes_read.sum"])
    wr = to_bytes(*m["dram__bytes_write.sum"])
    lines += ["", f"DRAM traffic of this launch: {rd + wr:.0f} bytes (read {rd:.0f} + write {wr:.0f})"]
    if a.rows:
        inst = float(m["smsp__inst_executed.sum"][0].replace(",", ""))
        lines.append(f"per row: {(rd + wr) / a.rows:.2f} DRAM bytes, {inst * 32 / a.rows:.1f} thread-instructions ({inst / (a.rows / 32):.1f} warp-instructions per 32 rows)")
    src = ncu_csv(a.rep, "source")
    if len(src) > 3:
        h = src[1]
        rows = src[2:]
        ie, sa, so = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
        tot_i = sum(int(r[ie]) for r in rows)
        tot_s = sum(int(r[sa]) for r in rows)
        lines += ["", f"SASS: {len(rows)} instructions, {tot_i} warp-level executions, {tot_s} stall samples", "top instructions by stall samples:"]
        for i in sorted(range(len(rows)), key=lambda i: -int(rows[i][sa]))[:15]:
            r = rows[i]
            lines.append(f"  #{i:5d} samples {int(r[sa]):6d} exec {int(r[ie]):9d}  {r[so].strip()[:80]}")
        mix = {}
        for r in rows:
            s = r[so].strip()
            if s.startswith("@"):
                s = s.split(None, 1)[1]
            op = s.split()[0].split(".")[0]
            mix[op] = mix.get(op, 0) + int(r[ie])
        lines += ["", "executed instruction mix (warp-level):"] + [f"  {k:10s} {v:12d} {100 * v / tot_i:5.1f}%" for k, v in sorted(mix.items(), key=lambda kv: -kv[1])[:14]]
    open(a.out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:40]))
    if a.json and a.key:
        try:
            d = json.load(open(a.json))
        except Exception:
            d = {}
        d[a.key] = {"dram_bytes_per_launch": rd + wr, "rows_per_launch": a.rows, "report": "profiles/" + __import__("os").path.basename(a.out), "kernel": m["Kernel Name"][0][:80]}
        json.dump(d, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
