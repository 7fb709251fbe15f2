"""List the hot loop of a profiled kernel: every SASS instruction executed at least --min times per warp-tile, with its
stall samples (from `ncu --set full --import-source on`).

    python tools/ncu_loop.py report.ncu-rep [--unit 2097152] [--min 0.25] [--out file]
"""
import argparse
import csv
import io
import subprocess


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--unit", type=float, default=2097152.0, help="executions of one loop-body instruction (warp-tiles of the launch)")
    ap.add_argument("--min", type=float, default=0.25)
    ap.add_argument("--out")
    a = ap.parse_args()
    out = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = rows[1]
    ie, sa, so = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
    lines, tot, samples = [], 0.0, 0
    for i, r in enumerate(rows[2:]):
        e = int(r[ie]) / a.unit
        if e >= a.min:
            lines.append(f"{i:5d} {e:6.2f} {int(r[sa]):5d}  {r[so].strip()[:100]}")
            tot += e
            samples += int(r[sa])
    head = f"{len(lines)} instructions, {tot:.1f} executions per warp-tile, {samples} samples"
    text = head + "\n" + "\n".join(lines) + "\n"
    if a.out:
        open(a.out, "w").write(text)
    print(head)


if __name__ == "__main__":
    main()
