#!/usr/bin/env python3
"""Top stall sites of a kernel from an ncu report's source page (SASS with stall-sample counts).

  python tools/ncu_hot.py gpurun_out/prof.ncu-rep [top_n] [launch_index]
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    # several launches are concatenated: split on the "Kernel Name" header rows
    blocks, cur = [], []
    for line in raw.splitlines():
        if line.startswith('"Kernel Name"'):
            if cur:
                blocks.append(cur)
            cur = [line]
        elif cur:
            cur.append(line)
    if cur:
        blocks.append(cur)
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    b = blocks[which]
    print(b[0][:160])
    rows = list(csv.DictReader(io.StringIO("\n".join(b[1:]))))
    tot = sum(int(r["# Samples"] or 0) for r in rows)
    stall_cols = [k for k in rows[0] if k.startswith("stall_") and "Not Issued" not in k]
    agg = {k: sum(int(r[k] or 0) for r in rows) for k in stall_cols}
    print("total samples", tot, " instructions", len(rows))
    print("stall mix:", ", ".join(f"{k[6:]}={100 * v / max(1, tot):.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    idx = sorted(range(len(rows)), key=lambda i: -int(rows[i]["# Samples"] or 0))[:top]
    for i in sorted(idx):
        r = rows[i]
        s = int(r["# Samples"] or 0)
        why = sorted(((int(r[k] or 0), k[6:]) for k in stall_cols), reverse=True)[:2]
        print(f"{i:5d} {100 * s / max(1, tot):5.1f}%  exec={r['Instructions Executed']:>9}  {r['Source'][:90]:90s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")


def window(rep, lo, hi, which=0):
    """SASS lines lo..hi of the source page with their sample counts (to read the code around a hot spot)."""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    blocks, cur = [], []
    for line in raw.splitlines():
        if line.startswith('"Kernel Name"'):
            if cur:
                blocks.append(cur)
            cur = [line]
        elif cur:
            cur.append(line)
    if cur:
        blocks.append(cur)
    rows = list(csv.DictReader(io.StringIO("\n".join(blocks[which][1:]))))
    for i in range(max(0, lo), min(len(rows), hi)):
        r = rows[i]
        print(f"{i:5d} {int(r['# Samples'] or 0):6d}  exec={r['Instructions Executed']:>9}  {r['Source'][:110]}")


if __name__ == "__main__" and len(sys.argv) > 2 and sys.argv[2] == "window":
    window(sys.argv[1], int(sys.argv[3]), int(sys.argv[4]))
    sys.exit(0)

if __name__ == "__main__":
    main()
