#!/usr/bin/env python3
"""Condenses gpurun_out ncu artefacts into the small text files committed under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches.csv profiles/r01_launches.txt
  python tools/summarize_ncu.py full gpurun_out/prof_gemm.ncu-rep profiles/r01_gemm_full.txt
"""
import collections
import csv
import re
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum"]


def launches(src, dst):
    lines = [ln for ln in open(src) if not ln.startswith("==")]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("unnamed>::", "")
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
        tot[name] += v
        cnt[name] += 1
    T = sum(tot.values())
    with open(dst, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none: {sum(cnt.values())} launches, {T:.1f} us total\n")
        f.write("# (cold-cache, serialised: compare SHARES, not absolutes)\n")
        f.write(f"{'us':>10} {'share':>6} {'n':>4} {'avg_us':>8}  kernel\n")
        for k, v in sorted(tot.items(), key=lambda x: -x[1]):
            f.write(f"{v:10.1f} {100 * v / T:5.1f}% {cnt[k]:4d} {v / cnt[k]:8.1f}  {k}\n")
    print(open(dst).read())


def full(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(dst, "w") as f:
        f.write(f"# ncu --set full --clock-control none, {len(data)} launches from {src}\n")
        for d in data:
            name = re.sub(r"\(.*", "", d[idx["Kernel Name"]]).replace("void ", "").replace("unnamed>::", "")
            f.write(f"\n{name}\n")
            for m in METRICS:
                if m in idx:
                    f.write(f"  {m} [{units[idx[m]]}] = {d[idx[m]]}\n")
    print(f"wrote {dst} ({len(data)} launches)")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
