#!/usr/bin/env python
"""Turns an ncu launch list (--metrics gpu__time_duration.sum --csv) and optional
`ncu --set full` reports into the Markdown summaries kept under profiles/.

    python tools/ncu_summary.py --title "..." --launches gpurun_out/launches.csv \
        --full gpurun_out/prof_a.ncu-rep gpurun_out/prof_b.ncu-rep > profiles/rN_summary.md
"""
import argparse
import collections
import csv
import subprocess

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic",
]


def launches(path, start_at=None, stop_at=None):
    """start_at = "kernel:n": count from the n-th launch (1-based) of the first kernel whose name
    contains `kernel` - to leave out an untimed first call of an entry point."""
    lines = [l for l in open(path) if not l.startswith("==")]
    agg, n = collections.OrderedDict(), 0
    want, seen = (start_at.split(":")[0], int(start_at.split(":")[1])) if start_at else (None, 0)
    stop, left = (stop_at.split(":")[0], int(stop_at.split(":")[1])) if stop_at else (None, 0)
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        if stop is not None and stop in row["Kernel Name"]:
            left -= 1
            if left <= 0:
                break
        if want is not None:
            if want in row["Kernel Name"]:
                seen -= 1
            if seen > 0:
                continue
            want = None
        v, u = float(row["Metric Value"].replace(",", "")), row["Metric Unit"]
        v = v / 1e6 if u == "ns" else v / 1e3 if u == "us" else v * 1e3 if u == "s" else v
        a = agg.setdefault(row["Kernel Name"], [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    return agg, n


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    if len(rows) < 3:
        return []
    hdr, units = rows[0], rows[1]
    res = []
    for row in rows[2:]:
        d = {"Kernel Name": row[hdr.index("Kernel Name")]}
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                d[m] = f"{row[i]} {units[i]}".strip()
        res.append(d)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--title", required=True)
    ap.add_argument("--command", default="")
    ap.add_argument("--launches")
    ap.add_argument("--stop-at", default=None, help="kernel:n - end the launch list before the n-th launch of that kernel")
    ap.add_argument("--start-at", default=None, help="kernel:n - start the launch list at the n-th launch of that kernel")
    ap.add_argument("--full", nargs="*", default=[])
    ap.add_argument("--note", default="")
    ap.add_argument("--only", default="", help="comma-separated launch indices to keep from each --full report")
    a = ap.parse_args()
    print(f"# {a.title}\n")
    if a.command:
        print(f"Command (B200, after the same command exited 0 without ncu): `{a.command}`\n")
    if a.launches:
        agg, n = launches(a.launches, a.start_at, a.stop_at)
        tot = sum(v[1] for v in agg.values())
        print(f"## Launch list: {n} launches, {tot:.1f} ms in total (cold-cache, serialised: compare shares)\n")
        print("| share | total ms | launches | kernel |\n|---|---|---|---|")
        for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:18]:
            print(f"| {100 * ms / tot:.2f}% | {ms:.3f} | {c} | `{k.split('(')[0][:90]}` |")
        print()
    for path in a.full:
        keep = {int(x) for x in a.only.split(",")} if a.only else None
        for idx, d in enumerate(full(path)):
            if keep is not None and idx not in keep:
                continue
            print(f"## `ncu --set full`: `{d['Kernel Name'].split('(')[0]}` ({path.split('/')[-1]})\n")
            print("| metric | value |\n|---|---|")
            for m in METRICS:
                if m in d:
                    print(f"| {m} | {d[m]} |")
            print()
    if a.note:
        print(a.note)


if __name__ == "__main__":
    main()
