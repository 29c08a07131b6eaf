"""Turns ncu CSV dumps into the markdown summaries kept under profiles/ (development tool).

  summarize_ncu.py launches <launch-list.csv> "<title>"   # from --metrics gpu__time_duration.sum --csv
  summarize_ncu.py full <raw.csv> "<title>"               # from ncu -i X.ncu-rep --page raw --csv
"""
import csv
import re
import sys
from collections import OrderedDict

FULL = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_local_op_st.sum",
        "launch__registers_per_thread", "launch__shared_mem_per_block", "launch__grid_size",
        "launch__block_size"]


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = name.replace("void ", "").replace("rt::<unnamed>::", "").replace("unnamed>::", "")
    name = re.sub(r"\(bool\)|\(int\)", "", name)
    return name.strip()[:100]


def launches(path, title):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr = next(r for r in rows if r and r[0] == "ID")
    ix = {h: i for i, h in enumerate(hdr)}
    agg = OrderedDict()
    total = 0.0
    for r in rows:
        if len(r) != len(hdr) or r[0] == "ID" or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        us = v / 1e3 if unit in ("ns", "nsecond") else (v * 1e3 if unit in ("ms", "msecond") else v)
        k = short(r[ix["Kernel Name"]])
        a = agg.setdefault(k, [0, 0.0, r[ix["Grid Size"]], r[ix["Block Size"]]])
        a[0] += 1
        a[1] += us
        total += us
    print(f"# {title}\n")
    print("`ncu --metrics gpu__time_duration.sum --clock-control none`; per-launch times are cold-cache "
          "and serialised: compare shares, not absolutes.\n")
    print("| kernel | launches | avg us | share of listed GPU time | grid | block |")
    print("|---|---|---|---|---|---|")
    for k, (n, us, grid, block) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {us / n:.1f} | {100 * us / total:.1f}% | {grid} | {block} |")


def full(path, title):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print(f"# {title}\n")
    seen = {}
    for r in rows[2:]:
        k = short(r[ix["Kernel Name"]])
        seen[k] = seen.get(k, 0) + 1
        if seen[k] > 2:
            continue  # the first two launches of each kernel are enough
        print(f"## `{k}` (launch {seen[k]})\n\n| metric | value |\n|---|---|")
        for m in FULL:
            if m in ix:
                print(f"| {m} | {r[ix[m]][:24]} {units[ix[m]]} |")
        print()


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
