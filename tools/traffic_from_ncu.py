"""ncu raw CSV of tools/prof_bench_trace.py (-k regex:trace_ -c 4) -> profiles/traffic.json entries.
usage: traffic_from_ncu.py <raw.csv> <out.json> [suffix]
The four trace launches are, in order: primary, ao, random (2^24 rays), random with RT_TRACE_REORDER."""
import csv
import json
import sys

NAMES = ["flake10m:primary", "flake10m:ao", "random1b", "random1b:reorder"]
RAYS = [3840 * 2160, 3840 * 2160 * 4, 1 << 24, 1 << 24]


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    rows = list(csv.reader(open(sys.argv[1], errors="replace")))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def col(r, name):
        for k, i in ix.items():
            if k == name or k.endswith("." + name):
                v = num(r[i])
                if v is None:
                    return None
                u = units[i]
                scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3,
                         "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3}.get(u, 1.0)
                return v * scale
        return None

    trace = [r for r in rows[2:] if "trace_" in r[ix["Kernel Name"]]]
    out = {}
    try:
        out = json.load(open(sys.argv[2]))
    except Exception:
        pass
    suffix = sys.argv[3] if len(sys.argv) > 3 else ""
    for name, n_rays, r in zip(NAMES, RAYS, trace):
        dram = (col(r, "dram__bytes_read.sum") or 0) + (col(r, "dram__bytes_write.sum") or 0)
        rec = {"kernel": r[ix["Kernel Name"]][:80], "ncu_us": col(r, "gpu__time_duration.sum"),
               "dram_bytes": dram, "dram_bytes_per_ray": dram / n_rays,
               "l1_sectors": col(r, "l1tex__t_sectors.sum"),
               "l1_sectors_miss": col(r, "l1tex__t_sectors_lookup_miss.sum"),
               "l1_lsu_wavefront_pct": col(r, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
               "l2_hit_pct": col(r, "lts__t_sector_hit_rate.pct"),
               "lts_pct": col(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
               "dram_pct": col(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
               "issue_pct": col(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
               "lanes_per_inst": col(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
               "warp_inst": col(r, "smsp__inst_executed.sum"),
               "warps_active_pct": col(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
               "registers": col(r, "launch__registers_per_thread"), "rays": n_rays}
        out[name + suffix] = rec
    json.dump(out, open(sys.argv[2], "w"), indent=1)
    for k in NAMES:
        if k + suffix in out:
            print(k + suffix, json.dumps(out[k + suffix]))


if __name__ == "__main__":
    main()
