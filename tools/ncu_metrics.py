"""Prints the metrics that matter from an ncu raw CSV (development tool).
usage: ncu -i X.ncu-rep --page raw --csv > x.csv; python tools/ncu_metrics.py x.csv [kernel-substring]"""
import csv
import re
import sys

PAT = (r'gpu__time_duration.sum|dram__bytes_(read|write).sum$|lts__t_bytes.sum$|issue_active.avg.pct|'
       r'warps_active.avg.pct|throughput.avg.pct_of_peak_sustained_elapsed|smsp__inst_executed.sum$|'
       r'registers_per_thread|warp_issue_stalled.*_per_warp_active|lts__t_sector_hit_rate|'
       r'occupancy_limit|l1tex__t_(sectors|requests)_pipe_lsu_mem_global_op_(ld|st).sum$|'
       r'bank_conflicts_pipe_lsu_mem_shared.sum$|launch__grid_size|launch__block_size|shared_mem_per_block')
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    name = r[idx['Kernel Name']]
    if len(sys.argv) > 2 and sys.argv[2] not in name:
        continue
    print('---', name[:90])
    for h in hdr:
        if re.search(PAT, h):
            v = r[idx[h]]
            try:
                if float(v) == 0:
                    continue
            except ValueError:
                pass
            print('  ', h, v[:40], units[idx[h]])
