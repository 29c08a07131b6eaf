"""Summarise an `ncu --page source --csv` dump: stall totals and the hottest SASS lines."""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
# the dump may hold several kernels: split on "Kernel Name" rows
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
for b in blocks[:1]:
    hdr = b["rows"][0]
    body = [r for r in b["rows"][1:] if len(r) == len(hdr)]
    print(b["name"][:120])
    si = hdr.index("# Samples")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[si] or 0) for r in body)
    print("total samples", tot)
    agg = {hdr[i]: sum(int(r[i] or 0) for r in body) for i in stall_cols}
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]:
        print(f"  {k:28s} {v:8d} {100.0 * v / max(tot, 1):5.1f}%")
    body.sort(key=lambda r: -int(r[si] or 0))
    for r in body[:top]:
        st = sorted(((int(r[i] or 0), hdr[i]) for i in stall_cols), reverse=True)[:2]
        print(f"{int(r[si]):7d} {100.0 * int(r[si]) / max(tot, 1):5.1f}%  {r[1][:90]:90s} {st}")
