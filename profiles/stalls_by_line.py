"""Aggregate warp-stall samples per CUDA source line from `ncu --page source --csv --print-source cuda,sass`."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cur_file = None
hdr = None
agg = {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        i_s = hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0] not in ("", "-"):                     # a CUDA source line row: totals for the SASS under it
        try:
            n = int(r[i_s] or 0)
        except ValueError:
            continue
        key = (cur_file, int(r[0]), r[1].strip()[:110])
        agg[key] = agg.get(key, 0) + n
tot = sum(agg.values()) or 1
print("total samples", tot)
for (f, ln, src), n in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    print("%6d %5.1f%%  %s:%d  %s" % (n, 100.0 * n / tot, f, ln, src))
