"""Print the top warp-stall instructions of one kernel from `ncu --page source --csv` output."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Source" in r and "Address" in r)
hdr = rows[h]
i_src, i_s, i_n = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
data = []
for k, r in enumerate(rows[h + 1:]):
    if len(r) > i_n:
        try:
            data.append((int(r[i_s] or 0), int(r[i_n] or 0), r[i_src].strip(), k))
        except ValueError:
            pass
tot = sum(d[0] for d in data) or 1
print("total samples", tot, "instructions", len(data), "warp-instr executed", sum(d[1] for d in data))
for s, n, src, k in sorted(data, reverse=True)[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print("%6d %5.1f%% n=%9d #%4d %s" % (s, 100 * s / tot, n, k, src[:100]))
