"""Top warp-stall locations of one kernel from `ncu -i X.ncu-rep --page source --csv [--kernel-name ...] > f.csv`.
usage: python tools/ncu_stalls.py f.csv [N] [section]"""
import csv
import sys

allrows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0          # section (kernel) index inside the file
starts = [i for i, r in enumerate(allrows) if r and r[0] == "Kernel Name"] + [len(allrows)]
rows = allrows[starts[which]:starts[which + 1]]
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[h], [r for r in rows[h + 1:] if len(r) == len(rows[h])]
si, src = hdr.index("# Samples"), hdr.index("Source")
cols = [i for i, x in enumerate(hdr) if x.startswith("stall_") and "Not Issued" not in x]
tot = sum(int(r[si] or 0) for r in data)
agg = {hdr[i]: sum(int(r[i] or 0) for r in data) for i in cols}
print(rows[0][1][:90] if len(rows[0]) > 1 else "", "| samples", tot)
print("  ".join("%s=%.0f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(data, key=lambda r: -int(r[si] or 0))[:n]:
    rs = sorted(((hdr[i][6:], int(r[i] or 0)) for i in cols), key=lambda kv: -kv[1])[:2]
    print("%6s  %-72s %s" % (r[si], r[src].strip()[:72], rs))
