"""Top stall sites of one kernel from an ncu report captured with --import-source on:
    python tools/ncu_hot.py report.ncu-rep <kernel regex> [launch index] [N]"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
idx = sys.argv[3] if len(sys.argv) > 3 else "0"
n = int(sys.argv[4]) if len(sys.argv) > 4 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat, "--launch-skip", idx,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
lines = raw.split("\n")
print(lines[0][:160])
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]
c = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") or "Stall" in h]
data = []
for r in rows[1:]:
    if r and r[0] == "Kernel Name":
        break                      # next kernel section
    if len(r) == len(hdr) and r[c["# Samples"]].isdigit():
        data.append(r)
tot = sum(int(r[c["# Samples"]] or 0) for r in data)
data.sort(key=lambda r: -int(r[c["# Samples"]] or 0))
print("total samples", tot)
# per-reason columns present in the page
reason = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in data[:n]:
    rs = sorted(((int(r[c[h]] or 0), h) for h in reason), reverse=True)[:2] if reason else []
    print("%6s %5.1f%%  %-70s %s" % (r[c["# Samples"]], 100.0 * int(r[c["# Samples"]]) / max(tot, 1), r[c["Source"]].strip()[:70],
                                 " ".join("%s=%d" % (h[6:], v) for v, h in rs if v)))
agg = {}
for r in data:
    for h in reason:
        agg[h] = agg.get(h, 0) + int(r[c[h]] or 0)
print("by reason:", " ".join("%s=%.1f%%" % (h[6:], 100.0 * v / max(tot, 1)) for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
