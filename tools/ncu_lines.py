"""Per-source-line stall samples of one kernel from an ncu report (--import-source on, -lineinfo):
    python tools/ncu_lines.py report.ncu-rep <kernel regex> [launch index] [N]"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
idx = sys.argv[3] if len(sys.argv) > 3 else "0"
n = int(sys.argv[4]) if len(sys.argv) > 4 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda", "--csv", "--kernel-name",
                      "regex:" + pat, "--launch-skip", idx, "--launch-count", "1"], capture_output=True, text=True).stdout
# sections: "File Name",<path> / header row / rows
sections, cur = [], None
for row in csv.reader(io.StringIO(raw)):
    if row and row[0] == "File Name":
        cur = {"file": row[1], "hdr": None, "rows": []}
        sections.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = row
    elif cur is not None:
        cur["rows"].append(row)
tot = 0
items = []
for s in sections:
    c = {h: i for i, h in enumerate(s["hdr"])}
    key = "# Samples" if "# Samples" in c else None
    if key is None:
        continue
    for r in s["rows"]:
        if len(r) == len(s["hdr"]) and r[c[key]].isdigit():
            v = int(r[c[key]])
            tot += v
            if v:
                items.append((v, s["file"].split("/")[-1], r[c["Line No"]], r[c["Source"]].strip()[:110]))
items.sort(reverse=True)
print("total samples", tot)
for v, f, ln, src in items[:n]:
    print("%6d %5.1f%%  %s:%s  %s" % (v, 100.0 * v / max(tot, 1), f, ln, src))
