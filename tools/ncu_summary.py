"""Summarise an `ncu --set full` report of `bench.py --ncu-slice` (one launch per kernel group) into
  profiles/<name>_summary.csv   time, DRAM bytes, DRAM %, tensor-pipe %, issue %, registers per launch
  profiles/ncu_traffic.json     {kernel-group tag: dram__bytes_read.sum + dram__bytes_write.sum per launch}  (read by bench.py)

usage: python tools/ncu_summary.py gpurun_out/X.ncu-rep gpurun_out/slice_ncu.log profiles/r1_sNN_ncu_full_slice
The slice log holds the JSON line {"ncu_slice": [tags in launch order]} printed by bench.py --ncu-slice."""
import csv
import io
import json
import subprocess
import sys

rep, log, out = sys.argv[1:4]
tags = None
for line in open(log):
    if line.startswith("{") and "ncu_slice" in line:
        tags = json.loads(line)["ncu_slice"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, data = rows[0], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
want = [("time_us", "gpu__time_duration.sum"), ("dram_read_MB", "dram__bytes_read.sum"),
        ("dram_write_MB", "dram__bytes_write.sum"),
        ("dram_pct_peak", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("tensor_pipe_pct_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
        ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("regs", "launch__registers_per_thread"),
        ("sm_throughput_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed")]
units = rows[1]
# expanded launches: unpack_linear_grad is two kernels; map by order, tolerating the extra kernel
names = [r[col["Kernel Name"]] for r in data]
traffic = {}
with open(out + "_summary.csv", "w") as f:
    f.write("# ncu --set full --clock-control none, one launch per kernel group (bench.py --ncu-slice); cold-cache, "
            "serialised: compare shares, not absolutes\n")
    f.write("tag,kernel," + ",".join(k for k, _ in want) + "\n")
    ti = 0
    for r, name in zip(data, names):
        tag = tags[ti] if tags and ti < len(tags) else ""
        if tags and ti + 1 < len(tags) and not name.startswith("unpack_linear_dln"):
            ti += 1
        vals = []
        for k, m in want:
            v = float(r[col[m]].replace(",", "")) if m in col and r[col[m]] else float("nan")
            u = units[col[m]] if m in col else ""
            if k.endswith("_MB"):
                v = v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
            if k == "time_us":
                v = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)
            vals.append(v)
        f.write('"%s","%s",' % (tag, name[:60].replace('"', "'")) + ",".join("%.3f" % v for v in vals) + "\n")
        if tag and tag not in traffic:
            traffic[tag] = (vals[1] + vals[2]) * 1e6
json.dump({"source": out.split("/")[-1] + "_summary.csv", "bytes_per_launch": traffic},
          open("profiles/ncu_traffic.json", "w"), indent=1)
print("wrote", out + "_summary.csv", "and profiles/ncu_traffic.json (%d tags)" % len(traffic))
