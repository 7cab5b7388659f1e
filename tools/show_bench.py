"""Print a compact summary of bench.py JSON lines: python tools/show_bench.py file.log [...]"""
import json
import sys

for path in sys.argv[1:]:
    lines = [x for x in open(path) if x.startswith("{")]
    if not lines:
        print(path, "NO JSON LINE")
        continue
    d = json.loads(lines[-1])
    r = d.get("roofline") or {}
    print("%s: %s = %.2f %s  (%.3f ms/step, e2e %.2f)  top=%s %s frac=%.3f whole=%.3f" % (
        path, d["metric"], d["value"], d["unit"], d["ms_per_step"] or 0, d["e2e"]["value"], r.get("kernel"),
        r.get("bound"), r.get("frac", 0), (r.get("whole_step") or {}).get("frac", 0)))
    for k, v in list((r.get("groups") or {}).items())[:int(10)]:
        print("     %-44s n=%-3d %8.3f ms %7.1f TF %7.0f GB/s" % (k, v["n"], v["ms"], v["tflops"], v.get("gbs", 0)))
