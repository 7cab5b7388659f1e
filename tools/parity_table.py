"""Measure full-architecture parity of the CUDA path against the CPU oracle at the BASELINE configs and write the
table the tolerances in tests/test_parity_full_gpu.py are derived from (profiles/parity_r2.json).

    python tools/parity_table.py [--precision bf16|fp32|both] [--configs cfg2,cfg3,cfg4,cfg5] [--out PATH]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--configs", default="cfg2,cfg3,cfg4,cfg5")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "parity_r2.json"))
    a = ap.parse_args()
    import torch
    import parity_full as pf
    precisions = ["bf16", "fp32"] if a.precision == "both" else [a.precision]
    table = {"device": torch.cuda.get_device_name(0), "torch_threads": torch.get_num_threads(), "rows": []}
    for prec in precisions:
        for c in a.configs.split(","):
            try:
                row, per_tensor = pf.MEASURES[c](precision=prec)
            except NotImplementedError as e:
                row, per_tensor = {"config": c, "precision": prec, "error": str(e)}, {}
            if per_tensor:
                top = sorted(per_tensor.items(), key=lambda kv: -kv[1][0])[:8]
                row["worst_tensors"] = [{"name": k, "rel_l2": v[0], "max_rel": v[1], "cos": v[2]} for k, v in top]
            row["id"] = c
            table["rows"].append(row)
            print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as fh:
        json.dump(table, fh, indent=1)


if __name__ == "__main__":
    main()
