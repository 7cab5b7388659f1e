"""Generate tests/golden/*.npz from the UNMODIFIED reference modules in /root/reference (CPU, fp32).

    python tests/golden/make_golden.py

Each fixture stores only what travels cheaply: the case description (constructor kwargs, seeds, input shape), the
reference output, the loss, and per-parameter gradient fingerprints (L2 norm + first 4 values).  Weights and inputs
are NOT stored: they are regenerated bit-identically from the seeds by oracle.*.deterministic_state / seeded CPU
generators (tests/golden_cases.py), so the fixtures pin  oracle(seeded weights, seeded input) == reference(same).
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import golden_cases as gc  # noqa: E402
import ref_loader  # noqa: E402


def fingerprint(t):
    t = t.detach().double().reshape(-1)
    return [float(t.norm())] + [float(v) for v in t[:4]]


def run_case(name, case):
    torch.manual_seed(0)
    model, sd, x, tgt = gc.build_reference(case, ref_loader)
    model.train(case.get("train_mode", False))
    if case.get("backward", True):
        y = model(x)
        loss = gc.loss_fn(case)(y, tgt)
        loss.backward()
        grads = {k: fingerprint(p.grad) for k, p in model.named_parameters() if p.grad is not None}
        lossv = float(loss)
    else:
        with torch.no_grad():
            y = model(x)
        grads, lossv = {}, float("nan")
    extra = {}
    if case.get("train_mode", False) and case["net"] == "dncnn":
        extra = {k: v.detach().numpy() for k, v in model.state_dict().items() if "running_" in k}
    np.savez_compressed(os.path.join(HERE, name + ".npz"), y=y.detach().numpy(), loss=np.float64(lossv),
                        grads=json.dumps(grads), case=json.dumps(case), **{"stat:" + k: v for k, v in extra.items()})
    print("%-28s y%s loss %.6f  %d grads" % (name, tuple(y.shape), lossv, len(grads)))


if __name__ == "__main__":
    if not ref_loader.have_reference():
        raise SystemExit("reference tree not mounted")
    torch.set_num_threads(8)
    only = sys.argv[1:]
    for name, case in gc.CASES.items():
        if not only or name in only:
            run_case(name, case)
