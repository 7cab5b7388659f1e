"""Pins the oracle (oracle/*.py): against the committed golden vectors generated from the real reference modules
(tests/golden/*.npz, made by tests/golden/make_golden.py) and, where /root/reference is mounted, against the live
reference modules including the model_zoo/dncnn_25.pth known-answer values of SURVEY.md section 8(c).
fp32 on both sides: outputs within 2e-5 max-abs, gradient fingerprints within 1e-4 relative."""
import glob
import json
import os

import numpy as np
import pytest
import torch

import golden_cases as gc
import ref_loader
from oracle import convnets_oracle as co

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def test_every_case_has_a_fixture():
    assert set(NAMES) == set(gc.CASES), (NAMES, list(gc.CASES))


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    case = json.loads(str(z["case"]))
    assert case == json.loads(json.dumps(gc.CASES[name])), "fixture is stale: rerun tests/golden/make_golden.py"
    sd, x, tgt = gc.state_and_io(case)
    for v in sd.values():
        if v.dtype.is_floating_point:
            v.requires_grad_(True)
    stats = {}
    y = gc.oracle_forward(case, sd, x, new_stats=stats)
    loss = gc.loss_fn(case)(y, tgt)
    loss.backward()
    assert np.abs(y.detach().numpy() - z["y"]).max() < 2e-5
    assert abs(loss.item() - float(z["loss"])) < 1e-6
    grads = json.loads(str(z["grads"]))
    assert grads
    for k, fp in grads.items():
        g = sd[k].grad.detach().double().reshape(-1)
        assert abs(float(g.norm()) - fp[0]) <= 1e-4 * fp[0] + 1e-6, k
        assert np.allclose(g[:4].numpy(), np.array(fp[1:]), rtol=1e-3, atol=1e-4 * fp[0] + 1e-6), k
    for k in z.files:
        if k.startswith("stat:"):
            assert np.abs(stats[k[5:]].detach().numpy() - z[k]).max() < 1e-5, k


needs_ref = pytest.mark.skipif(not ref_loader.have_reference(), reason="/root/reference not mounted")


@needs_ref
@pytest.mark.parametrize("name", ["swinir_m_gray32", "dncnn_br_train", "unetres_nb2"])
def test_oracle_matches_live_reference(name):
    case = gc.CASES[name]
    model, sd, x, tgt = gc.build_reference(case, ref_loader)
    model.train(case.get("train_mode", False))
    y_ref = model(x)
    y = gc.oracle_forward(case, {k: v.clone() for k, v in sd.items()}, x)
    assert (y - y_ref).abs().max().item() < 2e-5


@needs_ref
def test_dncnn25_known_answer():
    """SURVEY 8(c) anchor: DnCNN(1,1,64,17,'R') + model_zoo/dncnn_25.pth on torch.manual_seed(0) rand(1,1,256,256)."""
    sd = torch.load(os.path.join(ref_loader.REF, "model_zoo", "dncnn_25.pth"), map_location="cpu")
    assert [k for k, _, _ in co.dncnn_state_spec(1, 1, 64, 17, "R")] == list(sd.keys())
    torch.manual_seed(0)
    x = torch.rand(1, 1, 256, 256)
    with torch.no_grad():
        y = co.dncnn_forward(sd, x, nb=17, act_mode="R")
    assert abs(y.mean().item() - 0.5027428) < 1e-5
    assert abs(y.std().item() - 0.2438776) < 1e-5
    assert torch.allclose(y[0, 0, 0, :4], torch.tensor([0.4639804, 0.7212471, 0.1741740, 0.1720521]), atol=1e-5)
    assert abs(y[0, 0, 128, 128].item() - 0.7606996) < 1e-5


@needs_ref
def test_merge_bn_restatement():
    """utils/utils_bnorm.py:32-63 on DnCNN 'BR': oracle's folded state == reference merge_bn result."""
    import importlib
    case = gc.CASES["dncnn_br_eval"]
    model, sd, x, _ = gc.build_reference(case, ref_loader)
    model.eval()
    with torch.no_grad():
        y0 = model(x)
    merged = co.merge_bn_state(sd)
    bn = importlib.import_module("utils.utils_bnorm")
    bn.merge_bn(model)
    ref_sd = model.state_dict()
    assert list(ref_sd.keys()) == list(merged.keys())
    for k in merged:
        assert torch.allclose(ref_sd[k], merged[k], atol=1e-6), k
    with torch.no_grad():
        y1 = co.dncnn_forward(merged, x, nb=case["kw"]["nb"], act_mode="BR")
    assert (y1 - y0).abs().max().item() < 1e-4
