"""Generate tests/golden/kat/dncnn_25.npz: the known-answer test of SURVEY.md section 8(c), made with the UNMODIFIED
reference module (models/network_dncnn.DnCNN, act_mode='R') and the only checkpoint the reference ships
(model_zoo/dncnn_25.pth, sha256 0451a70d...58e5b0), on torch.manual_seed(0); x = rand(1,1,256,256) (cfg1 shape).

    python tests/golden/make_dncnn25_kat.py

The checkpoint's tensors are stored with the fixture because /root/reference does not exist on the GPU box; they are
data (a weight file), not reference source code."""
import hashlib
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

if __name__ == "__main__":
    path = os.path.join(ref_loader.REF, "model_zoo", "dncnn_25.pth")
    sha = hashlib.sha256(open(path, "rb").read()).hexdigest()
    sd = torch.load(path, map_location="cpu")
    model = ref_loader.load("network_dncnn").DnCNN(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="R")
    model.load_state_dict(sd, strict=True)
    model.eval()
    torch.manual_seed(0)
    x = torch.rand(1, 1, 256, 256)
    with torch.no_grad():
        y = model(x)
    assert abs(y.mean().item() - 0.5027428) < 1e-5 and abs(y[0, 0, 128, 128].item() - 0.7606996) < 1e-5
    # cfg1 denoising protocol: clean synthetic image + sigma=25 noise (np.random.seed(0)), PSNR on uint8
    clean, noisy = gc.dncnn_cfg1_images()
    with torch.no_grad():
        y_dn = model(noisy)
    p_in, p_out = gc.psnr_uint8(noisy, clean), gc.psnr_uint8(y_dn, clean)
    print("cfg1 denoise: %.2f dB -> %.2f dB" % (p_in, p_out))
    os.makedirs(os.path.join(HERE, "kat"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "kat", "dncnn_25.npz"), y=y.numpy(), sha256=sha, y_denoise=y_dn.numpy(),
                        psnr_in=np.float64(p_in), psnr_out=np.float64(p_out),
                        keys=np.array(list(sd.keys())), **{"w:" + k: v.numpy() for k, v in sd.items()})
    print("wrote kat/dncnn_25.npz  sha256", sha[:8], " y.mean %.7f" % y.mean().item())
