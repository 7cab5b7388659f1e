"""Generate tests/golden/kat/spect_golden.npz from the UNMODIFIED reference functions (run where /root/reference is
mounted): data/dataset_spect.py (DatasetSPECT on two temporary .dat frame pairs, seeded `random`), utils/utils_spect.py,
utils/utils_image.py (calculate_psnr, calculate_ssim, pll) and the quantisation of models/model_base.py:358-366.

    python tests/golden/make_spect_golden.py
"""
import os
import random
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ref_loader  # noqa: E402


from spect_cases import synth_frames  # noqa: E402


def main():
    util, uspect, dset = ref_loader.load_utils()
    L, H = synth_frames(2, 7, 256, 128)     # small frames keep the fixture small; the code path has no size constants
    out = {"frames_L": L, "frames_H": H}
    with tempfile.TemporaryDirectory() as td:
        dl, dh = os.path.join(td, "L"), os.path.join(td, "H")
        os.makedirs(dl); os.makedirs(dh)
        # DatasetSPECT reshapes to (2, 1024, 256): pad the small frames into that shape for the file round trip
        Lf, Hf = synth_frames(2, 7, 1024, 256)
        for i in range(2):
            Lf[i].tofile(os.path.join(dl, "f%d.dat" % i)); Hf[i].tofile(os.path.join(dh, "f%d.dat" % i))
        # util.get_image_paths only lists image extensions: give the dataset its path lists directly
        for method in ("log", "anscombe", "linear"):
            opt = {"n_channels": 2, "H_size": 64, "normalization": {"type": method, "max_pixel": 150}, "phase": "train",
                   "dataroot_H": dh, "dataroot_L": dl}
            orig = util.get_image_paths
            util.get_image_paths = lambda d: sorted(os.path.join(d, f) for f in os.listdir(d))
            try:
                ds = dset.DatasetSPECT(opt)
            finally:
                util.get_image_paths = orig
            random.seed(1234)
            draws, pl, ph = [], [], []
            for idx in (0, 1, 1, 0, 1, 0, 0, 1):
                state = random.getstate()
                item = ds[idx]
                random.setstate(state)
                top = random.randint(0, 1024 - 64); left = random.randint(0, 256 - 64); mode = random.randint(0, 7)
                draws.append((idx, top, left, mode))
                pl.append(item["L"].numpy()); ph.append(item["H"].numpy())
            out["draws_" + method] = np.array(draws, dtype=np.int32)
            out["patch_L_" + method], out["patch_H_" + method] = np.stack(pl), np.stack(ph)
            # the drawn windows, cut out of the 1024 x 256 frames, so the fixture need not store whole frames
            opt["phase"] = "test"
            util.get_image_paths = lambda d: sorted(os.path.join(d, f) for f in os.listdir(d))
            try:
                dt = dset.DatasetSPECT(opt)
            finally:
                util.get_image_paths = orig
            out["test_L_" + method] = dt[0]["L"].numpy()[:, :160, :96]
        out["full_frames_L"], out["full_frames_H"] = Lf, Hf
    # metrics on de-normalised image pairs
    rng = np.random.RandomState(3)
    Hc = H[0].transpose(1, 2, 0)                                   # (h, w, 2) counts
    Ec = np.maximum(Hc + rng.normal(0, 3.0, Hc.shape), -1.0).astype(np.float32)
    out["metric_E"], out["metric_H"] = Ec, Hc
    mv_g, mv_l = np.float32(Hc.max() * 1.3), np.float32(Hc.max())
    vals = []
    for mv in (mv_g, mv_l):
        for ch in range(2):
            e8 = (np.clip(Ec[:, :, ch], 0, mv) / mv * 255).astype(np.uint8)
            h8 = (np.clip(Hc[:, :, ch], 0, mv) / mv * 255).astype(np.uint8)
            E_rgb, H_rgb = np.stack([e8] * 3, axis=2), np.stack([h8] * 3, axis=2)
            vals.append((util.calculate_psnr(E_rgb, H_rgb), util.calculate_ssim(E_rgb, H_rgb)))
    out["metric_maxvals"] = np.array([mv_g, mv_l], dtype=np.float32)
    out["metric_psnr_ssim"] = np.array(vals, dtype=np.float64)     # rows: (global ch0, global ch1, local ch0, local ch1)
    out["metric_pll"] = np.array([util.pll(Ec, Hc)])
    x = np.linspace(0, 200, 97).astype(np.float32)
    for method in ("log", "anscombe", "linear"):
        n = uspect.normalize_spect(x, method, 150)
        out["norm_" + method] = np.asarray(n, dtype=np.float64)
        out["denorm_" + method] = np.asarray(uspect.denormalize_spect(np.asarray(n, dtype=np.float32), method, 150), dtype=np.float64)
    out["norm_x"] = x
    del out["full_frames_L"], out["full_frames_H"]
    np.savez_compressed(os.path.join(HERE, "kat", "spect_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "kat", "spect_golden.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
