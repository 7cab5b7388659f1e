"""TEST INFRASTRUCTURE: deterministic synthetic SPECT frame pairs shared by tests/golden/make_spect_golden.py (reference
side) and the SPECT tests (oracle / CUDA side), so the fixture need not store whole 2 x 1024 x 256 frames."""
import numpy as np


def synth_frames(n, seed, fh=1024, fw=256):
    """Poisson counts around a smooth 2-view activity map: H at full dose, L at 1/8 dose (scaled back)"""
    rng = np.random.RandomState(seed)
    yy, xx = np.meshgrid(np.arange(fh), np.arange(fw), indexing="ij")
    H, L = [], []
    for i in range(n):
        lam = []
        for v in range(2):
            cy, cx = fh * (0.3 + 0.4 * rng.rand()), fw * (0.3 + 0.4 * rng.rand())
            a = 40 * np.exp(-(((yy - cy) / (fh * 0.25)) ** 2 + ((xx - cx) / (fw * 0.3)) ** 2)) + 2.0
            a += 60 * np.exp(-(((yy - cy * 0.6) / 30.0) ** 2 + ((xx - cx * 1.1) / 20.0) ** 2))
            lam.append(a)
        lam = np.stack(lam)
        H.append(rng.poisson(lam).astype(np.float32))
        L.append((rng.poisson(lam / 8.0) * 8.0).astype(np.float32))
    return np.stack(L), np.stack(H)
