"""TEST INFRASTRUCTURE: import the unmodified reference modules from /root/reference (CPU, read-only).

`timm` is an un-vendored, unpinned dependency of models/network_swinir.py:11 that is not installed here; the three
symbols it supplies carry no hot-path arithmetic and are shimmed per SURVEY.md Appendix D.  /root/reference does not
exist on the GPU box: every caller must guard with `have_reference()`.
"""
import os
import sys
import types

import torch

REF = "/root/reference"


def have_reference():
    return os.path.isdir(os.path.join(REF, "models"))


def _install_timm_shim():
    if "timm.layers" in sys.modules:
        return

    def to_2tuple(x):
        return tuple(x) if isinstance(x, (tuple, list)) else (x, x)

    class DropPath(torch.nn.Module):
        """per-sample stochastic depth, floor(keep + U[0,1)) form (reference's own copy: models/network_vrt.py:181-205)"""

        def __init__(self, drop_prob=0.0):
            super().__init__()
            self.drop_prob = drop_prob

        def forward(self, x):
            if self.drop_prob == 0.0 or not self.training:
                return x
            keep = 1.0 - self.drop_prob
            mask = torch.floor(keep + torch.rand((x.shape[0],) + (1,) * (x.ndim - 1), dtype=x.dtype, device=x.device))
            return x / keep * mask

    timm = types.ModuleType("timm")
    layers = types.ModuleType("timm.layers")
    layers.to_2tuple, layers.DropPath = to_2tuple, DropPath
    layers.trunc_normal_ = torch.nn.init.trunc_normal_
    timm.layers = layers
    sys.modules["timm"], sys.modules["timm.layers"] = timm, layers


def load(name):
    """import `models.<name>` from the reference tree"""
    if not have_reference():
        raise RuntimeError("reference tree not mounted")
    sys.dont_write_bytecode = True
    _install_timm_shim()
    if REF not in sys.path:
        sys.path.append(REF)
    import importlib
    return importlib.import_module("models." + name)


def _stub(name, **attrs):
    if name not in sys.modules:
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
    return sys.modules[name]


def load_utils():
    """(utils.utils_image, utils.utils_spect, data.dataset_spect) of the reference.  utils_image.py:10-13 imports
    matplotlib / mpl_toolkits / lpips at module level (plotting and the LPIPS metric, neither on the path tested here);
    they are not installed in this image and are stubbed with empty modules -- no arithmetic lives in them."""
    if not have_reference():
        raise RuntimeError("reference tree not mounted")
    sys.dont_write_bytecode = True
    try:
        import matplotlib  # noqa: F401
    except ImportError:
        mpl = _stub("matplotlib")
        mpl.pyplot = _stub("matplotlib.pyplot")
        tk = _stub("mpl_toolkits")
        tk.mplot3d = _stub("mpl_toolkits.mplot3d", Axes3D=object)
    try:
        import lpips  # noqa: F401
    except ImportError:
        _stub("lpips")
    if REF not in sys.path:
        sys.path.append(REF)
    import importlib
    return (importlib.import_module("utils.utils_image"), importlib.import_module("utils.utils_spect"),
            importlib.import_module("data.dataset_spect"))
