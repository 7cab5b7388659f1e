"""kair_b200: sm_100a (B200) kernels behind KAIR's SwinIR / DnCNN / FDnCNN / UNetRes / RRDBNet_noUp module API (see
DESIGN.md)."""
import importlib
import sys

__version__ = "0.2.0"
__all__ = ["install"]

_TARGETS = (("models.network_swinir", "SwinIR", "kair_b200.network_swinir"),
            ("models.network_dncnn", "DnCNN", "kair_b200.network_dncnn"),
            ("models.network_dncnn", "FDnCNN", "kair_b200.network_dncnn"),
            ("models.network_dncnn", "IRCNN", "kair_b200.network_dncnn"),
            ("models.network_unet", "UNetRes", "kair_b200.network_unet"),
            ("models.network_rrdbnet", "RRDBNet_noUp", "kair_b200.network_rrdbnet"),
            ("models.network_ffdnet", "FFDNet", "kair_b200.network_ffdnet"))


def install():
    """Route KAIR's network classes through the sm_100a engines without editing any reference file.

    `models/select_network.py:16-274` (define_G) and `main_test_swinir.py:127-174` / `main_test_dncnn.py:117`
    look the classes up as attributes of `models.network_swinir / network_dncnn / network_unet` at call time, so
    replacing those attributes (or, when the reference module cannot be imported -- `network_swinir.py:11` needs the
    un-vendored `timm` --, registering ours under its name in `sys.modules`) is all it takes.  Call it once, with the
    KAIR checkout on `sys.path`, before the model is built.  Returns {module name: "patched" | "registered"}."""
    done = {}
    # the fused Charbonnier loss: model_plain.py imports the class by name from models.loss, so it must
    # be swapped before model_plain is imported -- hence "call install() first"
    try:
        ref_loss = importlib.import_module("models.loss")
        from .loss import CharbonnierLoss, PoissonLLLoss
        ref_loss.CharbonnierLoss = CharbonnierLoss
        if hasattr(ref_loss, "PoissonLLLoss"):      # the SPECT fork's loss (models/loss.py:294-345)
            ref_loss.PoissonLLLoss = PoissonLLLoss
        done["models.loss"] = "patched"
    except Exception:  # noqa: BLE001 - models.loss needs torchvision / other optional packages in some forks
        done["models.loss"] = "skipped"
    for ref_name, cls, mine_name in _TARGETS:
        mine = importlib.import_module(mine_name)
        try:
            ref = importlib.import_module(ref_name)
            setattr(ref, cls, getattr(mine, cls))
            done[ref_name] = "patched"
        except Exception:  # noqa: BLE001 - the reference module is absent or lacks a dependency
            sys.modules[ref_name] = mine
            pkg = sys.modules.get(ref_name.rsplit(".", 1)[0])
            if pkg is not None:
                setattr(pkg, ref_name.rsplit(".", 1)[1], mine)
            done[ref_name] = "registered"
    return done
