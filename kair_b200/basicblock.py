"""Module-tree builders for the conv nets, mirroring the subset of models/basicblock.py the hot path uses:
`sequential` (:15-35), `conv` (:61-98), `ResBlock` (:211-223), `upsample_convtranspose` (:471-477) and
`downsample_strideconv` (:495-501).

The trees are made of stock `nn.Conv2d / nn.ConvTranspose2d / nn.BatchNorm2d / nn.ReLU / nn.LeakyReLU` children in
the reference's order, so `state_dict()` keys, `init_weights` (class-name based, models/select_network.py:385-433),
`merge_bn` / `tidy_sequential` (utils/utils_bnorm.py:32-91) and the regularizers keep working.  The arithmetic is not
done by these children: `DnCNN.forward` / `UNetRes.forward` hand the whole tree to the sm_100a engines.
"""
import torch.nn as nn


def sequential(*args):
    """Flatten nested nn.Sequential arguments into one nn.Sequential (a single module is returned as is)."""
    if len(args) == 1:
        if isinstance(args[0], dict):
            raise NotImplementedError("sequential does not support OrderedDict input.")
        return args[0]
    flat = []
    for m in args:
        if isinstance(m, nn.Sequential):
            flat.extend(m.children())
        elif isinstance(m, nn.Module):
            flat.append(m)
    return nn.Sequential(*flat)


_UNSUPPORTED = "I234UuvMA"


def conv(in_channels=64, out_channels=64, kernel_size=3, stride=1, padding=1, bias=True, mode="CBR",
         negative_slope=0.2):
    """One module per letter of `mode`: C conv, T transposed conv, B BatchNorm2d(momentum=0.9, eps=1e-4), R/r ReLU,
    L/l LeakyReLU(negative_slope)."""
    layers = []
    for t in mode:
        if t == "C":
            layers.append(nn.Conv2d(in_channels, out_channels, kernel_size, stride, padding, bias=bias))
        elif t == "T":
            layers.append(nn.ConvTranspose2d(in_channels, out_channels, kernel_size, stride, padding, bias=bias))
        elif t == "B":
            layers.append(nn.BatchNorm2d(out_channels, momentum=0.9, eps=1e-04, affine=True))
        elif t in "Rr":
            layers.append(nn.ReLU(inplace=(t == "R")))
        elif t in "Ll":
            layers.append(nn.LeakyReLU(negative_slope=negative_slope, inplace=(t == "L")))
        elif t in _UNSUPPORTED:
            raise NotImplementedError("kair_b200.basicblock.conv: mode letter '%s' is outside the DnCNN / DRUNet hot "
                                      "path and has no sm_100a kernel" % t)
        else:
            raise NotImplementedError("Undefined type: %s" % t)
    return sequential(*layers)


class ResBlock(nn.Module):
    """x + conv(act(conv(x))) -- parameter container; evaluated by the UNetRes engine."""

    def __init__(self, in_channels=64, out_channels=64, kernel_size=3, stride=1, padding=1, bias=True, mode="CRC",
                 negative_slope=0.2):
        super().__init__()
        assert in_channels == out_channels, "Only support in_channels==out_channels."
        if mode[0] in ("R", "L"):
            mode = mode[0].lower() + mode[1:]
        self.res = conv(in_channels, out_channels, kernel_size, stride, padding, bias, mode, negative_slope)

    def forward(self, x):
        raise RuntimeError("kair_b200: ResBlock is a parameter container; the computation is fused into UNetRes.forward")


def upsample_convtranspose(in_channels=64, out_channels=3, kernel_size=2, stride=2, padding=0, bias=True, mode="2R",
                           negative_slope=0.2):
    assert len(mode) < 4 and mode[0] in "234", "mode examples: 2, 2R, 2BR, 3, ..., 4BR."
    k = int(mode[0])
    return conv(in_channels, out_channels, k, k, padding, bias, mode.replace(mode[0], "T"), negative_slope)


def downsample_strideconv(in_channels=64, out_channels=64, kernel_size=2, stride=2, padding=0, bias=True, mode="2R",
                          negative_slope=0.2):
    assert len(mode) < 4 and mode[0] in "234", "mode examples: 2, 2R, 2BR, 3, ..., 4BR."
    k = int(mode[0])
    return conv(in_channels, out_channels, k, k, padding, bias, mode.replace(mode[0], "C"), negative_slope)
