"""ORACLE (test infrastructure, not product code): CPU restatement of the reference's DnCNN and UNetRes (DRUNet).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
Functional PyTorch on plain state_dict tensors, following /root/reference models/network_dncnn.py:40-71,
models/network_unet.py:21-76 and models/basicblock.py:61-98, 211-223, 471-501 (file:line cited per function).
Pinned against the real reference modules by tests/test_oracle_vs_reference.py (where /root/reference is mounted)
and against tests/golden/*.npz (generated from the real reference by tests/golden/make_golden.py).
"""
from collections import OrderedDict

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------------ DnCNN

def dncnn_state_spec(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="BR"):
    """Ordered [(key, shape, dtype)] of DnCNN.state_dict() (network_dncnn.py:61-67; B.sequential flattens the
    per-layer Sequentials, basicblock.py:15-35, so indices count conv / BN / ReLU modules)."""
    spec = []
    idx = 0

    def conv(i, cout, cin):
        spec.append(("model.%d.weight" % i, (cout, cin, 3, 3), torch.float32))
        spec.append(("model.%d.bias" % i, (cout,), torch.float32))

    def bn(i, c):
        for k in ("weight", "bias", "running_mean", "running_var"):
            spec.append(("model.%d.%s" % (i, k), (c,), torch.float32))
        spec.append(("model.%d.num_batches_tracked" % i, (), torch.int64))

    conv(idx, nc, in_nc)
    idx += 2                      # head: conv + activation
    for _ in range(nb - 2):
        conv(idx, nc, nc)
        idx += 1
        if "B" in act_mode:
            bn(idx, nc)
            idx += 1
        idx += 1                  # activation
    conv(idx, out_nc, nc)
    return spec


def _act(x, act_mode, slope=0.2):
    """last letter of act_mode: 'R' ReLU, 'L' LeakyReLU(0.2) (basicblock.py:70-77)"""
    return F.relu(x) if act_mode[-1] == "R" else F.leaky_relu(x, slope)


def dncnn_forward(sd, x, nb=17, act_mode="BR", training=False, momentum=0.9, eps=1e-4, new_stats=None, residual=True):
    """DnCNN.forward (network_dncnn.py:69-71): x - model(x) (residual=False: FDnCNN.forward, :146-148, model(x));
    body layers conv -> [BN(eps 1e-4, momentum .9)] -> act
    (basicblock.py:65-71).  With training=True BatchNorm uses biased batch statistics; `new_stats` (dict) receives the
    running-stat updates PyTorch would make (running = (1-m)*running + m*batch, unbiased variance)."""
    has_bn = "B" in act_mode
    idx = 0
    h = F.conv2d(x, sd["model.0.weight"], sd["model.0.bias"], padding=1)
    h = _act(h, act_mode)
    idx = 2
    for _ in range(nb - 2):
        h = F.conv2d(h, sd["model.%d.weight" % idx], sd["model.%d.bias" % idx], padding=1)
        idx += 1
        if has_bn and ("model.%d.running_mean" % idx) in sd:
            p = "model.%d." % idx
            if training:
                mean = h.mean(dim=(0, 2, 3))
                var = h.var(dim=(0, 2, 3), unbiased=False)
                if new_stats is not None:
                    n = h.numel() // h.shape[1]
                    new_stats[p + "running_mean"] = (1 - momentum) * sd[p + "running_mean"].detach() + momentum * mean.detach()
                    new_stats[p + "running_var"] = (1 - momentum) * sd[p + "running_var"].detach() + \
                        momentum * var.detach() * n / max(n - 1, 1)
            else:
                mean, var = sd[p + "running_mean"], sd[p + "running_var"]
            h = (h - mean.view(1, -1, 1, 1)) * torch.rsqrt(var.view(1, -1, 1, 1) + eps)
            h = h * sd[p + "weight"].view(1, -1, 1, 1) + sd[p + "bias"].view(1, -1, 1, 1)
            idx += 1
        elif has_bn:
            idx += 1              # BN child deleted by merge_bn; indices are not renumbered (utils_bnorm.py:61)
        h = _act(h, act_mode)
        idx += 1
    n = F.conv2d(h, sd["model.%d.weight" % idx], sd["model.%d.bias" % idx], padding=1)
    return x - n if residual else n


# ------------------------------------------------------------------------------------------------ UNetRes (DRUNet)

def unetres_state_spec(in_nc=3, out_nc=3, nc=(64, 128, 256, 512), nb=4, bias=False):
    """Ordered [(key, shape, dtype)] of UNetRes.state_dict() (network_unet.py:25-57)."""
    spec = []

    def add(prefix, shape):
        spec.append((prefix + ".weight", tuple(shape), torch.float32))
        if bias:
            spec.append((prefix + ".bias", (shape[0],), torch.float32))

    def res(prefix, c):
        add(prefix + ".res.0", (c, c, 3, 3))
        add(prefix + ".res.2", (c, c, 3, 3))

    add("m_head", (nc[0], in_nc, 3, 3))
    for lvl in range(3):
        for j in range(nb):
            res("m_down%d.%d" % (lvl + 1, j), nc[lvl])
        add("m_down%d.%d" % (lvl + 1, nb), (nc[lvl + 1], nc[lvl], 2, 2))
    for j in range(nb):
        res("m_body.%d" % j, nc[3])
    for lvl in (3, 2, 1):
        # ConvTranspose2d weight is stored (in, out, kh, kw) (basicblock.py:67)
        spec.append(("m_up%d.0.weight" % lvl, (nc[lvl], nc[lvl - 1], 2, 2), torch.float32))
        if bias:
            spec.append(("m_up%d.0.bias" % lvl, (nc[lvl - 1],), torch.float32))
        for j in range(nb):
            res("m_up%d.%d" % (lvl, j + 1), nc[lvl - 1])
    add("m_tail", (out_nc, nc[0], 3, 3))
    return spec


def _b(sd, k):
    return sd.get(k + ".bias")


def _resblock(x, sd, p):
    """ResBlock 'CRC' (basicblock.py:211-223): x + conv(relu(conv(x)))."""
    h = F.relu(F.conv2d(x, sd[p + ".res.0.weight"], _b(sd, p + ".res.0"), padding=1))
    return x + F.conv2d(h, sd[p + ".res.2.weight"], _b(sd, p + ".res.2"), padding=1)


def unetres_forward(sd, x0, nb=4):
    """UNetRes.forward (network_unet.py:59-76): strided 2x2 conv down (basicblock.py:495-501), 2x2 transposed conv up
    (:471-477), skip additions before every up stage and before the tail."""
    x1 = F.conv2d(x0, sd["m_head.weight"], _b(sd, "m_head"), padding=1)
    feats = [x1]
    x = x1
    for lvl in (1, 2, 3):
        for j in range(nb):
            x = _resblock(x, sd, "m_down%d.%d" % (lvl, j))
        x = F.conv2d(x, sd["m_down%d.%d.weight" % (lvl, nb)], _b(sd, "m_down%d.%d" % (lvl, nb)), stride=2)
        feats.append(x)
    for j in range(nb):
        x = _resblock(x, sd, "m_body.%d" % j)
    for lvl in (3, 2, 1):
        x = x + feats[lvl]
        x = F.conv_transpose2d(x, sd["m_up%d.0.weight" % lvl], _b(sd, "m_up%d.0" % lvl), stride=2)
        for j in range(nb):
            x = _resblock(x, sd, "m_up%d.%d" % (lvl, j + 1))
    return F.conv2d(x + feats[0], sd["m_tail.weight"], _b(sd, "m_tail"), padding=1)


def merge_bn_state(sd, nb=17, eps=1e-4):
    """utils/utils_bnorm.py:32-63 on a DnCNN 'BR' state_dict: fold BN running stats + affine into the preceding conv;
    returns the state_dict of the merged model (BN entries removed, indices kept)."""
    keys = list(sd.keys())
    out = OrderedDict()
    for k in keys:
        parts = k.split(".")
        i = int(parts[1])
        is_bn = ("model.%d.running_mean" % i) in sd
        if is_bn:
            if parts[2] == "running_mean":
                p = "model.%d." % i
                cw, cb = "model.%d.weight" % (i - 1), "model.%d.bias" % (i - 1)
                invstd = (sd[p + "running_var"] + eps).pow(-0.5)
                w = out[cw] * invstd.view(-1, 1, 1, 1)
                b = (out[cb] - sd[p + "running_mean"]) * invstd
                out[cw] = w * sd[p + "weight"].view(-1, 1, 1, 1)
                out[cb] = b * sd[p + "weight"] + sd[p + "bias"]
            continue
        out[k] = sd[k].clone()
    return out


# --------------------------------------------------------------------------------------------- RRDBNet_noUp

def rrdbnet_state_spec(in_nc=3, out_nc=3, nf=64, nb=23, gc=32):
    """Ordered [(key, shape, dtype)] of RRDBNet_noUp.state_dict() (models/network_rrdbnet.py:106-118)."""
    spec = []

    def conv(name, o, i):
        spec.append((name + ".weight", (o, i, 3, 3), torch.float32))
        spec.append((name + ".bias", (o,), torch.float32))
    conv("conv_first", nf, in_nc)
    for b in range(nb):
        for r in (1, 2, 3):
            p = "RRDB_trunk.%d.RDB%d." % (b, r)
            for k in range(4):
                conv(p + "conv%d" % (k + 1), gc, nf + k * gc)
            conv(p + "conv5", nf, nf + 4 * gc)
    conv("trunk_conv", nf, nf)
    conv("HRconv", nf, nf)
    conv("conv_last", out_nc, nf)
    return spec


def _rdb(x, sd, p):
    """ResidualDenseBlock_5C.forward, models/network_rrdbnet.py:49-56."""
    import torch.nn.functional as F
    c = lambda name, t: F.conv2d(t, sd[p + name + ".weight"], sd[p + name + ".bias"], padding=1)
    l = lambda t: F.leaky_relu(t, 0.2)
    x1 = l(c("conv1", x))
    x2 = l(c("conv2", torch.cat((x, x1), 1)))
    x3 = l(c("conv3", torch.cat((x, x1, x2), 1)))
    x4 = l(c("conv4", torch.cat((x, x1, x2, x3), 1)))
    x5 = c("conv5", torch.cat((x, x1, x2, x3, x4), 1))
    return x5 * 0.2 + x


def rrdbnet_noup_forward(sd, x, nb=23):
    """RRDBNet_noUp.forward, models/network_rrdbnet.py:120-125 (RRDB.forward :67-72)."""
    import torch.nn.functional as F
    c = lambda name, t: F.conv2d(t, sd[name + ".weight"], sd[name + ".bias"], padding=1)
    fea = c("conv_first", x)
    t = fea
    for b in range(nb):
        out = t
        for r in (1, 2, 3):
            out = _rdb(out, sd, "RRDB_trunk.%d.RDB%d." % (b, r))
        t = out * 0.2 + t
    fea = fea + c("trunk_conv", t)
    return c("conv_last", F.leaky_relu(c("HRconv", fea), 0.2))


# --------------------------------------------------------------------------------------------- IRCNN

def ircnn_state_spec(in_nc=1, out_nc=1, nc=64):
    """Ordered [(key, shape, dtype)] of IRCNN.state_dict() (models/network_dncnn.py:97-112: model.0, .2, .., .12)."""
    spec = []
    chans = [(nc, in_nc)] + [(nc, nc)] * 5 + [(out_nc, nc)]
    for i, (o, c) in enumerate(chans):
        spec.append(("model.%d.weight" % (2 * i), (o, c, 3, 3), torch.float32))
        spec.append(("model.%d.bias" % (2 * i), (o,), torch.float32))
    return spec


def ircnn_forward(sd, x):
    """IRCNN.forward, models/network_dncnn.py:114-116 with the layers of :99-111 (dilations 1,2,3,4,3,2,1)."""
    h = x
    for i, d in enumerate((1, 2, 3, 4, 3, 2, 1)):
        h = F.conv2d(h, sd["model.%d.weight" % (2 * i)], sd["model.%d.bias" % (2 * i)], padding=d, dilation=d)
        if i < 6:
            h = F.relu(h)
    return x - h


# --------------------------------------------------------------------------------------------- FFDNet

def ffdnet_state_spec(in_nc=1, out_nc=1, nc=64, nb=15, act_mode="R"):
    """Ordered [(key, shape, dtype)] of FFDNet.state_dict() (models/network_ffdnet.py:45-49: model.0, .2, ..)."""
    spec = []
    chans = [(nc, in_nc * 4 + 1)] + [(nc, nc)] * (nb - 2) + [(out_nc * 4, nc)]
    for i, (o, c) in enumerate(chans):
        spec.append(("model.%d.weight" % (2 * i), (o, c, 3, 3), torch.float32))
        spec.append(("model.%d.bias" % (2 * i), (o,), torch.float32))
    return spec


def ffdnet_forward(sd, x, sigma, nb=15):
    """FFDNet.forward, models/network_ffdnet.py:54-69 (ReLU stages; PixelUnShuffle of basicblock.py:104-126)."""
    h, w = x.shape[-2:]
    x = F.pad(x, (0, int(-(-w // 2) * 2 - w), 0, int(-(-h // 2) * 2 - h)), mode="replicate")
    x = F.pixel_unshuffle(x, 2)
    m = sigma.repeat(1, 1, x.shape[-2], x.shape[-1])
    t = torch.cat((x, m), 1)
    for i in range(nb):
        t = F.conv2d(t, sd["model.%d.weight" % (2 * i)], sd["model.%d.bias" % (2 * i)], padding=1)
        if i < nb - 1:
            t = F.relu(t)
    return F.pixel_shuffle(t, 2)[..., :h, :w]
