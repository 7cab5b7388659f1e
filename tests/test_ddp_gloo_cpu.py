"""N > 1 path on CPU: world_size-2 `gloo` process group, one process per "GPU", DistributedDataParallel around the
drop-in modules (launch lists replayed through tests/emu_ops.py), per-rank half batches.  The averaged gradients must
equal the single-process gradients of the oracle on the full batch (models/model_base.py:113-119 semantics), and a
step of the same optimizer must leave both ranks with identical parameters."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _build(net):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import emu_ops
    from test_convnet_engine_cpu import _seed_params
    torch.manual_seed(0)
    if net == "dncnn":
        from kair_b200.network_dncnn import DnCNN
        model = DnCNN(in_nc=1, out_nc=1, nc=64, nb=4, act_mode="R")
        _seed_params(model, 7)
        shape, out_c = (4, 1, 16, 16), 1
    else:
        from kair_b200.network_swinir import SwinIR
        model = SwinIR(upscale=1, in_chans=1, img_size=16, window_size=8, img_range=1.0, depths=[2, 2], embed_dim=180,
                       num_heads=[6, 6], mlp_ratio=2, upsampler="", resi_connection="1conv", drop_path_rate=0.0)
        shape, out_c = (2, 1, 16, 16), 1
    model._ops = emu_ops
    g = torch.Generator().manual_seed(11)
    x, tgt = torch.rand(*shape, generator=g), torch.rand(shape[0], out_c, *shape[2:], generator=g)
    return model, x, tgt


def _worker(rank, world, port, net, out_dir, flat):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    model, x, tgt = _build(net)
    if flat:
        from kair_b200.parallel import FlatDataParallel
        ddp = FlatDataParallel(model)
    else:
        ddp = torch.nn.parallel.DistributedDataParallel(model, broadcast_buffers=False)
    n = x.shape[0] // world
    xs, ts = x[rank * n:(rank + 1) * n], tgt[rank * n:(rank + 1) * n]
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    loss = torch.mean((ddp(xs) - ts) ** 2)
    loss.backward()
    grads = {k: p.grad.clone() for k, p in model.named_parameters()}
    opt.step()
    torch.save({"grads": grads, "params": {k: p.detach().clone() for k, p in model.named_parameters()}},
               os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("net,flat", [("dncnn", False), ("swinir", False), ("dncnn", True), ("swinir", True)])
def test_ddp_world2_matches_full_batch(net, flat, tmp_path):
    """flat=False: torch DistributedDataParallel (the reference's wrapper); flat=True: kair_b200.parallel
    .FlatDataParallel (one all-reduce of the engine's flat gradient buffer)."""
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), net, str(tmp_path), flat), nprocs=world, join=True)
    r0, r1 = (torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r)) for r in range(world))
    for k in r0["grads"]:
        assert torch.equal(r0["grads"][k], r1["grads"][k]), k          # all-reduced: identical on both ranks
        assert torch.equal(r0["params"][k], r1["params"][k]), k
    # single process, full batch, same launch lists
    model, x, tgt = _build(net)
    torch.mean((model(x) - tgt) ** 2).backward()
    gmax = max(p.grad.norm().item() for p in model.parameters())
    for k, p in model.named_parameters():
        err = (r0["grads"][k] - p.grad).norm().item()
        assert err < 2e-2 * p.grad.norm().item() + 1e-3 * gmax, (k, err, p.grad.norm().item())
