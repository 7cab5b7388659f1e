"""kair_pixel_loss_noalloc (include/kair_b200.h): the allocation-free variant of kair_pixel_loss, called through the
C ABI and checked against the existing entry point on the same inputs, twice in a row (the static scratch array keeps
the previous call's contents; the kernel re-arms it).  Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest
import torch

MAX_N = 8192 * 8192   # KAIR_PIXEL_LOSS_NOALLOC_MAX_N


def _args(x, y, dx, loss, scratch, mode, eps, max_pixel, n=None):
    from kair_b200 import _lib
    a = _lib.STRUCTS["kair_pixel_loss_args"]()
    a.x, a.y, a.loss = x.data_ptr(), y.data_ptr(), loss.data_ptr()
    a.dx = dx.data_ptr() if dx is not None else None
    a.scratch = scratch.data_ptr() if scratch is not None else None
    a.n, a.eps, a.mode, a.max_pixel = (x.numel() if n is None else n), eps, mode, max_pixel
    return a


def _call(name, a):
    from kair_b200 import _lib
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return getattr(_lib.load(), name)(ctypes.byref(a), stream)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(8, 1, 128, 128), (3, 3, 37, 45), (1, 1, 1, 1), (4, 3, 512, 512)])
@pytest.mark.parametrize("mode,eps,max_pixel", [(0, 1e-9, 0.0), (1, 0.0, 0.0), (3, 1e-8, 150.0), (5, 0.0, 0.0)])
def test_noalloc_matches_pixel_loss(shape, mode, eps, max_pixel):
    g = torch.Generator().manual_seed(11)
    x = torch.rand(shape, generator=g).cuda()
    y = torch.rand(shape, generator=g).cuda()
    y.view(-1)[0] = x.view(-1)[0]                      # an exact zero difference
    # reference: the existing entry point with caller-owned scratch
    dx_ref, loss_ref = torch.empty_like(x), torch.zeros(1, device="cuda")
    scratch = torch.zeros(2, dtype=torch.float64, device="cuda")
    assert _call("kair_pixel_loss", _args(x, y, dx_ref, loss_ref, scratch, mode, eps, max_pixel)) == 0
    torch.cuda.synchronize()
    # the allocation-free variant: every buffer exists before the calls, two calls in a row on the shared array
    dx, loss = torch.empty_like(x), torch.zeros(2, device="cuda")
    seen = []
    for it in range(2):
        dx.fill_(float("nan"))
        a = _args(x, y, dx, loss[it:], None, mode, eps, max_pixel)
        assert _call("kair_pixel_loss_noalloc", a) == 0
        torch.cuda.synchronize()
        seen.append(loss[it].item())
        ref = loss_ref.item()
        assert abs(seen[-1] - ref) < 1e-6 * max(1.0, abs(ref)), (it, seen[-1], ref)
        assert torch.allclose(dx, dx_ref, rtol=1e-5, atol=1e-9)
    assert seen[0] == seen[1]                          # fixed summation order: bit-identical from call to call


@pytest.mark.gpu
def test_noalloc_without_gradient_and_back_to_back():
    """dx = NULL, and launches queued back to back on one stream without a host sync in between (stream order serialises
    them, which is the documented requirement)."""
    g = torch.Generator().manual_seed(5)
    xs = [torch.rand(n, generator=g).cuda() for n in (100_000, 7, 300_001)]
    ys = [torch.rand(x.shape, generator=g).cuda() for x in xs]
    loss = torch.zeros(len(xs), device="cuda")
    for i, (x, y) in enumerate(zip(xs, ys)):
        assert _call("kair_pixel_loss_noalloc", _args(x, y, None, loss[i:], None, 0, 1e-9, 0.0)) == 0
    torch.cuda.synchronize()
    for i, (x, y) in enumerate(zip(xs, ys)):
        ref = torch.sqrt((x.double() - y.double()) ** 2 + 1e-9).mean().item()
        assert abs(loss[i].item() - ref) < 2e-6 * max(1.0, abs(ref))


@pytest.mark.gpu
def test_noalloc_rejects_what_it_cannot_serve():
    from kair_b200 import _lib
    x = torch.rand(16).cuda()
    y = torch.rand(16).cuda()
    loss = torch.zeros(1, device="cuda")
    # beyond the static array's limit: refused before anything is launched (the 16-element buffers are never touched)
    assert _call("kair_pixel_loss_noalloc", _args(x, y, None, loss, None, 0, 1e-9, 0.0, n=MAX_N + 1)) != 0
    msg = _lib.load().kair_last_error().decode()
    assert "exceeds" in msg and str(MAX_N) in msg
    # caller-owned scratch belongs to kair_pixel_loss
    scratch = torch.zeros(2, dtype=torch.float64, device="cuda")
    assert _call("kair_pixel_loss_noalloc", _args(x, y, None, loss, scratch, 0, 1e-9, 0.0)) != 0
    assert "scratch" in _lib.load().kair_last_error().decode()
    # and the entry point still works afterwards
    assert _call("kair_pixel_loss_noalloc", _args(x, y, None, loss, None, 1, 0.0, 0.0)) == 0
    torch.cuda.synchronize()
    assert abs(loss.item() - (x - y).abs().mean().item()) < 1e-6


@pytest.mark.gpu
def test_noalloc_at_the_limit():
    """n = KAIR_PIXEL_LOSS_NOALLOC_MAX_N uses every slot of the array"""
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand(MAX_N, device="cuda", generator=g)
    y = torch.rand(MAX_N, device="cuda", generator=g)
    loss, loss_ref = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    scratch = torch.zeros(2, dtype=torch.float64, device="cuda")
    assert _call("kair_pixel_loss", _args(x, y, None, loss_ref, scratch, 0, 1e-9, 0.0)) == 0
    for _ in range(2):
        assert _call("kair_pixel_loss_noalloc", _args(x, y, None, loss, None, 0, 1e-9, 0.0)) == 0
        torch.cuda.synchronize()
        assert abs(loss.item() - loss_ref.item()) < 1e-6 * max(1.0, abs(loss_ref.item()))
