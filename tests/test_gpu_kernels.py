"""Kernel-level exactness of the tensor-core kernels against the FP32 CUDA-core kernels.
Operands are small integers (exactly representable in TF32, exact FP32 sums), so the persistent / paired-tile tcgen05 conv and the
two weight-gradient kernels (wgrad9, wgradm) must agree BIT FOR BIT with the FP32 kernels and with torch: any swizzle /
descriptor / tap-mapping error shows up as a hard mismatch."""
import ctypes as C
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    lib = pkg.lib.load()
    lib.dincae_debug_conv.restype = C.c_int32
    lib.dincae_debug_conv.argtypes = [C.c_void_p] * 5 + [C.c_int32] * 6 + [C.c_void_p, C.c_int32]
    lib.dincae_debug_wgrad.restype = C.c_int32
    lib.dincae_debug_wgrad.argtypes = [C.c_void_p] * 3 + [C.c_int32] * 5 + [C.c_void_p] * 2 + [C.c_int32]
    return pkg, lib, pkg.Context(0)


CONV_SHAPES = [(1, 16, 8, 32, 32), (2, 9, 13, 32, 64), (3, 14, 14, 64, 128), (2, 28, 28, 128, 256), (2, 7, 5, 256, 32), (4, 56, 56, 32, 32),
               (2, 12, 15, 192, 96), (3, 6, 4, 96, 192), (2, 23, 15, 64, 64)]       # channel counts that are not powers of two


@pytest.mark.parametrize("B,H,W,Ci,Co", CONV_SHAPES)
def test_conv_kernel_families_bit_exact(env, B, H, W, Ci, Co):
    pkg, lib, ctx = env
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(B * 1000 + H)
    x = torch.randint(-3, 4, (B, H, W, Ci), generator=g).float().to(dev)
    w = torch.randint(-2, 3, (9, Ci, Co), generator=g).float()              # [tap][ci][co]
    kpad = (9 * Ci + 15) // 16 * 16
    w_fwd = torch.zeros(kpad, Co); w_fwd[:9 * Ci] = w.reshape(9 * Ci, Co)
    w_k = w.reshape(9 * Ci, Co).t().contiguous()                             # K-major [co][tap*Ci+ci]
    bias = torch.randint(-5, 6, (Co,), generator=g).float()
    w_fwd, w_k, bias = w_fwd.to(dev), w_k.to(dev), bias.to(dev)
    outs = []
    for variant in (0, 4):
        out = torch.full((B, H, W, Co), 7.0, device=dev)
        torch.cuda.synchronize()
        pkg.lib.check(lib.dincae_debug_conv(ctx.handle, x.data_ptr(), w_fwd.data_ptr(), w_k.data_ptr(), bias.data_ptr(),
                                            B, H, W, Ci, Co, 1, out.data_ptr(), variant))
        ctx.sync()
        outs.append(out.cpu())
    # independent check of the FP32 kernel itself against torch (cross-correlation, pad 1, ReLU)
    ref = torch.nn.functional.conv2d(x.cpu().permute(0, 3, 1, 2), w.reshape(3, 3, Ci, Co).permute(3, 2, 0, 1), bias.cpu(), padding=1).relu().permute(0, 2, 3, 1)
    assert torch.equal(outs[0], ref)
    assert torch.equal(outs[1], outs[0]), "the tcgen05 conv differs from the FP32 kernel"


WGRAD_SHAPES = [(1, 4, 8, 32, 32), (3, 9, 13, 32, 32), (2, 112, 112, 32, 32), (1, 4, 8, 128, 32), (2, 9, 13, 64, 32), (2, 9, 13, 32, 64),
                (4, 28, 28, 64, 128), (2, 14, 14, 256, 128), (2, 14, 14, 128, 256), (1, 6, 7, 512, 256),
                (2, 12, 15, 192, 96), (3, 6, 4, 96, 192), (2, 23, 15, 64, 64), (2, 5, 3, 160, 32)]


@pytest.mark.parametrize("B,H,W,Ci,Co", WGRAD_SHAPES)
def test_wgrad_kernel_families_bit_exact(env, B, H, W, Ci, Co):
    pkg, lib, ctx = env
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(Ci + Co + H)
    x = torch.randint(-4, 5, (B, H, W, Ci), generator=g).float().to(dev)
    gout = torch.randint(-4, 5, (B, H, W, Co), generator=g).float().to(dev)
    res = []
    for use_tma in (0, 1):
        dW = torch.zeros(9 * Ci * Co, device=dev); db = torch.zeros(Co, device=dev)
        torch.cuda.synchronize()
        pkg.lib.check(lib.dincae_debug_wgrad(ctx.handle, x.data_ptr(), gout.data_ptr(), B, H, W, Ci, Co, dW.data_ptr(), db.data_ptr(), use_tma))
        ctx.sync()
        res.append((dW.cpu(), db.cpu()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
    assert torch.equal(res[0][1], gout.cpu().sum(dim=(0, 1, 2)))
    # independent check against torch autograd of the cross-correlation
    xt = x.cpu().permute(0, 3, 1, 2)
    wt = torch.zeros(Co, Ci, 3, 3, requires_grad=True)
    torch.nn.functional.conv2d(xt, wt, padding=1).backward(gout.cpu().permute(0, 3, 1, 2))
    ref = wt.grad.permute(2, 3, 1, 0).reshape(9, Ci, Co)                      # [tap=(ky,kx)][ci][co]
    assert torch.equal(res[0][0].reshape(9, Ci, Co), ref)


@pytest.mark.parametrize("B,H,W,Ci,Co", [(2, 9, 13, 32, 64), (3, 14, 14, 128, 256), (1, 20, 10, 32, 32)])
def test_persistent_conv_writes_only_its_output(env, B, H, W, Ci, Co):
    """Canaries around the output tensor: the TMA-store epilogue (edge patches are clipped by the unit) and the paired-tile kernel
    (phantom second tile of an odd tile count) must not touch a byte outside [B,H,W,Co]."""
    pkg, lib, ctx = env
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(7)
    x = torch.randint(-3, 4, (B, H, W, Ci), generator=g).float().to(dev)
    w = torch.randint(-2, 3, (9, Ci, Co), generator=g).float()
    kpad = (9 * Ci + 15) // 16 * 16
    w_fwd = torch.zeros(kpad, Co); w_fwd[:9 * Ci] = w.reshape(9 * Ci, Co)
    w_k = w.reshape(9 * Ci, Co).t().contiguous()
    bias = torch.zeros(Co)
    w_fwd, w_k, bias = w_fwd.to(dev), w_k.to(dev), bias.to(dev)
    n = B * H * W * Co
    pad = 1 << 16
    buf = torch.full((pad + n + pad,), -123.0, device=dev)
    out = buf[pad:pad + n]
    torch.cuda.synchronize()
    pkg.lib.check(lib.dincae_debug_conv(ctx.handle, x.data_ptr(), w_fwd.data_ptr(), w_k.data_ptr(), bias.data_ptr(),
                                        B, H, W, Ci, Co, 0, out.data_ptr(), 4))
    ctx.sync()
    assert torch.all(buf[:pad] == -123.0) and torch.all(buf[pad + n:] == -123.0)
    ref = torch.nn.functional.conv2d(x.cpu().permute(0, 3, 1, 2), w.reshape(3, 3, Ci, Co).permute(3, 2, 0, 1), None, padding=1).permute(0, 2, 3, 1)
    assert torch.equal(out.cpu().reshape(B, H, W, Co), ref)
