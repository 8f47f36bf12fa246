import sys, os, ctypes as C
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
lib = pkg.lib.load()
ctx = pkg.Context(0)
f = lib.dincae_debug_wgrad
f.restype = C.c_int32
f.argtypes = [C.c_void_p] * 3 + [C.c_int32] * 5 + [C.c_void_p] * 2 + [C.c_int32]
dev = torch.device("cuda")

def run(xin, g, use):
    B, H, W, Ci = xin.shape; Co = g.shape[3]
    dW = torch.zeros(9 * Ci * Co, device=dev); db = torch.zeros(Co, device=dev)
    torch.cuda.synchronize()
    pkg.lib.check(f(ctx.handle, xin.data_ptr(), g.data_ptr(), B, H, W, Ci, Co, dW.data_ptr(), db.data_ptr(), use))
    ctx.sync()
    return dW.cpu().numpy().reshape(9, Ci, Co), db.cpu().numpy()

def tf32(x):
    return (x.view(torch.int32) + 0x1000 & ~0x1FFF).view(torch.float32) if False else x

torch.manual_seed(0)
for (B, H, W, Ci, Co) in [(1, 4, 8, 32, 32), (3, 9, 13, 32, 32), (4, 112, 112, 32, 32), (1, 4, 8, 128, 32), (1, 4, 8, 32, 128), (2, 9, 13, 64, 32), (4, 28, 28, 64, 128), (2, 14, 14, 256, 128)]:
    xin = torch.randint(-4, 5, (B, H, W, Ci), device=dev).float()      # small integers: exact in TF32, exact sums
    g = torch.randint(-4, 5, (B, H, W, Co), device=dev).float()
    r0, b0 = run(xin, g, 0)
    r1, b1 = run(xin, g, 1)
    print((B, H, W, Ci, Co), "max|simt|", np.abs(r0).max(), "max diff", np.abs(r0 - r1).max(), "bias diff", np.abs(b0 - b1).max())
    if np.abs(r0 - r1).max() > 0 and Ci * Co <= 32 * 128:
        # structured probe: one-hot in at pixel (y,x), channel c ; g = one-hot at same pixel channel d
        for (y, x, c, d) in [(0, 0, 0, 0), (0, 1, 0, 0), (1, 0, 0, 0), (0, 0, 1, 0), (0, 0, 0, 1), (0, 0, 5, 7), (2, 3, 9, 17), (3, 7, 31, 30)]:
            xi = torch.zeros_like(xin); gg = torch.zeros_like(g)
            xi[0, y, x, c] = 1; gg[0, y, x, d] = 1
            r1, _ = run(xi, gg, 1)
            nz = np.argwhere(r1 != 0)
            print("   probe pixel(%d,%d) ci=%d co=%d -> expect [tap 4, %d, %d]; got %s vals %s" % (y, x, c, d, c, d, nz.tolist()[:6], r1[r1 != 0][:6]))
        break
