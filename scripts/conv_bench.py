"""Per-layer timing of the persistent tcgen05 conv on the cfg2 layer shapes (tuning aid)."""
import sys, os, ctypes, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load_package()
from dincae_b200 import lib as L
lib = L.load()
ctx = pkg.Context(0)
fn = lib.dincae_debug_conv
fn.restype = ctypes.c_int32
fn.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int32] * 6 + [ctypes.c_void_p, ctypes.c_int32]
B = 32
shapes = [(112, 32, 32), (56, 32, 64), (28, 64, 128), (14, 128, 256), (14, 256, 128), (28, 128, 64), (56, 64, 32),
          (14, 256, 256), (28, 128, 128), (56, 64, 64)]
variant = int(os.environ.get("VARIANT", "4"))
if os.environ.get("SHAPES"):
    shapes = [shapes[int(i)] for i in os.environ["SHAPES"].split(",")]
NREP = int(os.environ.get("NREP", "200"))
for (S, ci, co) in shapes:
    x = torch.randn(B, S, S, ci, device="cuda")
    w = torch.randn(9 * ci * co, device="cuda") * 0.05
    wK = torch.randn(co * 9 * ci, device="cuda") * 0.05
    bias = torch.randn(co, device="cuda")
    out = torch.empty(B, S, S, co, device="cuda")
    torch.cuda.synchronize()
    def run():
        st = fn(ctx.handle, x.data_ptr(), w.data_ptr(), wK.data_ptr(), bias.data_ptr(), B, S, S, ci, co, 1, out.data_ptr(), variant)
        assert st == 0, lib.dincae_last_error()
    for _ in range(5): run()
    ctx.sync()
    n = NREP
    t0 = time.perf_counter()
    for _ in range(n): run()
    ctx.sync()
    us = (time.perf_counter() - t0) / n * 1e6
    fl = 2.0 * 9 * ci * co * B * S * S
    print("S=%3d %3d->%3d  %7.2f us  %6.1f TFLOP/s" % (S, ci, co, us, fl / us / 1e6), flush=True)

if os.environ.get("TRACE"):
    tr = lib.dincae_debug_conv_trace
    tr.argtypes = [ctypes.c_void_p]; tr.restype = ctypes.c_int32
    for (S, ci, co) in shapes[:int(os.environ.get('NTRACE', '4'))]:
        x = torch.randn(B, S, S, ci, device="cuda"); w = torch.randn(9 * ci * co, device="cuda") * 0.05
        wK = torch.randn(co * 9 * ci, device="cuda") * 0.05; bias = torch.randn(co, device="cuda")
        out = torch.empty(B, S, S, co, device="cuda")
        trace = torch.zeros(3 * 64, dtype=torch.int64, device="cuda")
        for rep in range(3):
            trace.zero_(); torch.cuda.synchronize()
            tr(trace.data_ptr())
            fn(ctx.handle, x.data_ptr(), w.data_ptr(), wK.data_ptr(), bias.data_ptr(), B, S, S, ci, co, 1, out.data_ptr(), 4)
            ctx.sync(); tr(None)
        tv = trace.cpu().numpy().reshape(3, 64)
        t0 = tv[2, 62]
        f = lambda a: " ".join("%6d" % (v - t0) if v else "     ." for v in a)
        print("S=%d %d->%d (cycles since CTA0 start)" % (S, ci, co))
        print("  setup done  :", f(tv[2, 63:64]), " kernel end:", f(tv[0, 63:64]))
        print("  producer    :", f(tv[0, :14]))
        print("  mma         :", f(tv[1, :22]))
        print("  epilogue    :", f(tv[2, :14]))
        print("  epi detail  :", f(tv[2, 40:49]))
