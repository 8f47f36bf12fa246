"""Numerical check of the tensor-core (TF32) path against the FP32 CUDA-core path and the float64 oracle."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from oracle import model as OM

ctx = pkg.Context(0)
rng = np.random.default_rng(0)
for (W, H, Cin, F, refine, B) in [(24, 24, 10, (32, 64), False, 3), (21, 17, 10, (32, 64, 128), True, 4), (112, 112, 10, (32, 64, 128, 256), True, 4)]:
    lw = (0.3, 0.7) if refine else (1.0,)
    kw = dict(W=W, H=H, ninput=Cin, enc_nfilter_internal=F, loss_weights_refine=lw, regularization_L2_beta=1e-3, max_batch=B)
    m32 = pkg.Model(ctx, pkg.ModelConfig(precision="fp32", **kw), seed=5)
    mtc = pkg.Model(ctx, pkg.ModelConfig(precision="tf32", **kw), seed=5)
    flat = m32.get_params()
    mtc.set_params(flat)
    xin = np.asfortranarray(rng.standard_normal((W, H, Cin, B)).astype(np.float32))
    xt = np.asfortranarray(rng.standard_normal((W, H, 2, B)).astype(np.float32))
    xt[:, :, 1] = np.where(rng.random((W, H, B)) < 0.4, 0, 1).astype(np.float32)
    xt[:, :, 0] *= xt[:, :, 1]
    y32, ytc = m32(xin), mtc(xin)
    ocfg = OM.ModelConfig(W=W, H=H, ninput=Cin, nfilter=list(F), loss_weights_refine=lw, regularization_L2_beta=1e-3)
    yo = OM.forward(ocfg, flat, xin)
    def rel(a, b):
        return float(np.abs(a - b).max() / np.abs(b).max())
    print("case", W, H, F, refine, "fwd: fp32 vs oracle %.2e | tf32 vs oracle %.2e (m) %.2e (s2)" % (rel(y32, yo), rel(ytc[:, :, 0], yo[:, :, 0]), rel(ytc[:, :, 1], yo[:, :, 1])))
    lo, go = OM.loss_and_grad(ocfg, flat, xin, xt)
    l32, ltc = m32.train_step(xin, xt, 1e-3), mtc.train_step(xin, xt, 1e-3)
    g32 = m32.get_opt_state()[0] * 10; gtc = mtc.get_opt_state()[0] * 10
    gc = np.clip(go, -5, 5)
    print("   loss oracle %.6f fp32 %.6f (rel %.1e) tf32 %.6f (rel %.1e); grad rel-to-max: fp32 %.2e tf32 %.2e ; grad rms rel: tf32 %.2e"
          % (lo, l32, abs(l32 - lo) / abs(lo), ltc, abs(ltc - lo) / abs(lo), rel(g32, gc), rel(gtc, gc),
             float(np.sqrt(np.mean((gtc - gc) ** 2)) / np.sqrt(np.mean(gc ** 2)))))
