"""Randomised cross-check of the two precision modes of the library against each other (forward, loss, two training steps) over
odd sizes, channel counts that are / are not multiples of 32, batch sizes, up-sampling methods and refinement."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
ctx = pkg.Context(0)
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
ncase = int(sys.argv[2]) if len(sys.argv) > 2 else 30
worst = 0.0
for case in range(ncase):
    L = int(rng.integers(1, 5))
    W, H = int(rng.integers(2 ** L, 70)), int(rng.integers(2 ** L, 70))
    base = int(rng.choice([8, 25, 32, 40, 64]))
    F = tuple(int(base * (1 + l)) for l in range(L))
    Cin, O, B = int(rng.integers(3, 40)), int(rng.integers(1, 4)), int(rng.integers(1, 7))
    method = str(rng.choice(["nearest", "bilinear"]))
    refine = bool(rng.integers(0, 2))
    kw = dict(W=W, H=H, ninput=Cin, noutput=O, enc_nfilter_internal=F, upsampling_method=method,
              loss_weights_refine=(0.3, 0.7) if refine else (1.0,), regularization_L2_beta=1e-3, max_batch=B)
    xin = np.asfortranarray(rng.standard_normal((W, H, Cin, B)).astype(np.float32))
    xt = np.asfortranarray(rng.standard_normal((W, H, 2 * O, B)).astype(np.float32))
    xt[:, :, 1::2] = np.where(rng.random((W, H, O, B)) < 0.4, 0, rng.uniform(0.5, 2, (W, H, O, B))).astype(np.float32)
    xt[:, :, 0::2] *= (xt[:, :, 1::2] != 0)
    res = {}
    for prec in ("fp32", "tf32"):
        m = pkg.Model(ctx, pkg.ModelConfig(precision=prec, **kw), seed=case)
        if prec == "fp32":
            flat = m.get_params()
        else:
            m.set_params(flat)
        y = m(xin)
        l0 = m.train_step(xin, xt, 1e-3)
        l1 = m.train_step(xin, xt, 1e-3)
        res[prec] = (y, l0, l1, m.get_params())
        m.close()
    yr, yt = res["fp32"][0], res["tf32"][0]
    err = max(np.abs(yt[:, :, c] - yr[:, :, c]).max() / max(np.abs(yr[:, :, c]).max(), 1e-6) for c in range(2 * O))
    dl0 = abs(res["tf32"][1] - res["fp32"][1]) / abs(res["fp32"][1])
    dl1 = abs(res["tf32"][2] - res["fp32"][2]) / abs(res["fp32"][2])
    dw = np.abs(res["tf32"][3] - res["fp32"][3]).max()
    worst = max(worst, err)
    flag = "" if (err < 4e-3 and dl0 < 2e-3 and dl1 < 5e-3 and np.isfinite(dw)) else "   <-- CHECK"
    print("case %2d W=%2d H=%2d Cin=%2d O=%d F=%s B=%d %-8s refine=%d: out %.1e loss %.1e / %.1e dw %.1e%s" %
          (case, W, H, Cin, O, F, B, method, refine, err, dl0, dl1, dw, flag), flush=True)
print("worst output error / range:", worst)
