"""Measure the tensor-core mode's error against the float64 oracle on the parity cases (run on the GPU box).
Prints, per case: max |dm| / max|m|, max |dsigma| / max sigma, max relative error of sigma, rms ratios, loss error."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
from oracle import model as OM

pkg = ge.load_package()
ctx = pkg.Context(0)
CASES = [
    (24, 24, 10, 1, (32, 64), "nearest", False, 0.0, 0.0, 3),
    (21, 17, 10, 1, (32, 64, 128), "nearest", True, 1e-3, 0.0, 4),
    (20, 19, 16, 2, (32, 64), "bilinear", True, 1e-4, 0.0, 2),
    (45, 18, 60, 1, (25, 50, 75), "nearest", False, 1e-4, 0.0, 2),
    (112, 112, 10, 1, (32, 64, 128, 256), "nearest", True, 1e-3, 0.0, 4),
    (112, 112, 16, 2, (32, 64, 128, 256), "nearest", True, 1e-3, 0.0, 4),
    (177, 69, 60, 1, (25, 50, 75), "nearest", False, 1e-4, 0.0, 2),
]
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_parity import _problem
out = []
for (W, H, Cin, O, F, method, refine, l2, lap, B) in CASES:
    for seed in (5, 6, 7):
        rng = np.random.default_rng(W + H + seed)
        lw = (0.3, 0.7) if refine else (1.0,)
        kw = dict(W=W, H=H, ninput=Cin, noutput=O, enc_nfilter_internal=F, upsampling_method=method, loss_weights_refine=lw,
                  regularization_L2_beta=l2, laplacian_penalty=lap, max_batch=B)
        ocfg = OM.ModelConfig(W=W, H=H, ninput=Cin, noutput=O, nfilter=list(F), upsampling_method=method, loss_weights_refine=lw,
                              regularization_L2_beta=l2, laplacian_penalty=lap)
        model = pkg.Model(ctx, pkg.ModelConfig(precision=os.environ.get("PROBE_PREC", "tf32"), **kw), seed=seed)
        flat = model.get_params()
        xin, xt = _problem(rng, W, H, Cin, O, B)
        y = model(xin).astype(np.float64)
        y_o = OM.forward(ocfg, flat, xin)
        r = dict(case=[W, H, Cin, O, list(F), method, refine], seed=seed)
        for v in range(O):
            m, s2 = y[:, :, 2 * v], y[:, :, 2 * v + 1]
            m0, s20 = y_o[:, :, 2 * v], y_o[:, :, 2 * v + 1]
            sg, sg0 = np.sqrt(s2), np.sqrt(s20)
            r["m_max_of_range_%d" % v] = float(np.abs(m - m0).max() / np.abs(m0).max())
            r["m_rms_%d" % v] = float(np.sqrt(np.mean((m - m0) ** 2) / np.mean(m0 ** 2)))
            r["sig_max_of_range_%d" % v] = float(np.abs(sg - sg0).max() / sg0.max())
            r["sig_rel_max_%d" % v] = float((np.abs(sg - sg0) / sg0).max())
            r["s2_max_of_range_%d" % v] = float(np.abs(s2 - s20).max() / s20.max())
            big = np.abs(m0) > 0.25 * np.abs(m0).max()
            r["m_rel_max_big_%d" % v] = float((np.abs(m - m0)[big] / np.abs(m0)[big]).max())
        loss_o, g_o = OM.loss_and_grad(ocfg, flat, xin, xt)
        r["loss_rel"] = float(abs(model.loss(xin, xt) - loss_o) / abs(loss_o))
        model.train_step(xin, xt, 1e-3)
        mm, vv, t = model.get_opt_state()
        gc = np.clip(g_o, -5, 5)
        r["grad_max_of_range"] = float(np.abs(mm * 10 - gc).max() / np.abs(gc).max())
        r["grad_rms"] = float(np.sqrt(np.mean((mm * 10 - gc) ** 2) / np.mean(gc ** 2)))
        print(json.dumps(r), flush=True)
        out.append(r)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "tf32_probe_%s.json" % os.environ.get("PROBE_TAG", "base")), "w"), indent=1)
