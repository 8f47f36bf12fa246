"""cfg5-shaped smoke: 1024x1024 grid, 5 levels (32..512), refinement; one train step in TF32 vs FP32 mode on the GPU."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
ctx = pkg.Context(0)
W = H = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
rng = np.random.default_rng(0)
xin = np.asfortranarray(rng.standard_normal((W, H, 10, B)).astype(np.float32))
xt = np.asfortranarray(rng.standard_normal((W, H, 2, B)).astype(np.float32))
xt[:, :, 1] = (rng.random((W, H, B)) > 0.4).astype(np.float32)
xt[:, :, 0] *= xt[:, :, 1]
res = {}
for prec in ("fp32", "tf32"):
    cfg = pkg.ModelConfig(W=W, H=H, ninput=10, enc_nfilter_internal=(32, 64, 128, 256, 512), loss_weights_refine=(0.3, 0.7),
                          regularization_L2_beta=1e-3, max_batch=B, precision=prec)
    m = pkg.Model(ctx, cfg, seed=1)
    l0 = m.train_step(xin, xt, 1e-3)
    t0 = time.perf_counter()
    for _ in range(3):
        l1 = m.train_step(xin, xt, 1e-3)
    dt = (time.perf_counter() - t0) / 3
    y = m(xin)
    res[prec] = (l0, l1, y)
    print(prec, "loss step1 %.6f step4 %.6f  %.1f ms/step (host buffers) -> %.1f samples/s, %.1f Mpixel/s train" % (l0, l1, dt * 1e3, B / dt, B * W * H / dt / 1e6))
    m.close()
print("loss rel diff tf32 vs fp32:", abs(res["tf32"][0] - res["fp32"][0]) / abs(res["fp32"][0]),
      "output max diff / range:", np.abs(res["tf32"][2] - res["fp32"][2]).max() / np.abs(res["fp32"][2]).max())
