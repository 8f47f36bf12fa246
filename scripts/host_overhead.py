"""Host-side enqueue cost of one training step: the cfg2 topology on a 16x16 grid (GPU work ~ nothing), steps queued back to back."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
ctx = pkg.Context(0)
for (W, B) in ((16, 2), (112, 32)):
    cfg = pkg.ModelConfig(W=W, H=W, ninput=10, enc_nfilter_internal=(32, 64, 128, 256), loss_weights_refine=(0.3, 0.7),
                          regularization_L2_beta=1e-3, max_batch=B, precision="tf32")
    m = pkg.Model(ctx, cfg, seed=1)
    rng = np.random.default_rng(0)
    xin = np.asfortranarray(rng.standard_normal((W, W, 10, B)).astype(np.float32))
    xt = np.asfortranarray(rng.standard_normal((W, W, 2, B)).astype(np.float32))
    xt[:, :, 1] = (rng.random((W, W, B)) > 0.4).astype(np.float32)
    m.upload(0, xin, xt)
    for _ in range(5):
        m.train_step_slot(0, B, 1e-3, want_loss=False)
    ctx.sync()
    n = 200
    l0 = ctx.launches
    t0 = time.perf_counter()
    for _ in range(n):
        m.train_step_slot(0, B, 1e-3, want_loss=False)
    t1 = time.perf_counter()
    ctx.sync()
    t2 = time.perf_counter()
    print("W=%d B=%d: host enqueue %.3f ms/step (%d launches/step), with final sync %.3f ms/step" %
          (W, B, (t1 - t0) / n * 1e3, (ctx.launches - l0) // n, (t2 - t0) / n * 1e3))
