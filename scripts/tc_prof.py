"""Small driver for ncu: a few forward+backward steps of the cfg2 model (TF32 path)."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
ctx = pkg.Context(0)
W = H = 112; B = 32
cfg = pkg.ModelConfig(W=W, H=H, ninput=10, enc_nfilter_internal=(32, 64, 128, 256), loss_weights_refine=(0.3, 0.7),
                      regularization_L2_beta=1e-3, max_batch=B, precision=os.environ.get("DINCAE_PRECISION", "tf32"))
m = pkg.Model(ctx, cfg, seed=1)
rng = np.random.default_rng(0)
xin = np.asfortranarray(rng.standard_normal((W, H, 10, B)).astype(np.float32))
xt = np.asfortranarray(rng.standard_normal((W, H, 2, B)).astype(np.float32))
xt[:, :, 1] = (rng.random((W, H, B)) > 0.4).astype(np.float32)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    print(m.train_step(xin, xt, 1e-3))
