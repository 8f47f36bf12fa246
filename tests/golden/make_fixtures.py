"""Writes tests/golden/oracle_small.npz from the CPU oracle (see README.md in this directory)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import model as OM, data as OD, points as OP  # noqa: E402


def build():
    rng = np.random.default_rng(20261106)
    out = {}
    # ---- gridded assembly (train=False: no mask perturbation, no jitter) -> model forward / loss / grad / Adam
    W, H, V, N, T, B = 18, 13, 1, 9, 3, 3
    data = (rng.standard_normal((W, H, V, N)) + 12).astype(np.float32)
    data[rng.random((W, H, V, N)) < 0.35] = np.nan
    lon, lat, days = np.linspace(-7, -1, W), np.linspace(34, 38, H), 11000.0 + np.arange(N)
    od = OD.NCData(lon, lat, days, data, train=False, ntime_win=T)
    idx = np.array([0, 4, 8])
    xin, xtrue = od.batch(idx)
    cfg = OM.ModelConfig(W=W, H=H, ninput=xin.shape[2], nfilter=[8, 12], loss_weights_refine=(0.3, 0.7), regularization_L2_beta=1e-3)
    flat = OM.init_params(cfg, seed=5)
    y = OM.forward(cfg, flat, xin)
    loss, g = OM.loss_and_grad(cfg, flat, xin, xtrue)
    flat1 = OM.Adam(flat.size, clip_grad=5.0).step(flat, g, 1e-3)
    out.update(grid_data=data, grid_lon=lon, grid_lat=lat, grid_days=days, grid_idx=idx, grid_T=np.int64(T),
               xin=xin, xtrue=xtrue, nfilter=np.array([8, 12]), params=flat, forward=y, loss=np.float64(loss), grad=g, params_after_adam=flat1)
    # ---- point cloud: interp / adjoint on a random cloud
    P = 37
    pos = np.stack([rng.uniform(1, W - 1e-3, P), rng.uniform(1, H - 1e-3, P)], 1).astype(np.float32)
    A = rng.standard_normal((W, H)).astype(np.float32)
    vals = rng.standard_normal(P).astype(np.float32)
    out.update(pt_pos=pos, pt_A=A, pt_vals=vals, pt_interp=OP.interpnd(pos, A), pt_adjoint=OP.interp_adjn(pos, vals, (W, H)))
    return out


if __name__ == "__main__":
    fx = build()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_small.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path), "bytes")
