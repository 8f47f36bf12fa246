"""smoke(): one small invocation of the hot path on cuda:0, checked against the CPU oracle.

Lives beside __graft_entry__.py, OUTSIDE the product package: the oracle (oracle/) is test infrastructure and nothing under
dincae.jl_b200/ imports it; here it is the checker only."""
import numpy as np


def smoke(verbose=True):
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("dincae_b200.smoke needs a CUDA device; the library has no CPU path")
    import __graft_entry__ as ge
    api = ge.load_package().api
    from oracle import model as OM, data as OD, points as OP

    ctx = api.Context(0)
    rng = np.random.default_rng(0)
    # --- K4 golden vectors of test/test_interp_adjoint.jl:20-47
    pos = np.array([[1.5, 2.0], [5.0, 2.3]], np.float32)
    A = api.interp_adjn(pos, np.array([3, 4], np.float32), (6, 7), ctx)
    ref = OP.interp_adjn(pos, np.array([3, 4], np.float32), (6, 7))
    assert np.array_equal(A, ref), "interp_adjn differs from the oracle"
    # --- K1 assembly (bit-exact without jitter) + model train step (refinement + L2) on a 20x14 cube
    W, H, V, N, T, B = 20, 14, 1, 12, 3, 4
    data = (rng.standard_normal((W, H, V, N)) + 12).astype(np.float32)
    data[rng.random((W, H, V, N)) < 0.35] = np.nan
    lon, lat, days = np.linspace(-7, -1, W), np.linspace(34, 38, H), 11000.0 + np.arange(N)
    ds = api.GriddedDataset(ctx, lon, lat, days, data, train=True, ntime_win=T)
    od = OD.NCData(lon, lat, days, data, train=True, ntime_win=T)
    idx = np.array([0, 5, 11, 7])
    xin, xtrue = ds.assemble(idx)
    xin_o, xtrue_o = od.batch(idx)
    assert np.array_equal(xin, xin_o) and np.array_equal(xtrue, xtrue_o), "assembly differs from the oracle"
    cfg = api.ModelConfig(W=W, H=H, ninput=xin.shape[2], enc_nfilter_internal=(8, 12), loss_weights_refine=(0.3, 0.7),
                          regularization_L2_beta=1e-3, max_batch=B, precision="fp32")
    ocfg = OM.ModelConfig(W=W, H=H, ninput=xin.shape[2], nfilter=[8, 12], loss_weights_refine=(0.3, 0.7), regularization_L2_beta=1e-3)
    model = api.Model(ctx, cfg, seed=3)
    flat = model.get_params()
    y = model(xin)
    y_o = OM.forward(ocfg, flat, xin)
    err_f = np.abs(y - y_o).max() / np.abs(y_o).max()
    loss = model.train_step(xin, xtrue, 1e-3)
    opt = OM.Adam(flat.size, clip_grad=5.0)
    loss_o, g_o = OM.loss_and_grad(ocfg, flat, xin, xtrue)
    flat_o = opt.step(flat, g_o, 1e-3)
    err_l = abs(loss - loss_o) / abs(loss_o)
    big = np.abs(g_o) > 1e-5          # the first Adam step is lr*g/(|g|+eps): ill-conditioned where g ~ eps
    err_w = np.abs(model.get_params() - flat_o)[big].max()
    if verbose:
        print("smoke: forward rel err %.2e, loss %.6f vs oracle %.6f (rel %.2e), max |dw| after Adam %.2e, launches %d"
              % (err_f, loss, loss_o, err_l, err_w, ctx.launches))
    assert err_f < 1e-3 and err_l < 1e-3 and err_w < 2e-4, "train step differs from the oracle"
    # --- the tensor-core path (tcgen05 + TMA, TF32 operands): same check with 32/64 filters
    cfg2 = api.ModelConfig(W=W, H=H, ninput=xin.shape[2], enc_nfilter_internal=(32, 64), loss_weights_refine=(0.3, 0.7),
                           regularization_L2_beta=1e-3, max_batch=B, precision="tf32")
    ocfg2 = OM.ModelConfig(W=W, H=H, ninput=xin.shape[2], nfilter=[32, 64], loss_weights_refine=(0.3, 0.7), regularization_L2_beta=1e-3)
    m2 = api.Model(ctx, cfg2, seed=3)
    flat2 = m2.get_params()
    y2 = m2(xin)
    y2_o = OM.forward(ocfg2, flat2, xin)
    err_tc = max(np.abs(y2[:, :, c] - y2_o[:, :, c]).max() / np.abs(y2_o[:, :, c]).max() for c in range(2))
    loss2 = m2.train_step(xin, xtrue, 1e-3)
    loss2_o, _ = OM.loss_and_grad(ocfg2, flat2, xin, xtrue)
    err_l2 = abs(loss2 - loss2_o) / abs(loss2_o)
    if verbose:
        print("smoke: tcgen05/TMA path forward err %.2e of range, loss %.6f vs oracle %.6f (rel %.2e), launches %d"
              % (err_tc, loss2, loss2_o, err_l2, ctx.launches))
    assert err_tc < 2e-3 and err_l2 < 1e-3, "tensor-core train step differs from the oracle"
    return dict(forward_rel_err=float(err_f), loss_rel_err=float(err_l), max_dw=float(err_w), tc_forward_err=float(err_tc),
                tc_loss_rel_err=float(err_l2), launches=ctx.launches)
