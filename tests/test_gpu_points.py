"""Point-cloud path (reconstruct_points, src/points.jl) through the C-ABI against the oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import model as OM, points as OP, cbind


@pytest.fixture(scope="module")
def ctx(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return pkg.Context(0)


def synth_tracks(rng, grid, N, ntracks=4, npts=120):
    """Straight altimetry-like tracks, some points outside the domain, a few NaN values."""
    gx, gy = grid
    lon, lat, val, ids = [], [], [], []
    for n in range(N):
        lo, la, va, ii = [], [], [], []
        for t in range(ntracks):
            x0, x1 = rng.uniform(gx[0] - 2, gx[-1] + 2, 2); y0, y1 = rng.uniform(gy[0] - 2, gy[-1] + 2, 2)
            s = np.linspace(0, 1, npts)
            lo.append(x0 + (x1 - x0) * s); la.append(y0 + (y1 - y0) * s)
            v = np.sin(0.3 * lo[-1]) * np.cos(0.2 * la[-1]) + 0.02 * rng.standard_normal(npts)
            v[rng.random(npts) < 0.02] = np.nan
            va.append(v); ii.append(np.full(npts, t + 1))
        lon.append(np.concatenate(lo).astype(np.float32)); lat.append(np.concatenate(la).astype(np.float32))
        val.append(np.concatenate(va).astype(np.float32)); ids.append(np.concatenate(ii))
    return lon, lat, val, ids


GRID = (np.arange(-7, 37.001, 1.0), np.arange(29, 46.001, 1.0))          # 45 x 18


GRID_FULL = (np.arange(-7, 37.001, 0.25), np.arange(29, 46.001, 0.25))     # 177 x 69: test/test_reconstruct_points.jl:27-31 (cfg4)


def make(pkg, ctx, rng, N=7, T=3, naux=2, train=True, GRID=GRID, ntracks=4, npts=120):
    lon, lat, val, ids = synth_tracks(rng, GRID, N, ntracks, npts)
    W, H = len(GRID[0]), len(GRID[1])
    aux = rng.standard_normal((W, H, naux, N)).astype(np.float32) if naux else None
    doy = 50 + np.arange(N)
    kw = dict(ntime_win=T, probability_skip_for_training=0.5, jitter_std_pos=(0.17, 0.17), auxdata=aux, train=train)
    ds = pkg.PointDataset(ctx, GRID, lon, lat, val, ids, doy, **kw)
    od = OP.PointCloud(GRID, lon, lat, val, ids, doy, **kw)
    return ds, od


def test_point_dataset_filtering_bit_exact(pkg, ctx):
    rng = np.random.default_rng(0)
    ds, od = make(pkg, ctx, rng)
    for n in range(len(ds)):
        assert np.array_equal(ds.pos[n], od.pos[n]) and np.array_equal(ds.x[n], od.x[n]) and np.array_equal(ds.id[n], od.id[n])
    assert ds.sizex() == od.sizex()


def test_points_assemble_unperturbed_bit_exact(pkg, ctx):
    rng = np.random.default_rng(1)
    ds, od = make(pkg, ctx, rng, train=False)
    idx = np.array([0, 3, 6, 2])
    xin, truth = ds.assemble(idx)
    W, H, C = ds.sizex()
    T, naux = 3, 2
    for b, ind in enumerate(idx):
        ref, (tp, tx) = od.getxy(int(ind))
        assert np.array_equal(truth[b][0], tp) and np.array_equal(truth[b][1], tx)
        assert np.array_equal(xin[:, :, :4 + naux, b], ref[:, :, :4 + naux])         # meta + aux channels
        for k in range(T):
            n = min(max(ind - 1 + k, 0), len(ds) - 1)
            for e in range(2):        # gridded channels: bit-exact against the sequential C loop of interp_adjn!
                seq = cbind.interp_adjn(od.pos[n], od.x[n][:, e], (W, H))
                assert np.array_equal(xin[:, :, 4 + naux + 2 * k + e, b], seq)
        np.testing.assert_allclose(xin[..., b], ref, rtol=1e-5, atol=1e-5)


def test_points_assemble_perturbed_and_dropout(pkg, ctx):
    rng = np.random.default_rng(2)
    ds, od = make(pkg, ctx, rng, train=True)
    idx = np.array([1, 5, 0])
    keep = [ds.draw_track_keep(int(i), rng) for i in idx]
    assert all(k.any() for k in keep)                                               # at least one track stays (src/points.jl:296-299)
    xin, _ = ds.assemble(idx, seed=11, sample_ctr0=300, keep=keep)
    for b, ind in enumerate(idx):
        ref, _ = od.getxy(int(ind), seed=11, sample_ctr=300 + b, track_keep=keep[b])
        # bilinear scatter is continuous in the position: ulp differences of the Philox normals give tiny deviations only
        np.testing.assert_allclose(xin[..., b], ref, rtol=1e-4, atol=2e-4)
        # the weight channels sum to the number of gridded valid observations (integer check of the dropout)
        c = 4 + 2 + 2 * 1 + 1
        nvalid = od.x[ind][keep[b], 1].sum()
        assert abs(xin[:, :, c, b].sum() - nvalid) < 1e-2


@pytest.mark.parametrize("refine,lap,tu", [(False, 0.0, False), (True, 0.1, False), (False, 0.0, True)])
def test_point_loss_and_train_step_vs_oracle(pkg, ctx, refine, lap, tu):
    rng = np.random.default_rng(3)
    ds, od = make(pkg, ctx, rng, N=6, T=3, naux=0, train=True)
    W, H, C = ds.sizex()
    idx = np.array([0, 2, 5, 3])
    keep = [ds.draw_track_keep(int(i), rng) for i in idx]
    xin, truth = ds.assemble(idx, seed=5, sample_ctr0=0, keep=keep)
    lw = (0.3, 0.7) if refine else (1.0,)
    cfg = pkg.ModelConfig(W=W, H=H, ninput=C, enc_nfilter_internal=(6, 9), loss_weights_refine=lw, regularization_L2_beta=1e-4,
                          laplacian_penalty=lap, truth_uncertain=tu, max_batch=4, precision="fp32")
    ocfg = OM.ModelConfig(W=W, H=H, ninput=C, nfilter=[6, 9], loss_weights_refine=lw, regularization_L2_beta=1e-4,
                          laplacian_penalty=lap, truth_uncertain=tu)
    model = pkg.Model(ctx, cfg, seed=6)
    flat = model.get_params()
    loss_o, g_o = OM.loss_and_grad(ocfg, flat, xin, None, point_truth=truth)
    assert model.loss_points(xin, truth) == pytest.approx(loss_o, rel=1e-4)
    loss = model.train_step_points(xin, truth, 1e-3)
    assert loss == pytest.approx(loss_o, rel=1e-4)
    m, v, t = model.get_opt_state()
    gc = np.clip(g_o, -5, 5)
    np.testing.assert_allclose(m, 0.1 * gc, rtol=2e-3, atol=2e-6 * np.abs(gc).max())    # m = (1-beta1) g after the first step


def test_reconstruct_points_runs_and_is_deterministic(pkg, ctx):
    # test/test_reconstruct_points.jl:67-110 (shape/length checks) on synthetic tracks
    rng = np.random.default_rng(4)
    lon, lat, val, ids = synth_tracks(rng, GRID, 10)
    data = dict(lon=lon, lat=lat, values=val, id=ids, dayofyear=50 + np.arange(10))
    kw = dict(epochs=2, batch_size=4, enc_nfilter_internal=(5, 7), ntime_win=3, save_epochs=[2], jitter_std_pos=(0.17, 0.17),
              probability_skip_for_training=0.5, regularization_L2_beta=1e-4, laplacian_penalty=0.1, seed=3, precision="fp32")
    losses, res = pkg.reconstruct_points(ctx, data, GRID, **kw)
    assert len(losses) == 2 and np.all(np.isfinite(losses))
    assert res["mean"].shape == (45, 18, 10) and res["sigma"].shape == (45, 18, 10) and res["count"] == 1
    assert np.all(res["sigma"] > 0)
    losses2, res2 = pkg.reconstruct_points(ctx, data, GRID, **kw)
    assert losses == losses2 and np.array_equal(res["mean"], res2["mean"])             # no float atomics anywhere: bit-reproducible
    with pytest.raises(ValueError, match="No output will be saved"):
        pkg.reconstruct_points(ctx, data, GRID, epochs=2, save_epochs=[])


def test_reconstruct_gridded_driver(pkg, ctx):
    # test/test_DINCAE_SST.jl:62-96: output exists, length(losses) == epochs, for the five variants minus is3D
    rng = np.random.default_rng(5)
    W, H, N = 16, 12, 14
    d = (rng.standard_normal((W, H, N)) + 14).astype(np.float32)
    d[rng.random((W, H, N)) < 0.4] = np.nan
    entry = dict(lon=np.linspace(-7, -1, W), lat=np.linspace(34, 38, H), time=11000.0 + np.arange(N), data=d,
                 mask=rng.random((W, H)) > 0.1)
    for method, tu, lw in (("nearest", False, (1.0,)), ("bilinear", False, (1.0,)), ("nearest", True, (1.0,)), ("nearest", False, (0.3, 0.7))):
        losses, res = pkg.reconstruct(ctx, [[entry], [entry]], epochs=3, batch_size=5, enc_nfilter_internal=(6, 8), save_epochs=[2, 3],
                                      upsampling_method=method, truth_uncertain=tu, loss_weights_refine=lw, precision="fp32", seed=1)
        assert len(losses) == 3 and np.all(np.isfinite(losses))
        assert res[0]["count"] == 2 and res[0]["mean"].shape == (W, H, N, 1)
        assert np.all(res[0]["mean"][~entry["mask"]] == -9999)
    with pytest.raises(pkg.lib.DincaeError):
        pkg.reconstruct(ctx, [[entry], [entry]], epochs=1, save_epochs=[1], is3D=True)


def test_point_step_with_odd_point_counts_and_device_batch(pkg, ctx):
    """Regression: the int64 offsets behind the float point workspace were misaligned for some total point counts.  Also the
    device-resident batch (assemble(..., xin=device buffer) -> train_step_points) must equal the host-array path bit for bit."""
    for npts in (7, 50, 51):
        rng = np.random.default_rng(npts)
        lon, lat, val, ids = synth_tracks(rng, GRID, 8, ntracks=6, npts=npts)
        ds = pkg.PointDataset(ctx, GRID, lon, lat, val, ids, 50 + np.arange(8), ntime_win=3, probability_skip_for_training=0.5,
                              jitter_std_pos=(0.17, 0.17), auxdata=None, train=True)
        W, H, C = ds.sizex()
        idx = np.array([0, 3, 5, 6])
        keep = [ds.draw_track_keep(int(i), rng) for i in idx]
        cfg = pkg.ModelConfig(W=W, H=H, ninput=C, enc_nfilter_internal=(8, 12), regularization_L2_beta=1e-4, max_batch=4, precision="fp32")
        xin, truth = ds.assemble(idx, seed=5, sample_ctr0=0, keep=keep)
        a = pkg.Model(ctx, cfg, seed=2)
        la = [a.train_step_points(xin, truth, 1e-3) for _ in range(2)]
        buf = torch.empty(W * H * C * 4, dtype=torch.float32, device="cuda")
        xdev, truth2 = ds.assemble(idx, seed=5, sample_ctr0=0, keep=keep, xin=buf)
        b = pkg.Model(ctx, cfg, seed=2)
        lb = [b.train_step_points(xdev, truth2, 1e-3) for _ in range(2)]
        assert la == lb and np.isfinite(la).all()


@pytest.mark.parametrize("full", [False, True])
def test_point_loss_and_train_step_tf32_vs_oracle(pkg, ctx, full):
    """The point-cloud loss through the tensor-core model; full=True is BASELINE configs[3] at its real size: 177 x 69 grid,
    ntime_win = 27 (60 input channels), enc_nfilter_internal = [25, 50, 75], Laplacian penalty.  Stated tolerance: 1e-3 on the loss;
    gradients rms 1e-2 relative and 4e-2 of their range at worst (a few hundred observations drive the whole gradient, so TF32
    rounding of the backward products shows more than in the gridded loss)."""
    rng = np.random.default_rng(8)
    if full:
        ds, od = make(pkg, ctx, rng, N=6, T=27, naux=2, train=True, GRID=GRID_FULL, ntracks=6, npts=400)
        F, B, idx, lap = (25, 50, 75), 2, np.array([1, 4]), 0.1
    else:
        ds, od = make(pkg, ctx, rng, N=6, T=3, naux=2, train=True)
        F, B, idx, lap = (32, 64), 4, np.array([1, 2, 4, 3]), 0.0
    W, H, C = ds.sizex()
    assert not full or (W, H, C) == (177, 69, 60)
    keep = [ds.draw_track_keep(int(i), rng) for i in idx]
    xin, truth = ds.assemble(idx, seed=5, sample_ctr0=0, keep=keep)
    kw = dict(loss_weights_refine=(1.0,) if full else (0.3, 0.7), regularization_L2_beta=1e-4, laplacian_penalty=lap)
    cfg = pkg.ModelConfig(W=W, H=H, ninput=C, enc_nfilter_internal=F, max_batch=B, precision="tf32", clip_grad=None, **kw)
    ocfg = OM.ModelConfig(W=W, H=H, ninput=C, nfilter=list(F), **kw)
    model = pkg.Model(ctx, cfg, seed=6)
    flat = model.get_params()
    loss_o, g_o = OM.loss_and_grad(ocfg, flat, xin, None, point_truth=truth)
    assert model.loss_points(xin, truth) == pytest.approx(loss_o, rel=1e-3)
    assert model.train_step_points(xin, truth, 1e-3) == pytest.approx(loss_o, rel=1e-3)
    m, v, t = model.get_opt_state()              # clip_grad off: m = 0.1 * gradient, compared unclipped
    d = m * 10 - g_o
    assert np.abs(d).max() < 4e-2 * np.abs(g_o).max() and np.sqrt((d ** 2).mean()) < 1e-2 * np.sqrt((g_o ** 2).mean())
