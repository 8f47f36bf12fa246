"""SURVEY section 8f "next" rows on the GPU box: checkpoint round trip (N1) and the reconstruct driver fed from / writing
NetCDF-3 files (N2, N3).  The arrays that cross the file boundary must reproduce the in-memory run bit for bit."""
import numpy as np
import pytest
import torch
from scipy.io import netcdf_file

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return pkg.Context(0)


def _problem(rng, W, H, Cin, B):
    xin = np.asfortranarray(rng.standard_normal((W, H, Cin, B)).astype(np.float32))
    xt = np.asfortranarray(rng.standard_normal((W, H, 2, B)).astype(np.float32))
    xt[:, :, 1] = np.where(rng.random((W, H, B)) < 0.4, 0, rng.uniform(0.5, 2, (W, H, B))).astype(np.float32)
    xt[:, :, 0] *= (xt[:, :, 1] != 0)
    return xin, xt


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_checkpoint_resume_is_bit_exact(pkg, ctx, tmp_path, precision):
    rng = np.random.default_rng(4)
    W, H, Cin, B = 24, 20, 10, 4
    cfg = pkg.ModelConfig(W=W, H=H, ninput=Cin, enc_nfilter_internal=(32, 64), loss_weights_refine=(0.3, 0.7),
                          regularization_L2_beta=1e-3, max_batch=B, precision=precision)
    xin, xt = _problem(rng, W, H, Cin, B)
    a = pkg.Model(ctx, cfg, seed=1)
    for _ in range(3):
        a.train_step(xin, xt, 1e-3)
    fn = str(tmp_path / "ckpt.npz")
    a.save_checkpoint(fn, epoch=3)
    ref = [a.train_step(xin, xt, 1e-3) for _ in range(3)]
    b = pkg.Model(ctx, cfg, seed=99)                       # different initial weights: everything must come from the file
    b.load_checkpoint(fn)
    got = [b.train_step(xin, xt, 1e-3) for _ in range(3)]
    assert got == ref
    assert np.array_equal(a.get_params(), b.get_params())
    # the layout table covers the flat vector exactly, in Flux.trainables order
    lay = a.flux_layout()
    assert lay[0][0] == "chain1.enc1.weight" and lay[0][1] == (3, 3, Cin, 32) and lay[-1][0] == "chain2.dec1.bias"
    assert lay[-1][2] + lay[-1][1][0] == a.num_params


def test_reconstruct_from_and_to_netcdf(pkg, ctx, tmp_path):
    rng = np.random.default_rng(7)
    W, H, N = 24, 20, 10
    lon, lat = np.linspace(-7, -1, W), np.linspace(34, 38, H)
    days1900 = 36500.0 + np.arange(N)
    sst = (rng.standard_normal((W, H, N)) + 15).astype(np.float32)
    sst[rng.random((W, H, N)) < 0.4] = np.nan
    mask = np.ones((W, H), bool); mask[:3, :2] = False
    fn_in = str(tmp_path / "sst.nc")
    ds = netcdf_file(fn_in, "w", version=2)
    ds.createDimension("time", None); ds.createDimension("lat", H); ds.createDimension("lon", W)
    v = ds.createVariable("lon", "d", ("lon",)); v[:] = lon
    v = ds.createVariable("lat", "d", ("lat",)); v[:] = lat
    v = ds.createVariable("time", "d", ("time",)); v.units = "days since 1900-01-01 00:00:00"; v[0:N] = days1900
    v = ds.createVariable("mask", "i", ("lat", "lon")); v[:, :] = mask.T.astype(np.int32)
    v = ds.createVariable("SST", "f", ("time", "lat", "lon")); v._FillValue = np.float32(-9999.0)
    v[0:N, :, :] = np.where(np.isnan(sst), -9999.0, sst).astype(np.float32).transpose(2, 1, 0)
    ds.close()
    kw = dict(epochs=2, batch_size=4, enc_nfilter_internal=(8, 12), save_epochs=[1, 2], ntime_win=3, seed=3, precision="fp32")
    entry_file = [dict(filename=fn_in, varname="SST")]
    entry_mem = [dict(lon=lon, lat=lat, time=days1900 - 25567.0, data=sst, mask=mask, varname="SST")]
    fn_out = str(tmp_path / "rec.nc")
    losses_f, res_f = pkg.reconstruct(ctx, [entry_file, entry_file], [fn_out], **kw)
    losses_m, res_m = pkg.reconstruct(ctx, [entry_mem, entry_mem], None, **kw)
    assert losses_f == losses_m
    assert np.array_equal(res_f[0]["mean"], res_m[0]["mean"]) and np.array_equal(res_f[0]["sigma"], res_m[0]["sigma"])
    out = netcdf_file(fn_out, "r", mmap=False)
    try:
        rec = out.variables["SST"][...].transpose(2, 1, 0)
        err = out.variables["SST_error"][...].transpose(2, 1, 0)
        assert rec.shape == (W, H, N) and int(out.count) == 2
        assert np.array_equal(rec, res_m[0]["mean"][:, :, :, 0]) and np.array_equal(err, res_m[0]["sigma"][:, :, :, 0])
        assert np.all(rec[:3, :2, :] == np.float32(-9999.0))         # land points carry the fill value (src/data.jl:666-669)
        assert np.all(err[rec != np.float32(-9999.0)] > 0)
    finally:
        out.close()


def test_per_pixel_error_field_assembly_bit_exact(pkg, ctx):
    """errvarname (src/data.jl:34,66-67) = a 4-D obs_err_std (:323-326): both channels divided pixel by pixel, Inf -> 0, and the
    static valid counts drop the targets that carry no weight."""
    from oracle import data as OD
    rng = np.random.default_rng(11)
    W, H, V, N, T = 20, 14, 2, 9, 3
    data = (rng.standard_normal((W, H, V, N)) + 15).astype(np.float32)
    data[rng.random((W, H, V, N)) < 0.4] = np.nan
    err = rng.uniform(0.5, 2.0, (W, H, V, N)).astype(np.float32)
    err[rng.random((W, H, V, N)) < 0.1] = np.inf
    lon, lat, days = np.linspace(-7, -1, W), np.linspace(34, 38, H), 11000.0 + np.arange(N)
    kw = dict(train=True, ntime_win=T, isoutput=[True, False], obs_err_std=err, jitter_std=[0.05, 0.1])
    ds = pkg.GriddedDataset(ctx, lon, lat, days, data, **kw)
    od = OD.NCData(lon, lat, days, data, **kw)
    idx = np.array([0, N - 1, 4, 1])
    xin, xtrue = ds.assemble(idx)
    xin_o, xtrue_o = od.batch(idx)
    assert np.array_equal(xin, xin_o) and np.array_equal(xtrue, xtrue_o)
    imask = np.array([3, 0, N - 1, 2])
    xin, xtrue = ds.assemble(idx, imask, seed=7, sample_ctr0=10)
    xin_o, xtrue_o = od.batch(idx, imask, seed=7, sample_ctr0=10)
    assert np.array_equal(xtrue, xtrue_o) and np.array_equal(xin == 0, xin_o == 0)
    np.testing.assert_allclose(xin, xin_o, rtol=0, atol=2e-6)
    assert np.array_equal(ds.valid_counts(0), (od.x[:, :, 0, :, 1] != 0).sum(axis=(0, 1)))


def test_driver_errvarname_modeldir_paramfile(pkg, ctx, tmp_path):
    """reconstruct fed from a file with an error variable, writing checkpoints at the save epochs (src/model.jl:599-620) and the
    parameter file (:655-676); the file-fed run equals the in-memory run that is given the same error field."""
    rng = np.random.default_rng(9)
    W, H, N = 16, 12, 8
    lon, lat = np.linspace(-7, -1, W), np.linspace(34, 38, H)
    days1900 = 36500.0 + np.arange(N)
    sst = (rng.standard_normal((W, H, N)) + 15).astype(np.float32)
    sst[rng.random((W, H, N)) < 0.4] = np.nan
    err = rng.uniform(0.5, 1.5, (W, H, N)).astype(np.float32)
    fn_in = str(tmp_path / "sst.nc")
    ds = netcdf_file(fn_in, "w", version=2)
    ds.createDimension("time", None); ds.createDimension("lat", H); ds.createDimension("lon", W)
    v = ds.createVariable("lon", "d", ("lon",)); v[:] = lon
    v = ds.createVariable("lat", "d", ("lat",)); v[:] = lat
    v = ds.createVariable("time", "d", ("time",)); v.units = "days since 1900-01-01 00:00:00"; v[0:N] = days1900
    v = ds.createVariable("SST", "f", ("time", "lat", "lon")); v._FillValue = np.float32(-9999.0)
    v[0:N, :, :] = np.where(np.isnan(sst), -9999.0, sst).astype(np.float32).transpose(2, 1, 0)
    v = ds.createVariable("SST_err", "f", ("time", "lat", "lon")); v._FillValue = np.float32(-9999.0)
    v[0:N, :, :] = err.transpose(2, 1, 0)
    ds.close()
    mdir, pfile = str(tmp_path / "models"), str(tmp_path / "param.nc")
    kw = dict(epochs=3, batch_size=4, enc_nfilter_internal=(8, 12), save_epochs=[2, 3], ntime_win=3, seed=3, precision="fp32",
              learning_rate_decay_epoch=50.0, loss_weights_refine=(0.3, 0.7))
    entry_file = [dict(filename=fn_in, varname="SST", errvarname="SST_err")]
    mask = np.isnan(sst).mean(axis=2) <= 0.95                                 # the loader's minfrac rule (no mask variable in the file)
    entry_mem = [dict(lon=lon, lat=lat, time=days1900 - 25567.0, data=sst, mask=mask, err=err, varname="SST")]
    losses_f, res_f = pkg.reconstruct(ctx, [entry_file, entry_file], None, modeldir=mdir, paramfile=pfile, **kw)
    losses_m, res_m = pkg.reconstruct(ctx, [entry_mem, entry_mem], None, **kw)
    assert losses_f == losses_m and np.array_equal(res_f[0]["mean"], res_m[0]["mean"])
    import os
    assert sorted(os.listdir(mdir)) == ["model-checkpoint-00002.npz", "model-checkpoint-00003.npz"]
    z = np.load(os.path.join(mdir, "model-checkpoint-00003.npz"))
    assert np.array_equal(z["params"], pkg.reconstruct.last_model.get_params()) and int(z["e"]) == 3 and int(z["ntime_win"]) == 3
    assert z["train_mean_data"].shape == (W, H, 1)
    p = netcdf_file(pfile, "r", mmap=False)
    try:
        assert np.allclose(p.variables["losses"][...], losses_f) and int(p.epochs) == 3 and list(p.save_epochs) == [2, 3]
        assert list(p.enc_nfilter_internal) == [8, 12] and list(p.skipconnections) == [2, 3] and float(p.learning_rate_decay_epoch) == 50.0
    finally:
        p.close()


def test_reconstruct_points_from_ragged_file(pkg, ctx, tmp_path):
    """reconstruct_points(filename, varname, ...) through ncio.load_ragged_points == the same data passed in memory."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_ncio import _write_ragged
    from test_gpu_points import synth_tracks, GRID
    import datetime
    rng = np.random.default_rng(4)
    N = 8
    lon, lat, val, ids = synth_tracks(rng, GRID, N)
    dates1900 = 36500.0 + 40 + np.arange(N)
    fn = tmp_path / "sla.nc"
    _write_ragged(fn, val, lon, lat, [np.full(len(v), d) for v, d in zip(val, dates1900)], [i.astype(np.int32) for i in ids], dates1900)
    doy = [(datetime.date(1900, 1, 1) + datetime.timedelta(days=int(d))).timetuple().tm_yday for d in dates1900]
    kw = dict(epochs=2, batch_size=4, enc_nfilter_internal=(5, 7), ntime_win=3, save_epochs=[2], jitter_std_pos=(0.17, 0.17),
              probability_skip_for_training=0.5, regularization_L2_beta=1e-4, seed=3, precision="fp32")
    pfile = str(tmp_path / "pparam.nc")
    lf, rf = pkg.reconstruct_points(ctx, str(fn), GRID, varname="sla", paramfile=pfile, **kw)
    lm, rm = pkg.reconstruct_points(ctx, dict(lon=lon, lat=lat, values=val, id=ids, dayofyear=np.array(doy, float)), GRID, **kw)
    assert lf == lm and np.array_equal(rf["mean"], rm["mean"])
    p = netcdf_file(pfile, "r", mmap=False)
    try:
        assert int(p.ntime_win) == 3 and np.allclose(p.jitter_std_pos, [0.17, 0.17]) and float(p.probability_skip_for_training) == 0.5
    finally:
        p.close()
