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
