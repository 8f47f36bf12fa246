"""NetCDF-3 loader and reconstruction writer (host side, no GPU)."""
import numpy as np
import pytest
from scipy.io import netcdf_file

import __graft_entry__ as ge


@pytest.fixture(scope="module")
def ncio():
    ge.load_package()
    import importlib
    return importlib.import_module("dincae_b200.ncio")


def _write_input(path, lon, lat, days_since_1900, sst, mask=None, fill=-9999.0):
    ds = netcdf_file(str(path), "w", version=2)
    ds.createDimension("time", None); ds.createDimension("lat", len(lat)); ds.createDimension("lon", len(lon))
    v = ds.createVariable("lon", "d", ("lon",)); v[:] = lon
    v = ds.createVariable("lat", "d", ("lat",)); v[:] = lat
    v = ds.createVariable("time", "d", ("time",)); v.units = "days since 1900-01-01 00:00:00"; v[0:len(days_since_1900)] = days_since_1900
    if mask is not None:
        v = ds.createVariable("mask", "i", ("lat", "lon")); v[:, :] = mask.T.astype(np.int32)
    v = ds.createVariable("SST", "f", ("time", "lat", "lon")); v._FillValue = np.float32(fill)
    a = np.where(np.isnan(sst), fill, sst).astype(np.float32)
    v[0:sst.shape[2], :, :] = a.transpose(2, 1, 0)
    ds.close()


def test_load_gridded_nc_roundtrip(tmp_path, ncio):
    rng = np.random.default_rng(0)
    W, H, N = 7, 5, 4
    lon, lat = np.linspace(-7, -1, W), np.linspace(34, 38, H)
    t1900 = 36500.0 + np.arange(N)
    sst = (rng.standard_normal((W, H, N)) + 15).astype(np.float32)
    sst[rng.random((W, H, N)) < 0.3] = np.nan
    mask = rng.random((W, H)) > 0.2
    fn = tmp_path / "in.nc"
    _write_input(fn, lon, lat, t1900, sst, mask)
    lon2, lat2, time, data, err, missing, mask2 = ncio.load_gridded_nc(str(fn), "SST")
    assert np.array_equal(lon2, lon) and np.array_equal(lat2, lat)
    assert data.shape == (W, H, N) and np.array_equal(np.isnan(data), np.isnan(sst))
    assert np.array_equal(data[~np.isnan(sst)], sst[~np.isnan(sst)])
    assert np.array_equal(missing, np.isnan(sst)) and np.array_equal(mask2, mask)
    assert np.all(err == 1.0)
    # 1900-01-01 is 25567 days before 1970-01-01
    assert np.allclose(time, t1900 - 25567.0)
    # without a mask variable: sea points need >= minfrac valid data through time (src/data.jl:48-57)
    fn2 = tmp_path / "in2.nc"
    _write_input(fn2, lon, lat, t1900, sst, None)
    m3 = ncio.load_gridded_nc(str(fn2), "SST", minfrac=0.5)[6]
    assert np.array_equal(m3, np.isnan(sst).mean(axis=2) <= 0.5)


def test_non_netcdf3_is_refused(tmp_path, ncio):
    fn = tmp_path / "h5.nc"
    fn.write_bytes(b"\x89HDF\r\n\x1a\n" + b"\0" * 64)
    with pytest.raises(ValueError, match="NetCDF-3"):
        ncio.load_gridded_nc(str(fn), "SST")


def test_normalize2_golden(ncio):
    # test/test_utils.jl:4-8 of the reference: normalize2 of a constant is zero, otherwise zero mean / unit std
    assert np.all(ncio.normalize2(np.full((3, 4), 2.5)) == 0)
    x = np.random.default_rng(1).standard_normal((6, 7)) * 3 + 2
    y = ncio.normalize2(x)
    assert abs(y.mean()) < 1e-12 and abs(y.std(ddof=1) - 1) < 1e-12


def test_write_reconstruction_layout(tmp_path, ncio):
    rng = np.random.default_rng(2)
    W, H, N = 6, 4, 3
    lon, lat = np.linspace(0, 5, W), np.linspace(40, 43, H)
    mean = rng.standard_normal((W, H, N)).astype(np.float32); mean[0, 0, :] = -9999.0
    sigma = rng.random((W, H, N)).astype(np.float32); sigma[0, 0, :] = -9999.0
    meandata = rng.standard_normal((W, H, 1)).astype(np.float32); meandata[1, 1, 0] = np.nan
    fn = tmp_path / "rec.nc"
    ncio.write_reconstruction(str(fn), ["SST"], lon, lat, meandata, mean, sigma, count=2)
    raw = open(fn, "rb").read(4)
    assert raw == b"CDF\x02"                                     # :netcdf3_64bit_offset, src/data.jl:720
    ds = netcdf_file(str(fn), "r", mmap=False)
    try:
        assert ds.dimensions["time"] is None and ds.dimensions["lon"] == W and ds.dimensions["lat"] == H
        assert set(ds.variables) == {"lon", "lat", "SST", "SST_error", "SST_mean"}
        assert ds.variables["SST"].dimensions == ("time", "lat", "lon")
        assert np.array_equal(ds.variables["SST"][...].transpose(2, 1, 0), mean)
        assert np.array_equal(ds.variables["SST_error"][...].transpose(2, 1, 0), sigma)
        md = ds.variables["SST_mean"][...].T
        assert md[1, 1] == np.float32(-9999.0) and np.array_equal(np.delete(md.ravel(), 1 * H + 1), np.delete(meandata[:, :, 0].ravel(), 1 * H + 1))
        assert ds.variables["SST"]._FillValue == np.float32(-9999.0)
        assert int(ds.count) == 2
        assert np.allclose(ds.variables["lon"][...], lon) and np.allclose(ds.variables["lat"][...], lat)
    finally:
        ds.close()


def _write_ragged(path, values, lon, lat, dtime, ids, dates1900, varname="sla"):
    """CF contiguous ragged array file as documented in src/points.jl:478-501 (NetCDF-3: int instead of int64)."""
    ds = netcdf_file(str(path), "w", version=2)
    nobs = sum(len(v) for v in values)
    ds.createDimension("time_instances", len(values)); ds.createDimension("obs", nobs)
    v = ds.createVariable("size", "i", ("time_instances",)); v.sample_dimension = "obs"; v[:] = [len(x) for x in values]
    v = ds.createVariable("dates", "d", ("time_instances",)); v.units = "days since 1900-01-01 00:00:00"; v[:] = dates1900
    for name, arrs, t in ((varname, values, "f"), ("lon", lon, "f"), ("lat", lat, "f"), ("id", ids, "i"), ("dtime", dtime, "d")):
        v = ds.createVariable(name, t, ("obs",))
        if name == "dtime":
            v.units = "days since 1900-01-01 00:00:00"
        if name == varname:
            v._FillValue = np.float32(-9999.0)
        a = np.concatenate(arrs) if nobs else np.zeros(0)
        if name == varname:
            a = np.where(np.isnan(a), -9999.0, a)
        v[:] = a
    ds.close()


def test_load_ragged_points(tmp_path, ncio):
    # DINCAE.loaddata (src/points.jl:429-446): per-instance vectors out of the contiguous ragged representation, empty instance included
    rng = np.random.default_rng(3)
    counts = [4, 0, 7, 1]
    values = [rng.standard_normal(c).astype(np.float32) for c in counts]
    values[2][3] = np.nan
    lon = [rng.uniform(-7, 37, c).astype(np.float32) for c in counts]
    lat = [rng.uniform(29, 46, c).astype(np.float32) for c in counts]
    ids = [rng.integers(1, 5, c).astype(np.int32) for c in counts]
    dates = 36500.0 + 59 + np.arange(4)                       # 1999-12-08 + 59 ... day of year known below
    dtime = [np.full(c, d) + rng.uniform(-0.5, 0.5, c) for c, d in zip(counts, dates)]
    fn = tmp_path / "sla.nc"
    _write_ragged(fn, values, lon, lat, dtime, ids, dates)
    v2, (lo2, la2), dt2, id2, dates70, doy = ncio.load_ragged_points(str(fn), "sla")
    assert [len(x) for x in v2] == counts
    for a, b in zip(values, v2):
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)], b[~np.isnan(b)])
    assert all(np.array_equal(a, b) for a, b in zip(lon, lo2)) and all(np.array_equal(a, b) for a, b in zip(lat, la2))
    assert all(np.array_equal(a, b) for a, b in zip(ids, id2)) and id2[0].dtype == np.int64
    assert np.allclose(dates70, dates - 25567.0) and all(np.allclose(a - 25567.0, b) for a, b in zip(dtime, dt2))
    import datetime
    d0 = datetime.date(1900, 1, 1) + datetime.timedelta(days=int(dates[0]))
    assert doy[0] == d0.timetuple().tm_yday and np.all(np.diff(doy) == 1)


def test_write_paramfile_attributes(tmp_path, ncio):
    # src/model.jl:655-676: losses(epochs) + one global attribute per keyword
    fn = tmp_path / "param.nc"
    ncio.write_paramfile(str(fn), [3.0, 2.0, 1.5], dict(epochs=3, batch_size=50, truth_uncertain=False, enc_nfilter_internal=[16, 24, 36, 54],
                                                       skipconnections=[2, 3, 4, 5], clip_grad=5.0, save_epochs=[2, 3], upsampling_method="nearest",
                                                       learning_rate_decay_epoch=float("inf"), loss_weights_refine=[0.3, 0.7], is3D=False))
    ds = netcdf_file(str(fn), "r", mmap=False)
    try:
        assert np.array_equal(ds.variables["losses"][...], [3.0, 2.0, 1.5]) and ds.dimensions["epochs"] == 3
        assert int(ds.epochs) == 3 and int(ds.batch_size) == 50 and int(ds.truth_uncertain) == 0 and int(ds.is3D) == 0
        assert list(ds.enc_nfilter_internal) == [16, 24, 36, 54] and list(ds.skipconnections) == [2, 3, 4, 5] and list(ds.save_epochs) == [2, 3]
        assert float(ds.clip_grad) == 5.0 and np.isinf(ds.learning_rate_decay_epoch) and np.allclose(ds.loss_weights_refine, [0.3, 0.7])
        assert ds.upsampling_method in (b"nearest", "nearest")
    finally:
        ds.close()


def test_oracle_per_pixel_error_field():
    # src/data.jl:323-326: a 4-D obs_err_std divides both channels pixel by pixel; Inf (missing error) zeroes them
    from oracle import data as OD
    rng = np.random.default_rng(5)
    W, H, V, N = 6, 5, 2, 4
    data = (rng.standard_normal((W, H, V, N)) + 10).astype(np.float32)
    data[rng.random((W, H, V, N)) < 0.3] = np.nan
    err = rng.uniform(0.5, 2.0, (W, H, V, N)).astype(np.float32)
    err[0, 0, 0, :] = np.inf
    od = OD.NCData(np.arange(W), np.arange(H), 11000.0 + np.arange(N), data, obs_err_std=err, ntime_win=3)
    d = data - od.meandata[:, :, :, None]
    ok = ~np.isnan(d)
    assert np.allclose(od.x[..., 0][ok], (d / err ** 2)[ok]) and np.allclose(od.x[..., 1][ok], (1 / err ** 2)[ok])
    assert np.all(od.x[0, 0, 0, :, :] == 0) and np.all(od.x[..., 1][~ok] == 0)
