"""NetCDF-3 input loader and reconstruction writer: the callers and data formats on
either side of the hot path, host side only.

  load_gridded_nc   src/data.jl:34-98    (lon, lat, time, data with NaN, error, missingmask, mask)
  load_aux_data     src/data.jl:124-146  (time-mean removal + normalize2)
  ncsetup / write   src/data.jl:609-651, :712-769 (file layout, fill value, running mean bookkeeping through `count`)

Only the classic and 64-bit-offset NetCDF formats are handled (scipy.io.netcdf_file); the reference WRITES
`:netcdf3_64bit_offset` (src/data.jl:720), so its outputs are read and reproduced exactly; NetCDF-4/HDF5 inputs are
refused with a clear error because no HDF5 reader exists in this image.
"""
import datetime as _dt
import os
import re

import numpy as np
from scipy.io import netcdf_file

FILL_VALUE = -9999.0


def _days_since_1970(values, units):
    """CF time -> days since 1970-01-01 (what time_cos_sin expects, src/data.jl:286-297)."""
    m = re.match(r"\s*(\w+)\s+since\s+(\d{1,4})-(\d{1,2})-(\d{1,2})(?:[ T](\d{1,2}):(\d{1,2}):(\d{1,2}(?:\.\d*)?))?", units)
    if not m:
        raise ValueError("unsupported time units %r" % units)
    unit = m.group(1).lower()
    scale = {"days": 1.0, "day": 1.0, "hours": 1 / 24.0, "hour": 1 / 24.0, "minutes": 1 / 1440.0, "seconds": 1 / 86400.0}[unit]
    y, mo, d = int(m.group(2)), int(m.group(3)), int(m.group(4))
    hh, mi, ss = (int(m.group(5) or 0), int(m.group(6) or 0), float(m.group(7) or 0.0))
    origin = _dt.datetime(y, mo, d, hh, mi) + _dt.timedelta(seconds=ss)
    off = (origin - _dt.datetime(1970, 1, 1)).total_seconds() / 86400.0
    return np.asarray(values, np.float64) * scale + off


def _open(fname):
    with open(fname, "rb") as f:
        magic = f.read(4)
    if magic[:3] != b"CDF":
        raise ValueError("%s is not a NetCDF-3 (classic / 64-bit offset) file%s" %
                         (fname, " (NetCDF-4/HDF5 cannot be read here)" if magic == b"\x89HDF" else ""))
    return netcdf_file(fname, "r", mmap=False)


def _masked(var, fill_to):
    """nomissing(var, fill_to): scale/offset and _FillValue / missing_value handling as NCDatasets does."""
    a = np.array(var[...], copy=True)
    out = a.astype(np.float32)
    bad = np.zeros(a.shape, bool)
    for att in ("_FillValue", "missing_value"):
        if hasattr(var, att):
            bad |= a == np.asarray(getattr(var, att)).reshape(-1)[0]
    if hasattr(var, "scale_factor"):
        out = out * np.float32(var.scale_factor)
    if hasattr(var, "add_offset"):
        out = out + np.float32(var.add_offset)
    out[bad] = fill_to
    return out


def load_gridded_nc(fname, varname, errvarname=None, minfrac=0.05, obs_err_std=1.0):
    """lon, lat, time (days since 1970), data (W,H,N) with NaN, error (W,H,N), missingmask (W,H,N), mask (W,H)
    -- src/data.jl:34-98.  Arrays come back in the reference's (lon, lat, time) index order."""
    ds = _open(fname)
    try:
        lon = np.array(ds.variables["lon"][...], np.float64)
        lat = np.array(ds.variables["lat"][...], np.float64)
        tv = ds.variables["time"]
        time = _days_since_1970(tv[...], tv.units.decode() if isinstance(tv.units, bytes) else tv.units)
        data = _masked(ds.variables[varname], np.nan).transpose(2, 1, 0)          # file (time,lat,lon) -> (lon,lat,time)
        if "mask" in ds.variables:
            mask = np.array(ds.variables["mask"][...]).T == 1
        else:
            missingfraction = np.isnan(data).mean(axis=2)
            mask = missingfraction <= 1 - minfrac
        if errvarname is not None:
            err = _masked(ds.variables[errvarname], np.inf).transpose(2, 1, 0)
        else:
            err = np.full(data.shape, obs_err_std, np.float32)
    finally:
        ds.close()
    return lon, lat, time, np.ascontiguousarray(data), np.ascontiguousarray(err), np.isnan(data), mask


def normalize2(data):
    """src/data.jl:113-121 (std with the n-1 denominator, as Julia's `std`)."""
    data = np.asarray(data, np.float64)
    s, m = data.std(ddof=1), data.mean()
    return (data - m) if s == 0 else (data - m) / s


def load_aux_data(sz, auxdata_files):
    """src/data.jl:124-146: per file two channels, normalize2(data/err^2) and normalize2(1/err^2), data with its
    per-time-step spatial mean removed.  auxdata_files: [{filename, varname, errvarname (optional)}]."""
    W, H, N = sz
    aux = np.zeros((W, H, 2 * len(auxdata_files), N), np.float32)
    for i, f in enumerate(auxdata_files):
        ds = _open(f["filename"])
        try:
            data = _masked(ds.variables[f["varname"]], np.nan).transpose(2, 1, 0).astype(np.float64)
            ev = f.get("errvarname")
            err = np.ones(data.shape) if ev is None else _masked(ds.variables[ev], np.nan).transpose(2, 1, 0).astype(np.float64)
        finally:
            ds.close()
        data = data - data.mean(axis=(0, 1), keepdims=True)
        aux[:, :, 2 * i, :] = normalize2(data / err ** 2)
        aux[:, :, 2 * i + 1, :] = normalize2(1 / err ** 2)
    return aux


def write_reconstruction(fname, varnames, lon, lat, meandata, mean, sigma, count):
    """The file DINCAE.ncsetup + savesample leave behind after `count` save epochs (src/data.jl:609-651, :712-769):
    NetCDF-3 64-bit offset; dims time (unlimited), lon, lat; `var`, `var_error` (lon,lat,time in Julia order = (time,lat,lon) on
    disk), `var_mean` (lon,lat), all Float32 with _FillValue -9999; global attribute `count`.

    mean / sigma: (W,H,N) arrays or lists of them per variable, already averaged over the save epochs and already holding
    the fill value where the reference writes it (dincae_reconstruct_range does both); meandata: (W,H,nvar) or (W,H)."""
    if os.path.exists(fname):
        os.remove(fname)
    if isinstance(mean, np.ndarray):
        mean, sigma = [mean], [sigma]
    meandata = np.asarray(meandata, np.float32)
    if meandata.ndim == 2:
        meandata = meandata[:, :, None]
    W, H, N = mean[0].shape
    ds = netcdf_file(fname, "w", version=2)
    try:
        ds.createDimension("time", None)
        ds.createDimension("lon", W)
        ds.createDimension("lat", H)
        vlon = ds.createVariable("lon", "f", ("lon",))
        vlat = ds.createVariable("lat", "f", ("lat",))
        recs = []
        for i, name in enumerate(varnames):
            vm = ds.createVariable(name, "f", ("time", "lat", "lon")); vm._FillValue = np.float32(FILL_VALUE)
            ve = ds.createVariable(name + "_error", "f", ("time", "lat", "lon")); ve._FillValue = np.float32(FILL_VALUE)
            vmean = ds.createVariable(name + "_mean", "f", ("lat", "lon")); vmean._FillValue = np.float32(FILL_VALUE)
            md = meandata[:, :, i].astype(np.float32).copy()
            md[np.isnan(md)] = FILL_VALUE
            vmean[:, :] = md.T
            recs.append((vm, ve, i))
        vlon[:] = np.asarray(lon, np.float32)
        vlat[:] = np.asarray(lat, np.float32)
        for vm, ve, i in recs:
            vm[0:N, :, :] = np.nan_to_num(np.asarray(mean[i], np.float32).transpose(2, 1, 0), nan=FILL_VALUE)
            ve[0:N, :, :] = np.nan_to_num(np.asarray(sigma[i], np.float32).transpose(2, 1, 0), nan=FILL_VALUE)
        ds.count = np.int32(count)
    finally:
        ds.close()
    return fname


def load_ragged_points(fname, varname):
    """DINCAE.loaddata (src/points.jl:429-446): a CF contiguous-ragged-array file (a count variable with the attribute
    sample_dimension, here `size(time_instances)`; `varname`, lon, lat, dtime, id along the sample dimension; `dates` per
    instance) -> values, (lon, lat), dtime, id as lists of per-instance vectors, dates as days since 1970 and the day of the
    year of every instance (what PointCloud derives with Dates.dayofyear, src/points.jl:197-202)."""
    ds = _open(fname)
    try:
        count = None
        for name, v in ds.variables.items():
            if hasattr(v, "sample_dimension"):
                sd = v.sample_dimension
                sd = sd.decode() if isinstance(sd, bytes) else sd
                if ds.variables[varname].dimensions == (sd,):
                    count = np.array(v[...], np.int64)
        if count is None:
            raise ValueError("%s: no count variable with sample_dimension = the dimension of %r" % (fname, varname))
        ends = np.cumsum(count); starts = ends - count

        def ragged(name, dtype, fill_nan):
            var = ds.variables[name]
            a = _masked(var, np.nan) if fill_nan else np.array(var[...])
            a = a.astype(dtype)
            return [a[s:e].copy() for s, e in zip(starts, ends)]
        values = ragged(varname, np.float32, True)
        lon = ragged("lon", np.float32, True); lat = ragged("lat", np.float32, True)
        dv = ds.variables["dtime"]
        du = dv.units.decode() if isinstance(dv.units, bytes) else dv.units
        dt_all = _days_since_1970(np.array(dv[...], np.float64), du)
        dtime = [dt_all[s:e].copy() for s, e in zip(starts, ends)]
        ids = ragged("id", np.int64, False)
        tv = ds.variables["dates"]
        dates = _days_since_1970(np.array(tv[...], np.float64), tv.units.decode() if isinstance(tv.units, bytes) else tv.units)
    finally:
        ds.close()
    epoch = _dt.datetime(1970, 1, 1)
    doy = np.array([(epoch + _dt.timedelta(days=float(d))).timetuple().tm_yday for d in dates], np.float64)
    return values, (lon, lat), dtime, ids, dates, doy


def write_paramfile(fname, losses, attrs):
    """The parameter file of reconstruct / reconstruct_points (src/model.jl:655-676, src/points.jl:658-692): variable
    `losses(epochs)` plus one global attribute per keyword.  Booleans are written as Int, sequences as vectors, strings as text."""
    if os.path.exists(fname):
        os.remove(fname)
    ds = netcdf_file(fname, "w", version=2)
    try:
        ds.createDimension("epochs", len(losses))
        v = ds.createVariable("losses", "d", ("epochs",))
        v[:] = np.asarray(losses, np.float64)
        for k, val in attrs.items():
            if isinstance(val, bool):
                val = np.int32(val)
            elif isinstance(val, str):
                pass
            elif isinstance(val, (list, tuple, range, np.ndarray)):
                val = list(val)
                if len(val) and all(isinstance(x, str) for x in val):
                    val = ",".join(val)
                elif all(float(x).is_integer() for x in val) and not any(isinstance(x, float) for x in val):
                    val = np.asarray(val, np.int32)
                else:
                    val = np.asarray(val, np.float64)
            elif isinstance(val, (int, np.integer)):
                val = np.int32(val)
            else:
                val = np.float64(val)
            setattr(ds, k, val)
    finally:
        ds.close()
    return fname
