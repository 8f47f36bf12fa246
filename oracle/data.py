"""Oracle for the gridded data path (src/data.jl).

Arrays are indexed like Julia ([i,j,...] = [lon,lat,...]) and stored
Fortran-ordered so that `.ravel(order="F")` is the reference's memory order
(WHCN), the layout the C-ABI takes unchanged.
"""
import numpy as np
from . import philox


def time_cos_sin(days_since_origin, cycle_period=365.25):
    """src/data.jl:289-297.  NOTE: the reference divides the
    millisecond count by 1000*60*60, so `t` is in HOURS although the comment
    says days; restated literally."""
    t = np.asarray(days_since_origin, dtype=np.float64) * 24.0
    return (np.cos(2 * np.pi * t / cycle_period).astype(np.float32),
            np.sin(2 * np.pi * t / cycle_period).astype(np.float32))


def normalize2(data):
    """src/data.jl:111-119 (std with the n-1 denominator, as Statistics.std)."""
    data = np.asarray(data, dtype=np.float64)
    s = np.std(data, ddof=1) if data.size > 1 else 0.0
    m = np.mean(data)
    return data - m if s == 0 else (data - m) / s


class NCData:
    """src/data.jl:148-170 + ctor :250-369 (2-D, scalar variables only)."""

    def __init__(self, lon, lat, days, data_full, train=False, obs_err_std=None,
                 jitter_std=None, ntime_win=3, isoutput=None, remove_mean=True,
                 mean_data=None, mask=None):
        data_full = np.asfortranarray(data_full, dtype=np.float32)
        W, H, V, N = data_full.shape
        self.shape = (W, H, V, N)
        self.missingmask = np.isnan(data_full[:, :, 0, :])            # src/data.jl:91
        if remove_mean:
            if mean_data is None:                                     # :266-270 sequential f32 sum over time
                s = np.zeros((W, H, V), np.float32)
                cnt = np.zeros((W, H, V), np.float32)
                for n in range(N):
                    d = data_full[:, :, :, n]
                    nan = np.isnan(d)
                    s = (s + np.where(nan, np.float32(0), d)).astype(np.float32)
                    cnt = cnt + (~nan)
                with np.errstate(invalid="ignore", divide="ignore"):
                    meandata = (s / cnt.astype(np.float32)).astype(np.float32)
            else:
                meandata = np.asarray(mean_data, np.float32)
        else:
            meandata = np.zeros((W, H, V), np.float32)
        self.meandata = meandata
        obs_err_std = np.ones(V, np.float32) if obs_err_std is None else np.asarray(obs_err_std, np.float32)
        self.obs_err_std = obs_err_std
        data = (data_full - meandata[:, :, :, None]).astype(np.float32)   # :299
        nan = np.isnan(data)
        x = np.zeros((W, H, V, N, 2), np.float32, order="F")
        if obs_err_std.ndim == 1:                                         # :318-322 one value per variable
            s2 = (obs_err_std * obs_err_std).astype(np.float32)[None, None, :, None]
        else:                                                             # :323-326 per-pixel error field (errvarname), Inf = missing
            s2 = (obs_err_std * obs_err_std).astype(np.float32)
        x[..., 0] = np.where(nan, np.float32(0), data) / s2               # :315,:321
        x[..., 1] = (1 - nan.astype(np.float32)) / s2                     # :316,:322
        self.x = x
        lon = np.asarray(lon, np.float64); lat = np.asarray(lat, np.float64)
        self.lon_scaled = (2 * (lon - lon.min()) / (lon.max() - lon.min()) - 1).astype(np.float32)   # :342
        self.lat_scaled = (2 * (lat - lat.min()) / (lat.max() - lat.min()) - 1).astype(np.float32)   # :343
        self.time_cos, self.time_sin = time_cos_sin(days)
        self.train = train
        self.jitter_std = (np.full(V, 0.05, np.float32) if jitter_std is None
                           else np.asarray(jitter_std, np.float32))
        self.ntime_win = ntime_win
        self.isoutput = (np.arange(V) == 0) if isoutput is None else np.asarray(isoutput, bool)
        self.mask = np.ones((W, H), bool) if mask is None else np.asarray(mask, bool)

    def sizex(self):
        W, H, V, N = self.shape
        return (W, H, 4 + 2 * V * self.ntime_win)              # src/data.jl:199

    def sizey(self):
        W, H, V, N = self.shape
        return (W, H, 2 * int(self.isoutput.sum()))            # src/data.jl:208

    def window(self, ind):
        """src/data.jl:405 (0-based)."""
        N = self.shape[3]
        half = (self.ntime_win - 1) // 2
        return [min(max(ind - half + k, 0), N - 1) for k in range(self.ntime_win)]

    def getxy(self, ind, imask=None, seed=0, sample_ctr=0):
        """src/data.jl:391-517.  `ind`, `imask` 0-based.  imask=None => no perturbation
        (train=false).  Jitter = jitter_std[v]*Philox normal, added AFTER zeroing and at
        every pixel."""
        W, H, V, N = self.shape
        T = self.ntime_win
        xin = np.zeros((W, H, 4 + 2 * V * T), np.float32, order="F")
        xin[:, :, 0] = self.lon_scaled[:, None]
        xin[:, :, 1] = self.lat_scaled[None, :]
        xin[:, :, 2] = self.time_cos[ind]
        xin[:, :, 3] = self.time_sin[ind]
        nrange = self.window(ind)
        centraln = (T + 1) // 2 - 1
        off = 4
        for v in range(V):
            for localn, n in enumerate(nrange):
                xin[:, :, off] = self.x[:, :, v, n, 0]
                xin[:, :, off + 1] = self.x[:, :, v, n, 1]
                if imask is not None:
                    if localn == centraln:
                        mm = self.missingmask[:, :, imask]
                        xin[:, :, off][mm] = 0
                        xin[:, :, off + 1][mm] = 0
                    z = jitter_noise(W * H, v * T + localn, seed, sample_ctr).reshape((W, H), order="F")
                    xin[:, :, off] += self.jitter_std[v] * z
                off += 2
        xtrue = np.zeros(self.sizey(), np.float32, order="F")
        o = 0
        for v in np.flatnonzero(self.isoutput):
            xtrue[:, :, o] = self.x[:, :, v, ind, 0]
            xtrue[:, :, o + 1] = self.x[:, :, v, ind, 1]
            o += 2
        return xin, xtrue

    def batch(self, time_idx, imask=None, seed=0, sample_ctr0=0):
        """One mini-batch as DataBatches.iterate assembles it (src/data.jl:831-868)."""
        xs, ys = [], []
        for b, ind in enumerate(time_idx):
            xi, yi = self.getxy(int(ind), None if imask is None else int(imask[b]), seed, sample_ctr0 + b)
            xs.append(xi); ys.append(yi)
        return (np.asfortranarray(np.stack(xs, axis=3)), np.asfortranarray(np.stack(ys, axis=3)))


def jitter_noise(npix, slot, seed, sample_ctr):
    """Philox draws of the assembly kernel: element = pixel (i + W*j), stream =
    (slot = v*T+localn, sample counter)."""
    return philox.normal_field(npix, np.uint32(slot), np.uint32(sample_ctr & 0xFFFFFFFF),
                               np.uint32((sample_ctr >> 32) & 0x7FFFFFFF), seed)


def savesample_accumulate(old_m, old_s, xrec, meandata, mask, count, fill_value=-9999.0):
    """The arithmetic of savesample (src/data.jl:615-619, :653-675) for one output
    variable.  xrec: (W,H,2,B) = (m, sigma2); old_*: (W,H,B) current file content
    (zeros when count==1).  Returns the new (m, sigma) slices."""
    m = xrec[:, :, 0, :]
    s2 = xrec[:, :, 1, :]
    rec = (m + meandata[:, :, None]).astype(np.float32)
    with np.errstate(invalid="ignore"):
        sig = np.sqrt(s2).astype(np.float32)
    sig[np.isnan(rec)] = np.nan
    rec = rec.copy()
    rec[~mask, :] = np.nan
    sig[~mask, :] = np.nan

    def acc(old, new):
        v = ((np.float32(count - 1) * old.astype(np.float32) + new.astype(np.float32)) / np.float32(count))
        v = np.where(np.isnan(v), np.float32(fill_value), v)      # Float32 arithmetic, src/data.jl:616-618
        return v.astype(np.float32)
    return acc(old_m, rec), acc(old_s, sig)
