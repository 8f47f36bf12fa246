"""Oracle for the point-cloud path (src/points.jl, ext/CUDAExt.jl).

Positions are 1-based fractional grid indices exactly as in the reference
(src/points.jl:108-121); arrays are indexed [i, j] = [lon, lat] like Julia.
"""
import numpy as np
from . import philox


def interpnd(pos, A):
    """Bilinear gather, src/points.jl:6-34 (CPU) / ext/CUDAExt.jl:8-29 (GPU).

    pos: (P,2) float32 1-based positions; A: (W,H).  float32 arithmetic in the
    reference's order: corners visited as CartesianIndices((0:1,0:1)) =
    (0,0),(1,0),(0,1),(1,1) and accumulated sequentially.
    """
    pos = np.asarray(pos, dtype=np.float32).reshape(-1, 2)
    A = np.asarray(A)
    out = np.zeros(len(pos), dtype=A.dtype)
    ind = np.floor(pos).astype(np.int64)
    c = (pos - ind.astype(np.float32)).astype(np.float32)
    one = np.float32(1)
    for oj in (0, 1):
        for oi in (0, 1):
            cc = (c[:, 0] if oi else one - c[:, 0]) * (c[:, 1] if oj else one - c[:, 1])
            out = out + A[ind[:, 0] - 1 + oi, ind[:, 1] - 1 + oj] * cc.astype(A.dtype)
    return out


def interp_cells(pos):
    """Integer cell (1-based) of every point = floor(pos): the bit-exact part."""
    return np.floor(np.asarray(pos, dtype=np.float32).reshape(-1, 2)).astype(np.int64)


def interp_adjn(pos, values, sz, dtype=None):
    """Adjoint scatter-add onto a zeroed grid, src/points.jl:52-77.

    Sequential over points then corners, like the reference's CPU loop.
    `dtype=np.float64` gives the order-independent sum used to judge the GPU's
    sorted segmented reduction.
    """
    pos = np.asarray(pos, dtype=np.float32).reshape(-1, 2)
    values = np.asarray(values)
    dt = dtype or values.dtype
    A2 = np.zeros(sz, dtype=dt)
    ind = np.floor(pos).astype(np.int64)
    c = (pos - ind.astype(np.float32)).astype(np.float32)
    one = np.float32(1)
    for p in range(len(pos)):
        for oj in (0, 1):
            for oi in (0, 1):
                cc = (c[p, 0] if oi else one - c[p, 0]) * (c[p, 1] if oj else one - c[p, 1])
                A2[ind[p, 0] - 1 + oi, ind[p, 1] - 1 + oj] += dt(values[p]) * dt(cc) if dtype else values[p] * cc
    return A2


def interp_adjn_fast(pos, values, sz, dtype=np.float64):
    """Vectorised (np.add.at) form of interp_adjn for large P."""
    pos = np.asarray(pos, dtype=np.float32).reshape(-1, 2)
    A2 = np.zeros(sz, dtype=dtype)
    ind = np.floor(pos).astype(np.int64)
    c = (pos - ind.astype(np.float32)).astype(np.float32)
    one = np.float32(1)
    for oj in (0, 1):
        for oi in (0, 1):
            cc = (c[:, 0] if oi else one - c[:, 0]) * (c[:, 1] if oj else one - c[:, 1])
            contrib = (np.asarray(values, dtype=np.float32) * cc).astype(dtype)
            np.add.at(A2, (ind[:, 0] - 1 + oi, ind[:, 1] - 1 + oj), contrib)
    return A2


def posindices(grid, lon, lat):
    """src/points.jl:98-134 for one time step.  grid = (gx, gy) coordinate vectors.

    pos = length(grid)*(coord-grid[1])/(grid[end]-grid[1]) + 1  (uses length,
    not length-1: SURVEY T15); inside = 1 <= pos < sz for every coordinate.
    Evaluated in Float64 (the grids of the reference's callers are Float64
    ranges, test/test_reconstruct_points.jl:27-31) from Float32 coordinates;
    the inside test is on the Float64 value, pos is then stored as Float32.
    """
    out = []
    inside = np.ones(len(lon), dtype=bool)
    for g, co in ((grid[0], lon), (grid[1], lat)):
        g = np.asarray(g, dtype=np.float64)
        co = np.asarray(co, dtype=np.float32).astype(np.float64)
        p = len(g) * (co - g[0]) / (g[-1] - g[0]) + 1.0
        inside &= (p >= 1) & (p < len(g))
        out.append(p.astype(np.float32))
    pos = np.stack(out, axis=1)
    return pos, inside


def pertpos(pos, jitter_std_pos, maxpos, noise):
    """src/points.jl:229-235 with externally supplied N(0,1) draws `noise` (P,2)."""
    pos = np.asarray(pos, dtype=np.float32)
    pp = pos + np.asarray(jitter_std_pos, dtype=np.float32)[None, :] * np.asarray(noise, dtype=np.float32)
    lo = np.float32(1)
    hi = np.asarray(maxpos, dtype=np.float32)[None, :]
    return np.minimum(np.maximum(pp, lo), hi).astype(np.float32)


def pos_noise(P, seed, sample_ctr, localn):
    """The Philox draws the CUDA gridding kernel uses for position jitter.

    element e = 2*p + l (l = coordinate) of stream (c1=localn, c2/c3 = sample counter).
    """
    z = philox.normal_field(2 * P, np.uint32(localn), np.uint32(sample_ctr & 0xFFFFFFFF),
                            np.uint32((sample_ctr >> 32) & 0xFFFFFFFF) | np.uint32(0x80000000), seed)
    return z.reshape(P, 2)


class PointCloud:
    """src/points.jl:137-227 (construction) with host-precomputed day-of-year."""

    def __init__(self, grid, lon, lat, values, ids, dayofyear, ntime_win=9,
                 probability_skip_for_training=0.2, jitter_std_pos=(5.0, 5.0),
                 auxdata=None, train=True):
        self.grid = grid
        self.sz = (len(grid[0]), len(grid[1]))
        ntime = len(values)
        self.pos, self.x, self.id = [], [], []
        for n in range(ntime):
            p, inside = posindices(grid, lon[n], lat[n])
            v = np.asarray(values[n], dtype=np.float32)[inside]
            self.pos.append(p[inside])
            xx = np.zeros((len(v), 2), dtype=np.float32)
            xx[:, 0] = np.where(np.isnan(v), 0, v)          # src/points.jl:188
            xx[:, 1] = 1 - np.isnan(v).astype(np.float32)   # src/points.jl:189 (obs_err_std = 1)
            self.x.append(xx)
            self.id.append(np.asarray(ids[n], dtype=np.int64)[inside])
        self.grid_scaled = tuple(
            (2 * (np.asarray(g, np.float64) - np.min(g)) / (np.max(g) - np.min(g)) - 1).astype(np.float32)
            for g in grid)
        d = np.asarray(dayofyear, dtype=np.float64)
        self.dayofyear_cos = np.cos(2 * np.pi * d / 365.25).astype(np.float32)
        self.dayofyear_sin = np.sin(2 * np.pi * d / 365.25).astype(np.float32)
        self.ntime_win = ntime_win
        self.pskip = probability_skip_for_training
        self.jitter_std_pos = jitter_std_pos
        self.train = train
        W, H = self.sz
        self.auxdata = (np.zeros((W, H, 0, ntime), np.float32) if auxdata is None
                        else np.asarray(auxdata, np.float32))

    def sizex(self):
        return (self.sz[0], self.sz[1], 4 + 2 * self.ntime_win + self.auxdata.shape[2])

    def getxy(self, ind, seed=0, sample_ctr=0, track_keep=None, dtype=np.float64):
        """src/points.jl:237-313; `ind` 0-based.  track_keep: bool per point of
        step `ind` (the externally drawn track dropout, integer work); None keeps all."""
        W, H = self.sz
        ntime = len(self.x)
        naux = self.auxdata.shape[2]
        C = 4 + naux + 2 * self.ntime_win
        xin = np.zeros((W, H, C), dtype=np.float32)
        xin[:, :, 0] = self.grid_scaled[0][:, None]
        xin[:, :, 1] = self.grid_scaled[1][None, :]
        xin[:, :, 2] = self.dayofyear_cos[ind]
        xin[:, :, 3] = self.dayofyear_sin[ind]
        for l in range(naux):
            xin[:, :, 4 + l] = self.auxdata[:, :, l, ind]
        half = (self.ntime_win - 1) // 2
        centraln = half
        for localn in range(self.ntime_win):
            n = min(max(ind - half + localn, 0), ntime - 1)
            off = 4 + naux + 2 * localn
            p = self.pos[n]
            if self.train:
                p = pertpos(p, self.jitter_std_pos, (W - 1, H - 1),
                            pos_noise(len(p), seed, sample_ctr, localn))
            x = self.x[n]
            if localn == centraln and self.train and track_keep is not None:
                sel = np.asarray(track_keep, dtype=bool)
                p, x = p[sel], x[sel]
            xin[:, :, off] = interp_adjn_fast(p, x[:, 0], (W, H), dtype).astype(np.float32)
            xin[:, :, off + 1] = interp_adjn_fast(p, x[:, 1], (W, H), dtype).astype(np.float32)
        return xin, (self.pos[ind], self.x[ind])
