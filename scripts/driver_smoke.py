"""cfg3-shaped end-to-end run of the reconstruct driver (two variables, refinement, L2; 112x112, 256 time steps, 2 epochs) in the
default TF32 mode: losses must fall and the reconstruction must be finite on sea points and carry the fill value on land."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
W = H = 112; N = 256
lon, lat, days = np.linspace(-7, -0.8, W), np.linspace(33.8, 38.2, H), 11000.0 + np.arange(N)
base = 15 + 8 * np.cos(2 * np.pi * days / 365.25)[None, None, :] + np.cumsum(rng.standard_normal((W, H, N)), axis=0) * 0.05
sst = base.astype(np.float32); chl = (0.5 * base + rng.standard_normal((W, H, N))).astype(np.float32)
mask = np.ones((W, H), bool); mask[:10, :20] = False
for a in (sst, chl):
    a[rng.random((W, H, N)) < 0.5] = np.nan
    a[~mask] = np.nan
entry = [dict(lon=lon, lat=lat, time=days, data=sst, mask=mask, varname="SST"),
         dict(lon=lon, lat=lat, time=days, data=chl, mask=mask, varname="CHL", isoutput=False, jitter_std=0.1)]
t0 = time.perf_counter()
losses, res = pkg.reconstruct(ctx, [entry, entry], None, epochs=3, batch_size=32, enc_nfilter_internal=(32, 64, 128, 256), save_epochs=[2, 3],
                              ntime_win=3, loss_weights_refine=(0.3, 0.7), regularization_L2_beta=1e-3, seed=1, precision="tf32")
dt = time.perf_counter() - t0
m, s = res[0]["mean"][:, :, :, 0], res[0]["sigma"][:, :, :, 0]
print("losses", [round(x, 4) for x in losses], "count", res[0]["count"], "wall %.2f s" % dt)
print("land fill ok:", bool(np.all(m[~mask] == -9999.0)), " sea finite:", bool(np.isfinite(m[mask]).all() and (s[mask] > 0).all()),
      " rms(rec - truth) on observed sea pixels: %.3f" % np.sqrt(np.nanmean((m[mask] - sst[mask]) ** 2)))
assert losses[-1] < losses[0] and np.all(m[~mask] == -9999.0) and np.isfinite(m[mask]).all()
