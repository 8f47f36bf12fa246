"""cfg4-shaped timing (BASELINE configs[3]): point-cloud training on the 177x69 grid, T=27, F=(25,50,75), 1 aux cube, Laplacian
penalty, batch 32, ~2000 points per time step.  Times batch assembly (gridding on the device) and the training step."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
grid = (np.arange(-7, 37.001, 0.25), np.arange(29, 46.001, 0.25))        # 177 x 69
W, H = len(grid[0]), len(grid[1])
N, T, B, ntracks, npts = 128, 27, 32, 6, 350
lon, lat, val, ids = [], [], [], []
for n in range(N):
    lo, la, va, ii = [], [], [], []
    for t in range(ntracks):
        x0, x1 = rng.uniform(grid[0][0], grid[0][-1], 2); y0, y1 = rng.uniform(grid[1][0], grid[1][-1], 2)
        s = np.linspace(0, 1, npts)
        lo.append(x0 + (x1 - x0) * s); la.append(y0 + (y1 - y0) * s)
        va.append(np.sin(0.3 * lo[-1]) * np.cos(0.2 * la[-1]) + 0.02 * rng.standard_normal(npts)); ii.append(np.full(npts, t + 1))
    lon.append(np.concatenate(lo).astype(np.float32)); lat.append(np.concatenate(la).astype(np.float32))
    val.append(np.concatenate(va).astype(np.float32)); ids.append(np.concatenate(ii))
aux = rng.standard_normal((W, H, 2, N)).astype(np.float32)
ds = pkg.PointDataset(ctx, grid, lon, lat, val, ids, 50 + np.arange(N), ntime_win=T, probability_skip_for_training=0.877,
                      jitter_std_pos=(0.1715, 0.1715), auxdata=aux, train=True)
Wd, Hd, C = ds.sizex()
for prec, DEVICE_BATCH in (("tf32", False), ("tf32", True), ("fp32", True)):
    cfg = pkg.ModelConfig(W=Wd, H=Hd, ninput=C, enc_nfilter_internal=(25, 50, 75), regularization_L2_beta=1e-4, laplacian_penalty=0.1,
                          max_batch=B, precision=prec)
    model = pkg.Model(ctx, cfg, seed=1)
    import torch
    xbuf = torch.empty(Wd * Hd * C * B, dtype=torch.float32, device="cuda") if DEVICE_BATCH else None
    def step(k):
        idx = rng.integers(0, N, B)
        keep = [ds.draw_track_keep(int(i), rng) for i in idx]
        t0 = time.perf_counter()
        xin, truth = ds.assemble(idx, seed=3, sample_ctr0=k * B, keep=keep, xin=xbuf)
        t1 = time.perf_counter()
        loss = model.train_step_points(xin, truth, 1e-3)
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1, loss
    for k in range(3): step(k)
    ta = ts = 0.0; n = 20
    for k in range(n):
        a, s, loss = step(3 + k); ta += a; ts += s
    print("%s device_batch=%s: Cin=%d, assembly %.2f ms/batch, train step %.2f ms/batch -> %.0f samples/s end to end, loss %.4f"
          % (prec, DEVICE_BATCH, C, ta / n * 1e3, ts / n * 1e3, B * n / (ta + ts), loss))

# per-kernel CUDA-event profile of three TF32 steps (library profile hooks)
import ctypes, json
lib = pkg.lib.load()
cfg = pkg.ModelConfig(W=Wd, H=Hd, ninput=C, enc_nfilter_internal=(25, 50, 75), regularization_L2_beta=1e-4, laplacian_penalty=0.1, max_batch=B, precision="tf32")
model = pkg.Model(ctx, cfg, seed=1)
import torch
xbuf = torch.empty(Wd * Hd * C * B, dtype=torch.float32, device="cuda")
def one(k):
    idx = rng.integers(0, N, B)
    keep = [ds.draw_track_keep(int(i), rng) for i in idx]
    xin, truth = ds.assemble(idx, seed=3, sample_ctr0=k * B, keep=keep, xin=xbuf)
    return model.train_step_points(xin, truth, 1e-3)
one(0)
pkg.lib.check(lib.dincae_profile_begin(ctx.handle))
t0 = time.perf_counter()
for k in range(3): one(1 + k)
wall = (time.perf_counter() - t0) / 3
buf = ctypes.create_string_buffer(1 << 16)
pkg.lib.check(lib.dincae_profile_end(ctx.handle, buf, len(buf)))
tab = json.loads(buf.value.decode())
tab.sort(key=lambda k: -k["ms"])
print("profiled step wall %.2f ms; kernel total %.2f ms/step" % (wall * 1e3, sum(k["ms"] for k in tab) / 3))
for k in tab[:14]:
    print("  %-28s %3d launches/step %7.3f ms/step" % (k["kernel"], k["launches"] // 3, k["ms"] / 3))
