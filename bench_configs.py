"""Short measured runs of the BASELINE.json configs other than the headline (bench.py imports this; results land in the
bench line's `configs` block).  Every config is timed like the headline: warm-up, blocks of exactly K steps between a barrier +
synchronize, CUDA events on the library's stream, max over ranks, median over blocks; inputs resident in HBM (cubes are
generated on the device).

  cfg1  configs[0]  tutorial SST 112^2, one chain, batch 32 per GPU (the reference's CPU-runnable case, here on the GPU)
  cfg3  configs[2]  SST + auxiliary field (V = 2 => 16/18 input channels), refinement, data-parallel: weak (32 per GPU) AND the
                    strong-scaling point at a fixed global batch of 256
  cfg4  configs[3]  reconstruct_points: 177 x 69 grid, ntime_win = 27, [25,50,75] filters, Laplacian penalty, ~2000 points per step
  cfg5  configs[4]  1024^2, 5 levels (32..512), refinement, global batch 64 (N >= 2; one GPU holds 32) and the time-sharded
                    reconstruction of 10 000 fields (N >= 2; one GPU does a 1024-field sample)
"""
import os
import time

import numpy as np


def _median_block(timer, fn, steps, repeats):
    ms, _, allms = timer.median(fn, steps, repeats)
    return ms / steps, allms


def _top_kernels(pkg, ctx, fn, rank):
    """Per-launch CUDA-event profile of one step; kernels >= 5 % of it (their rooflines are filled in by finalize())."""
    import bench
    tab = bench.kernel_profile(pkg, pkg.lib.load(), ctx, fn, 1, rank)
    if rank != 0:
        return None
    return [k for k in tab if k["share"] >= 0.05]


def finalize(configs, roofs, world):
    """Roofline fractions once the roofs are known (the TF32 roof is measured at the very end of the run)."""
    import bench
    for c in configs.values():
        c["frac_of_tf32_roof"] = c["achieved_tflops"] / world / roofs["tf32_tflops"]
        if "recon" in c:
            c["recon"]["frac_of_tf32_roof"] = c["recon"]["achieved_tflops"] / world / roofs["tf32_tflops"]
        tab = c.pop("_kernels", None)
        if tab is not None:
            c["rooflines"] = [{kk: (round(v, 4) if isinstance(v, float) else v) for kk, v in bench.roofline_of(k, roofs, None).items() if kk != "traffic"}
                              for k in tab]


def gridded_config(pkg, ctx, timer, torch, dev, rank, world, precision, *, name, W, H, V, T, F, refine, l2, batch_per_gpu,
                   ncube, steps, repeats, recon_fields, recon_seed_per_rank=False, isoutput=None, note=None):
    import bench
    B = batch_per_gpu
    lon, lat, days, cube, sea = bench.synth_cube_device(W, H, V, ncube, dev, seed=11)          # same cube on every rank
    ds = pkg.GriddedDataset(ctx, lon, lat, days, cube, train=True, ntime_win=T, mask=sea, isoutput=isoutput)
    del cube
    cin = 4 + 2 * T * V
    cfg = pkg.ModelConfig(W=W, H=H, ninput=cin, enc_nfilter_internal=F, loss_weights_refine=refine, regularization_L2_beta=l2,
                          max_batch=B, precision=precision)
    model = pkg.Model(ctx, cfg, seed=1)
    counts = ds.valid_counts(0)
    npix = W * H
    rng = np.random.default_rng(99)
    k = [0]

    def step():
        idx = rng.integers(0, ncube, B * world); imask = rng.integers(0, ncube, B * world)
        lo = rank * B
        ctr0 = k[0] * B * world + lo
        k[0] += 1
        nglob = None if world == 1 else np.array([counts[idx].sum()], np.int64)
        return model.train_step_gridded(ds, idx[lo:lo + B], imask[lo:lo + B], 1e-3, seed=7, sample_ctr0=ctr0, n_valid_global=nglob, want_loss=False)
    for _ in range(2):
        step()
    ms_step, blocks = _median_block(timer, step, steps, repeats)
    fl = bench.train_flops_per_sample(W, H, cin, 1, F, len(refine))
    sps = B * world / (ms_step * 1e-3)
    out = {"workload": name, "W": W, "H": H, "ninput": cin, "filters": list(F), "refine": list(refine), "batch_per_gpu": B, "global_batch": B * world,
           "train_samples_per_s": sps, "ms_per_step": ms_step, "steps": steps, "ms_blocks": blocks,
           "train_gflop_per_sample": fl["train"] / 1e9, "achieved_tflops": sps * fl["train"] / 1e12}
    if note:
        out["note"] = note
    tab = _top_kernels(pkg, ctx, step, rank)
    if rank == 0:
        out["_kernels"] = tab
    div = model.replica_divergence() if world > 1 else None
    if div is not None:
        out["replica_max_abs_diff"] = div
    # ---- reconstruction, time-sharded: this rank's fields only, no communication
    if recon_fields:
        per = recon_fields // world
        if recon_seed_per_rank:                   # a big record: every rank generates just its own shard on the device
            ds.close()
            torch.cuda.empty_cache()
            lon, lat, days, cube, sea = bench.synth_cube_device(W, H, V, per, dev, seed=100 + rank, chunk=64)
            ds = pkg.GriddedDataset(ctx, lon, lat, days, cube, train=False, ntime_win=T, mask=sea, isoutput=isoutput)
            del cube
            torch.cuda.empty_cache()
            t0, t1 = 0, per
        else:
            per = min(per, ncube // world)
            t0, t1 = rank * per, (rank + 1) * per
        m_out = torch.empty(per * npix, dtype=torch.float32, device=dev); s_out = torch.empty_like(m_out)

        def rec():
            pkg.api.reconstruct_range(model, ds, t0, t1, 1, -9999.0, m_out, s_out)
        if per <= 1024:
            rec()
        ms_rec, _ = _median_block(timer, rec, 1, 1 if per > 1024 else 3)
        mpx = per * world * npix / (ms_rec * 1e-3) / 1e6
        out["recon"] = {"value": mpx, "unit": "Mpixel/s", "fields": per * world, "ms": ms_rec,
                        "achieved_tflops": mpx * 1e6 / npix * fl["fwd"] / 1e12}
    model.close(); ds.close()
    torch.cuda.empty_cache()
    return out


def synth_tracks(rng, grid, N, ntracks=6, npts=350):
    gx, gy = grid
    lon, lat, val, ids = [], [], [], []
    for n in range(N):
        lo, la, va, ii = [], [], [], []
        for t in range(ntracks):
            x0, x1 = rng.uniform(gx[0], gx[-1], 2); y0, y1 = rng.uniform(gy[0], gy[-1], 2)
            s = np.linspace(0, 1, npts)
            lo.append(x0 + (x1 - x0) * s); la.append(y0 + (y1 - y0) * s)
            va.append(np.sin(0.3 * lo[-1]) * np.cos(0.2 * la[-1]) + 0.02 * rng.standard_normal(npts)); ii.append(np.full(npts, t + 1))
        lon.append(np.concatenate(lo).astype(np.float32)); lat.append(np.concatenate(la).astype(np.float32))
        val.append(np.concatenate(va).astype(np.float32)); ids.append(np.concatenate(ii))
    return lon, lat, val, ids


def points_config(pkg, ctx, timer, torch, dev, rank, world, precision, cpu):
    """cfg4 (test/test_reconstruct_points.jl shapes): gridding of the scattered observations on the device
    (K4 sorted scatter) + model + point-cloud loss (K4 gather, sorted adjoint) + Adam; data-parallel over the mini-batch."""
    import bench
    grid = (np.arange(-7, 37.001, 0.25), np.arange(29, 46.001, 0.25))        # 177 x 69
    W, H = len(grid[0]), len(grid[1])
    N, T, B, F, lap = 128, 27, 32, (25, 50, 75), 0.1
    rng = np.random.default_rng(0)
    lon, lat, val, ids = synth_tracks(rng, grid, N)
    aux = rng.standard_normal((W, H, 2, N)).astype(np.float32)
    kw = dict(ntime_win=T, probability_skip_for_training=0.877, jitter_std_pos=(0.1715, 0.1715), auxdata=aux, train=True)
    ds = pkg.PointDataset(ctx, grid, lon, lat, val, ids, 50 + np.arange(N), **kw)
    Wd, Hd, C = ds.sizex()
    cfg = pkg.ModelConfig(W=Wd, H=Hd, ninput=C, enc_nfilter_internal=F, regularization_L2_beta=1e-4, laplacian_penalty=lap, max_batch=B, precision=precision)
    model = pkg.Model(ctx, cfg, seed=1)
    xbuf = torch.empty(Wd * Hd * C * B, dtype=torch.float32, device=dev)
    k = [0]

    nvalid_step = np.array([int((x[:, 1] != 0).sum()) for x in ds.x], np.int64)      # valid observations per time step (static)

    def step():
        idx = rng.integers(0, N, B * world)                                      # the same draw on every rank
        lo = rank * B
        # track dropout of this rank's own samples (integer work on the host); one generator per (step, sample) so that no rank
        # has to draw for the others
        keep = [ds.draw_track_keep(int(i), np.random.default_rng([17, k[0], lo + j])) for j, i in enumerate(idx[lo:lo + B])]
        xin, truth = ds.assemble(idx[lo:lo + B], seed=3, sample_ctr0=k[0] * B * world + lo, keep=keep, xin=xbuf)
        k[0] += 1
        if world == 1:
            return model.train_step_points(xin, truth, 1e-3, want_loss=False)
        nglob = int(nvalid_step[idx].sum())                                      # valid observations of the GLOBAL batch
        model.grad_step_points(xin, truth, n_valid_global=nglob)
        model.allreduce_grads()
        return model.apply_step(1e-3, want_loss=False)
    for _ in range(3):
        step()
    ms_step, blocks = _median_block(timer, step, 10, 3)
    fl = bench.train_flops_per_sample(Wd, Hd, C, 1, F, 1)
    sps = B * world / (ms_step * 1e-3)
    npts = float(np.mean([len(p) for p in ds.pos]))
    out = {"workload": "BASELINE configs[3]: reconstruct_points, 177x69 grid, ntime_win=27 (%d input channels), filters [25,50,75], laplacian_penalty=0.1, "
                       "batch %d per GPU, %.0f observations per time step" % (C, B, npts),
           "batch_per_gpu": B, "global_batch": B * world, "train_samples_per_s": sps, "ms_per_step": ms_step, "steps": 10, "ms_blocks": blocks,
           "timed_region": "host track dropout draw + device gridding (27 sorted scatters per sample) + forward + point loss + backward + Adam",
           "train_gflop_per_sample": fl["train"] / 1e9, "achieved_tflops": sps * fl["train"] / 1e12}
    tab = _top_kernels(pkg, ctx, step, rank)
    if rank == 0:
        out["_kernels"] = tab
    model.close(); ds.close()
    if cpu and rank == 0:
        out["cpu_baseline"] = points_cpu_baseline(grid, lon, lat, val, ids, aux, N, T, F, lap)
    torch.cuda.empty_cache()
    return out


def points_cpu_baseline(grid, lon, lat, val, ids, aux, N, T, F, lap, B=8, budget_s=12.0):
    """The oracle port of the same points step on the host cores (bounded sample: mini-batches of 8)."""
    import torch
    from oracle import model as OM, points as OP
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    od = OP.PointCloud(grid, lon, lat, val, ids, 50 + np.arange(N), ntime_win=T, probability_skip_for_training=0.877,
                       jitter_std_pos=(0.1715, 0.1715), auxdata=aux, train=True)
    W, H, C = od.sizex()
    ocfg = OM.ModelConfig(W=W, H=H, ninput=C, nfilter=list(F), regularization_L2_beta=1e-4, laplacian_penalty=lap)
    flat = OM.init_params(ocfg, 1)
    opt = OM.Adam(flat.size)
    rng = np.random.default_rng(0)
    n, used = 0, 0.0
    while True:
        t0 = time.perf_counter()
        idx = rng.integers(0, N, B)
        xs, truth = [], []
        for b, i in enumerate(idx):
            x, tr = od.getxy(int(i), seed=3, sample_ctr=n * B + b, dtype=np.float32)
            xs.append(x); truth.append(tr)
        xin = np.asfortranarray(np.stack(xs, axis=3))
        _, flat = OM.train_step(ocfg, flat, opt, xin, None, 1e-3, dtype=torch.float32, point_truth=truth)
        dt = time.perf_counter() - t0
        if n > 0:
            used += dt
        n += 1
        if (n >= 3 and used > budget_s) or n >= 8:
            break
    return {"value": (n - 1) * B / used, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%d point-cloud train steps of batch %d after 1 warm-up (numpy gridding + torch-CPU float32 fwd/bwd + numpy Adam)" % (n - 1, B)}


def run_all(pkg, ctx, timer, torch, dev, rank, world, precision, cpu=True):
    F4 = (32, 64, 128, 256)
    out = {}
    common = dict(pkg=pkg, ctx=ctx, timer=timer, torch=torch, dev=dev, rank=rank, world=world, precision=precision)
    out["cfg1"] = gridded_config(**common, name="BASELINE configs[0]: tutorial SST 112x112, ntime_win=3, filters 32*2^(0:3), one chain, batch 32 per GPU",
                                 W=112, H=112, V=1, T=3, F=F4, refine=(1.0,), l2=0.0, batch_per_gpu=32, ncube=1024, steps=20, repeats=3, recon_fields=512)
    out["cfg3_weak"] = gridded_config(**common, name="BASELINE configs[2]: SST + auxiliary field (V=2, 16/18 input channels), refine (0.3,0.7), batch 32 per GPU",
                                      W=112, H=112, V=2, T=3, F=F4, refine=(0.3, 0.7), l2=1e-3, batch_per_gpu=32, ncube=1024, steps=20, repeats=3,
                                      recon_fields=512, isoutput=[True, False])
    if 256 % world == 0:
        out["cfg3_strong"] = gridded_config(**common, name="BASELINE configs[2] at a FIXED global batch of 256 (strong scaling): %d per GPU" % (256 // world),
                                            W=112, H=112, V=2, T=3, F=F4, refine=(0.3, 0.7), l2=1e-3, batch_per_gpu=256 // world, ncube=1024,
                                            steps=10, repeats=3, recon_fields=0, isoutput=[True, False])
        out["cfg3_strong"]["scaling"] = "strong"
    out["cfg4"] = points_config(pkg, ctx, timer, torch, dev, rank, world, precision, cpu)
    b5 = 64 // world if world >= 2 else 32
    out["cfg5"] = gridded_config(**common, name="BASELINE configs[4]: 1024x1024, 5 levels 32*2^(0:4), refine (0.3,0.7), global batch %d" % (b5 * world),
                                 W=1024, H=1024, V=1, T=3, F=(32, 64, 128, 256, 512), refine=(0.3, 0.7), l2=1e-3, batch_per_gpu=b5, ncube=48,
                                 steps=2, repeats=3, recon_fields=10000 if world >= 2 else 1024, recon_seed_per_rank=True,
                                 note=None if world >= 2 else "one GPU: batch 32 (the config's global batch 64 is defined for 2/4/8 GPUs) and a 1024-field "
                                                               "sample of the 10 000-field reconstruction")
    return out
