#!/usr/bin/env python
"""bench.py -- DINCAE hot-path benchmark on B200 (contract: see the round prompt / DESIGN.md section 6).

    python bench.py --gpus N --steps K --warmup W            # this framework (CUDA library through the C-ABI)
    python bench.py --impl reference --gpus N --steps K ...  # CPU oracle port on the box's host cores

Workload (BASELINE.json configs[1]): tutorial SST gap-filling model, 112x112 grid, ntime_win=3 (10 input channels),
enc_nfilter_internal=32*2^(0:3), refinement stage loss_weights_refine=(0.3,0.7), regularization_L2_beta=1e-3,
batch 32 per GPU, synthetic cloud-masked SST cube resident in HBM.  One "step" = one training mini-batch:
batch assembly (K1) -> forward (K2) -> head+loss (K3) -> backward (K2) -> [all-reduce] -> L2+clip+Adam (K3).
metric = training samples/s (whole job).  The reconstruction throughput (Mpixel/s) is reported in "recon".
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = dict(W=112, H=112, V=1, T=3, N=4096, B=32, F=(32, 64, 128, 256), refine=(0.3, 0.7), l2=1e-3, lr=1e-3)


def synth_cube(W, H, N, seed=20261106):
    """Synthetic cloud-masked SST cube: seasonal cycle + smooth noise, land mask, moving clouds."""
    from scipy.ndimage import uniform_filter
    rng = np.random.default_rng(seed)
    j = np.arange(H, dtype=np.float32)[None, :, None]; n = np.arange(N, dtype=np.float32)[None, None, :]

    def smooth(a, r):      # separable box blur over (lon, lat), renormalised to unit variance
        a = uniform_filter(a, size=(2 * r + 1, 2 * r + 1, 1), mode="nearest")
        return a / a.std()
    sst = 15 + 8 * np.cos(2 * np.pi * n / 365.25 - 0.02 * j) + 1.5 * smooth(rng.standard_normal((W, H, N), dtype=np.float32), 4)
    g = smooth(rng.standard_normal((W, H, 1), dtype=np.float32), 8)[:, :, 0]
    land = g <= -0.8
    c = smooth(rng.standard_normal((W, H, N), dtype=np.float32), 6)
    frac = rng.uniform(0.3, 0.9, N)
    thr = np.array([np.quantile(c[:, :, k], 1 - frac[k]) for k in range(N)], dtype=np.float32)
    missing = c > thr[None, None, :]
    data = sst.astype(np.float32)
    data[missing] = np.nan
    data[land, :] = np.nan
    lon = np.linspace(-7, -0.8, W); lat = np.linspace(33.8, 38.2, H)
    days = 11012.0 + np.arange(N)               # daily from 2000-02-25 (examples/DINCAE_tutorial.jl:56)
    return lon, lat, days, data.reshape(W, H, 1, N), ~land


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu=0):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_flops_per_sample(wl):
    """Train FLOPs per sample: fwd + wgrad (= fwd) + dgrad (= fwd minus the first conv of chain 1)."""
    W, H, F = wl["W"], wl["H"], wl["F"]
    cin0 = 4 + 2 * wl["T"] * wl["V"]
    total, first = 0.0, None
    for ch in range(len(wl["refine"])):
        enc = [cin0 + (2 if ch else 0)] + list(F); dec = [2] + list(F)
        h, w = H, W
        for l in range(len(F)):
            fe = 2 * 9 * enc[l] * enc[l + 1] * h * w
            fd = 2 * 9 * dec[l + 1] * dec[l] * h * w
            if first is None:
                first = fe
            total += fe + fd
            h, w = h // 2 + h % 2, w // 2 + w % 2
    return dict(fwd=total, train=3 * total - first)


def run_reference(args, rank, world):
    """--impl reference: the CPU oracle port (PyTorch-CPU ops = oneDNN convs, numpy assembly/Adam) on the host cores.
    The reference's own Atype=Array path is Julia/Flux and cannot run here (no Julia runtime): kind = "port"."""
    if rank != 0:
        return
    import torch
    from oracle import data as OD, model as OM
    wl = dict(WORKLOAD)
    wl["N"] = 64                                     # bounded sample of the same workload: 64 time steps resident on the host
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    lon, lat, days, data, sea = synth_cube(wl["W"], wl["H"], wl["N"])
    od = OD.NCData(lon, lat, days, data, train=True, ntime_win=wl["T"], mask=sea)
    ocfg = OM.ModelConfig(W=wl["W"], H=wl["H"], ninput=10, nfilter=list(wl["F"]), loss_weights_refine=wl["refine"], regularization_L2_beta=wl["l2"])
    flat = OM.init_params(ocfg, 1)
    opt = OM.Adam(flat.size)
    rng = np.random.default_rng(0)
    B = wl["B"]

    def step(k):
        nonlocal flat
        idx = rng.integers(0, wl["N"], B); imask = rng.integers(0, wl["N"], B)
        xin, xt = od.batch(idx, imask, 1, k * B)
        loss, flat = OM.train_step(ocfg, flat, opt, xin, xt, wl["lr"], dtype=torch.float32)
        return loss
    for k in range(args.warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        loss = step(args.warmup + k)
    dt = time.perf_counter() - t0
    val = args.steps * B / dt
    line = {"impl": "reference", "metric": "train_samples_per_s", "value": val, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(dict(WORKLOAD), 1, "hbm"),
            "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": "%d train steps of batch %d (oracle: numpy assembly + torch-CPU fwd/bwd + numpy Adam), cube of %d steps" % (args.steps, B, wl["N"])},
            "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "final_loss": float(loss)}
    emit(line)


def workload_config(wl, world, where):
    return {"workload": "BASELINE configs[1]: tutorial SST 112x112, ntime_win=3, enc_nfilter_internal=32*2^(0:3), refine (0.3,0.7), "
                        "L2 beta=1e-3, batch %d per GPU, %d-step synthetic cube" % (wl["B"], wl["N"]),
            "W": wl["W"], "H": wl["H"], "ntime_win": wl["T"], "batch_per_gpu": wl["B"], "global_batch": wl["B"] * world,
            "nsteps_cube": wl["N"], "parallelism": "dp%d" % world, "resident": where,
            "l2_policy": "inputs larger than L2: %.0f MB cube sampled at random + >1 GB of activations rewritten every step (L2 = 126 MB)"
                         % (wl["W"] * wl["H"] * wl["N"] * 4 / 1e6)}


def cpu_baseline(wl, budget_s=20.0):
    """Oracle port timed on the host cores on a bounded sample (rank 0, N=1 only)."""
    import torch
    from oracle import data as OD, model as OM
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    N = 32
    lon, lat, days, data, sea = synth_cube(wl["W"], wl["H"], N)
    od = OD.NCData(lon, lat, days, data, train=True, ntime_win=wl["T"], mask=sea)
    ocfg = OM.ModelConfig(W=wl["W"], H=wl["H"], ninput=10, nfilter=list(wl["F"]), loss_weights_refine=wl["refine"], regularization_L2_beta=wl["l2"])
    flat = OM.init_params(ocfg, 1)
    opt = OM.Adam(flat.size)
    rng = np.random.default_rng(0)
    B = wl["B"]
    n, t_used = 0, 0.0
    while True:
        idx = rng.integers(0, N, B); imask = rng.integers(0, N, B)
        t0 = time.perf_counter()
        xin, xt = od.batch(idx, imask, 1, n * B)
        _, flat = OM.train_step(ocfg, flat, opt, xin, xt, wl["lr"], dtype=torch.float32)
        dt = time.perf_counter() - t0
        if n > 0:                       # first step = warm-up
            t_used += dt
        n += 1
        if (n >= 3 and t_used > budget_s) or n >= 12:
            break
    return {"value": (n - 1) * B / t_used, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%d train steps of batch %d after 1 warm-up (numpy assembly + torch-CPU float32 fwd/bwd + numpy Adam)" % (n - 1, B)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="dincae_b200", choices=["dincae_b200", "reference"])
    ap.add_argument("--precision", default=os.environ.get("DINCAE_PRECISION", "tf32"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--recon-steps", type=int, default=512)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load_package()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; dincae_b200 has no CPU path (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = dict(WORKLOAD)
    B = wl["B"]
    stream = torch.cuda.Stream()
    ctx = pkg.Context(local, torch_stream=stream)
    lon, lat, days, data, sea = synth_cube(wl["W"], wl["H"], wl["N"])
    ds = pkg.GriddedDataset(ctx, lon, lat, days, data, train=True, ntime_win=wl["T"], mask=sea)
    cfg = pkg.ModelConfig(W=wl["W"], H=wl["H"], ninput=10, enc_nfilter_internal=wl["F"], loss_weights_refine=wl["refine"],
                          regularization_L2_beta=wl["l2"], max_batch=B, precision=args.precision)
    model = pkg.Model(ctx, cfg, seed=1)          # same seed on every rank: replicas start (and stay) identical
    counts = ds.valid_counts(0)
    dev = torch.device("cuda", local)
    npix = wl["W"] * wl["H"]
    xin_d = torch.empty(B * 10 * npix, dtype=torch.float32, device=dev)
    xt_d = torch.empty(B * 2 * npix, dtype=torch.float32, device=dev)
    gview = model.grad_tensor() if world > 1 else None
    rng = np.random.default_rng(1234)            # identical on all ranks: every rank knows the whole global batch

    def draw():
        idx = rng.integers(0, wl["N"], B * world); imask = rng.integers(0, wl["N"], B * world)
        return idx, imask

    step_no = [0]

    def step_device(want_loss=False):
        """device-resident step: assemble this rank's shard on the GPU, train, (all-reduce), Adam."""
        idx, imask = draw()
        lo = rank * B
        ds.assemble(idx[lo:lo + B], imask[lo:lo + B], seed=7, sample_ctr0=step_no[0] * B * world + lo, xin=xin_d, xtrue=xt_d)
        step_no[0] += 1
        if world == 1:
            return model.train_step(xin_d, xt_d, wl["lr"], B=B, want_loss=want_loss)
        nglob = np.array([counts[idx].sum()], np.int64)       # n_noncloud of the GLOBAL batch from static per-step counts: no collective
        model.grad_step(xin_d, xt_d, B=B, n_valid_global=nglob)
        with torch.cuda.stream(stream):
            dist.all_reduce(gview)                             # SUM of normalised gradient sums + loss term, NCCL over NVLink
        return model.apply_step(wl["lr"], want_loss=want_loss)

    def timed(fn, steps):
        with torch.cuda.stream(stream):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = ctx.launches
            e0.record(stream)
            for _ in range(steps):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, ctx.launches - l0

    # ---- device-resident arm -------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms, launches = timed(step_device, args.steps)
    clk = clocks.stop() if rank == 0 else None
    value = args.steps * B * world / (ms / 1e3)
    final_loss = step_device(want_loss=True)

    # ---- end-to-end arm: the reference-facing call with HOST buffers (Atype(inputs_) per batch, src/data.jl:858) ----
    xin_h = torch.empty((B, 10, wl["H"], wl["W"]), dtype=torch.float32).pin_memory()
    xt_h = torch.empty((B, 2, wl["H"], wl["W"]), dtype=torch.float32).pin_memory()
    idx, imask = draw()
    a, b = ds.assemble(idx[rank * B:rank * B + B], imask[rank * B:rank * B + B], seed=7, sample_ctr0=0)
    xin_h.numpy()[...] = np.transpose(a, (3, 2, 1, 0)); xt_h.numpy()[...] = np.transpose(b, (3, 2, 1, 0))
    xin_np = np.transpose(xin_h.numpy(), (3, 2, 1, 0)); xt_np = np.transpose(xt_h.numpy(), (3, 2, 1, 0))   # F-ordered views of the pinned memory
    assert xin_np.flags.f_contiguous and xt_np.flags.f_contiguous

    e2e_slot = [0]
    model.upload(0, xin_np, xt_np)

    def step_e2e():
        if world == 1:
            # one-batch-ahead prefetch like the reference's PrefetchDataIter: the H2D copy of the NEXT batch (19.3 MB from pinned
            # memory) overlaps this step's kernels; every step still pays its own H2D copy and the D2H read of its loss
            s_ = e2e_slot[0]
            model.upload(s_ ^ 1, xin_np, xt_np)
            loss = model.train_step_slot(s_, B, wl["lr"], want_loss=True)
            e2e_slot[0] = s_ ^ 1
            return loss
        nglob = np.array([counts[idx].sum()], np.int64)
        s_ = e2e_slot[0]
        model.upload(s_ ^ 1, xin_np, xt_np)                                        # next shard's H2D overlaps this step, as above
        model.grad_step_slot(s_, B, nglob)
        e2e_slot[0] = s_ ^ 1
        with torch.cuda.stream(stream):
            dist.all_reduce(gview)
        return model.apply_step(wl["lr"], want_loss=True)
    for _ in range(3):
        step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps)
    e2e = {"value": args.steps * B * world / (ms_e2e / 1e3), "unit": "samples/s",
           "h2d_bytes_per_step": int(xin_h.numel() * 4 + xt_h.numel() * 4), "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps}

    # ---- per-kernel CUDA-event profile of the same step (separate pass, so the timed region stays unperturbed) ----
    roof, kern_table = None, None
    lib = pkg.lib.load()
    import ctypes
    nprof = 5
    if rank == 0:
        pkg.lib.check(lib.dincae_profile_begin(ctx.handle))
    for _ in range(nprof):          # every rank runs the steps (they contain the all-reduce); only rank 0 records events
        step_device()
    if rank == 0:
        buf = ctypes.create_string_buffer(1 << 16)
        pkg.lib.check(lib.dincae_profile_end(ctx.handle, buf, len(buf)))
        kern_table = json.loads(buf.value.decode())
        tot = sum(k["ms"] for k in kern_table)
        for k in kern_table:
            k["share"] = k["ms"] / tot
            k["ms_per_launch"] = k["ms"] / k["launches"]
        kern_table.sort(key=lambda k: -k["ms"])
        top = kern_table[0]
        peaks, traffic = {}, {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")))
        except Exception:
            pass
        pk_t = peaks.get("bf16_tflops_sustained", 1400.0)
        pk_b = peaks.get("hbm_gbs", 6650.0)
        src = "MEASURED_PEAKS.json" if peaks else "fallback (6.65 TB/s, 1.4 PFLOP/s sustained)"
        t_s = top["ms"] * 1e-3
        t_flops = top["flops"] / (pk_t * 1e12)           # time the algorithmic FLOPs need at the tensor peak
        t_bytes = top["bytes"] / (pk_b * 1e9)            # time the algorithmic bytes need at the HBM peak
        tr = traffic.get(top["kernel"])
        tr_val = max(tr, key=lambda e: e["dram_bytes"])["dram_bytes"] if tr else None
        common = {"kernel": top["kernel"], "share_of_step": top["share"], "launches_per_step": top["launches"] / nprof,
                  "algorithmic_flops_per_launch": top["flops"] / top["launches"], "algorithmic_bytes_per_launch": top["bytes"] / top["launches"],
                  "avg_launch_us": 1e3 * top["ms"] / top["launches"], "traffic": tr_val,
                  "traffic_note": "dram read+write bytes of the largest launch of this kernel in profiles/r01_final_ncu_full_top_kernels.csv (ncu --set full)" if tr else None,
                  "note": "achieved = algorithmic work of all launches of this kernel in %d steps / their CUDA-event time (library per-launch profile)" % nprof}
        if t_bytes >= t_flops:
            ach = top["bytes"] / t_s / 1e9
            roof = dict(common, bound="hbm", achieved=ach, peak=pk_b, unit="GB/s", frac=ach / pk_b, peak_source=src + " hbm_gbs")
        else:
            ach = top["flops"] / t_s / 1e12
            roof = dict(common, bound="tensor", achieved=ach, peak=pk_t, unit="TFLOP/s", frac=ach / pk_t,
                        peak_source=src + " bf16_tflops_sustained (kernel timed inside a long step; operands are TF32)")

    # ---- reconstruction throughput (time-sharded, no communication) ----
    recon = None
    nrec = min(args.recon_steps, wl["N"])
    per = nrec // world
    t0r, t1r = rank * per, (rank + 1) * per
    m_out = torch.empty(per * npix, dtype=torch.float32, device=dev); s_out = torch.empty_like(m_out)
    test_ds = ds
    def rec():
        pkg.api.reconstruct_range(model, test_ds, t0r, t1r, 1, -9999.0, m_out, s_out)
    rec()
    ms_rec, _ = timed(rec, 2)
    recon = {"value": 2 * per * world * npix / (ms_rec / 1e3) / 1e6, "unit": "Mpixel/s", "steps": per * world, "ms": ms_rec / 2}

    if rank == 0:
        fl = algorithmic_flops_per_sample(wl)
        line = {"metric": "train_samples_per_s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "tf32" if args.precision == "tf32" else "f32", "data": "synthetic",
                "config": workload_config(wl, world, "hbm"), "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clk, "roofline": roof, "recon": recon, "final_loss": float(final_loss),
                "train_gflop_per_sample": fl["train"] / 1e9, "achieved_tflops_whole_step": value * fl["train"] / 1e12,
                "kernels": [{kk: (round(v, 6) if isinstance(v, float) else v) for kk, v in k.items()} for k in (kern_table or [])[:12]]}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def emit(line):
    """The JSON line is the only thing written to the real stdout (libraries such as NCCL print banners on fd 1)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    main()
