#!/usr/bin/env python
"""bench.py -- DINCAE hot-path benchmark on B200 (contract: see the round prompt / DESIGN.md section 6).

    python bench.py --gpus N --steps K --warmup W            # this framework (CUDA library through the C-ABI)
    python bench.py --impl reference --gpus N --steps K ...  # CPU oracle port on the box's host cores

Headline workload (BASELINE.json configs[1]): tutorial SST gap-filling model, 112x112 grid, ntime_win=3 (10 input channels),
enc_nfilter_internal=32*2^(0:3), refinement stage loss_weights_refine=(0.3,0.7), regularization_L2_beta=1e-3,
batch 32 per GPU, synthetic cloud-masked SST cube resident in HBM.  One "step" = one training mini-batch:
batch assembly (K1) -> forward (K2) -> head+loss (K3) -> backward (K2) -> [all-reduce, inside the library] -> L2+clip+Adam (K3),
issued as ONE library call (dincae_train_step_gridded / _dp_gridded) that replays the step as a CUDA graph.
metric = training samples/s (whole job); `value` = median of REPEATS timed blocks of exactly K steps each.
Beside the headline the line carries: `e2e` (host buffers through the C-ABI), `rooflines` (every kernel >= 3 % of the step
against the measured HBM / TF32 tensor roofs), `recon` (Mpixel/s), `configs` (short runs of the other BASELINE configs:
cfg1, cfg3 incl. the fixed-global-256 strong-scaling point, cfg4 points, cfg5 1024^2) and `cpu_baseline`.
"""
import argparse
import ctypes
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
REPEATS = 5


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


def synth_cube_device(W, H, V, N, dev, seed=7, chunk=256):
    """The same kind of cube generated ON the device, shard by shard (the side configs and the 10k-field reconstruction never
    touch the host): torch tensor (N,V,H,W) = the memory of a Julia (W,H,V,N) array; NaN = cloud / land."""
    import torch
    import torch.nn.functional as Fn
    g = torch.Generator(device=dev); g.manual_seed(seed)
    out = torch.empty((N, V, H, W), dtype=torch.float32, device=dev)
    jj = torch.arange(H, device=dev, dtype=torch.float32)[None, None, :, None]

    def smooth(a, r):
        a = Fn.avg_pool2d(a, 2 * r + 1, 1, r, count_include_pad=False)
        return a / a.std()
    land = smooth(torch.randn((1, 1, H, W), device=dev, generator=g), 8)[0, 0] <= -0.8
    for n0 in range(0, N, chunk):
        nn = min(chunk, N - n0)
        t = torch.arange(n0, n0 + nn, device=dev, dtype=torch.float32)[:, None, None, None]
        x = 15 + 8 * torch.cos(2 * math.pi * t / 365.25 - 0.02 * jj) + 1.5 * smooth(torch.randn((nn, V, H, W), device=dev, generator=g), 4)
        c = smooth(torch.randn((nn, V, H, W), device=dev, generator=g), 6)
        frac = 0.3 + 0.6 * torch.rand((nn, V, 1, 1), device=dev, generator=g)
        # c ~ N(0,1) after renormalisation: the (1 - frac) quantile is sqrt(2) * erfinv(1 - 2 frac)
        thr = math.sqrt(2.0) * torch.erfinv(1 - 2 * frac)
        x[c > thr] = float("nan")
        x[:, :, land] = float("nan")
        out[n0:n0 + nn] = x
    lon = np.linspace(-7, -0.8, W); lat = np.linspace(33.8, 38.2, H); days = 11012.0 + np.arange(N)
    return lon, lat, days, out, ~land.cpu().numpy().T          # mask as (W,H)


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed regions."""
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


def train_flops_per_sample(W, H, cin0, nout, F, nchains):
    """Algorithmic train FLOPs per sample: fwd + wgrad (= fwd) + dgrad (= fwd minus the first conv of chain 1)."""
    total, first = 0.0, None
    for ch in range(nchains):
        enc = [cin0 + (2 * nout if ch else 0)] + list(F); dec = [2 * nout] + list(F)
        h, w = H, W
        for l in range(len(F)):
            fe = 2 * 9 * enc[l] * enc[l + 1] * h * w
            fd = 2 * 9 * dec[l + 1] * dec[l] * h * w
            if first is None:
                first = fe
            total += fe + fd
            h, w = h // 2 + h % 2, w // 2 + w % 2
    return dict(fwd=total, train=3 * total - first)


def workload_config(wl, world, where, ncube=None):
    ncube = wl["N"] if ncube is None else ncube
    return {"workload": "BASELINE configs[1]: tutorial SST 112x112, ntime_win=3, enc_nfilter_internal=32*2^(0:3), refine (0.3,0.7), "
                        "L2 beta=1e-3, batch %d per GPU, %d-step synthetic cube" % (wl["B"], ncube),
            "W": wl["W"], "H": wl["H"], "ntime_win": wl["T"], "batch_per_gpu": wl["B"], "global_batch": wl["B"] * world,
            "nsteps_cube": ncube, "parallelism": "dp%d" % world, "resident": where,
            "l2_policy": "inputs larger than L2: %.0f MB cube sampled at random + >1 GB of activations rewritten every step (L2 = 126 MB)"
                         % (wl["W"] * wl["H"] * ncube * 4 / 1e6)}


# ------------------------------------------------------------------------------------------------- CPU arms (oracle port)
def _oracle_gridded_stepper(wl, N):
    import torch
    from oracle import data as OD, model as OM
    lon, lat, days, data, sea = synth_cube(wl["W"], wl["H"], N)
    od = OD.NCData(lon, lat, days, data, train=True, ntime_win=wl["T"], mask=sea)
    ocfg = OM.ModelConfig(W=wl["W"], H=wl["H"], ninput=10, nfilter=list(wl["F"]), loss_weights_refine=wl["refine"], regularization_L2_beta=wl["l2"])
    st = {"flat": OM.init_params(ocfg, 1)}
    opt = OM.Adam(st["flat"].size)
    rng = np.random.default_rng(0)
    B = wl["B"]

    def step(k):
        idx = rng.integers(0, N, B); imask = rng.integers(0, N, B)
        xin, xt = od.batch(idx, imask, 1, k * B)
        loss, st["flat"] = OM.train_step(ocfg, st["flat"], opt, xin, xt, wl["lr"], dtype=torch.float32)
        return loss
    return step


def run_reference(args, rank, world):
    """--impl reference: the CPU oracle port (PyTorch-CPU ops = oneDNN convs, numpy assembly/Adam) on the host cores.
    The reference's own Atype=Array path is Julia/Flux and cannot run here (no Julia runtime): kind = "port"."""
    if rank != 0:
        return
    import torch
    wl = dict(WORKLOAD)
    ncube = 64                                       # bounded sample of the same workload: 64 time steps resident on the host
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step = _oracle_gridded_stepper(wl, ncube)
    B = wl["B"]
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
            "config": workload_config(wl, 1, "host", ncube),
            "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": "%d train steps of batch %d (oracle: numpy assembly + torch-CPU fwd/bwd + numpy Adam), cube of %d steps" % (args.steps, B, ncube)},
            "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "final_loss": float(loss)}
    emit(line)


def cpu_baseline(wl, budget_s=20.0):
    """Oracle port timed on the host cores on a bounded sample (rank 0, N=1 only)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step = _oracle_gridded_stepper(wl, 32)
    n, t_used = 0, 0.0
    while True:
        t0 = time.perf_counter()
        step(n)
        dt = time.perf_counter() - t0
        if n > 0:                       # first step = warm-up
            t_used += dt
        n += 1
        if (n >= 3 and t_used > budget_s) or n >= 12:
            break
    return {"value": (n - 1) * wl["B"] / t_used, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%d train steps of batch %d after 1 warm-up (numpy assembly + torch-CPU float32 fwd/bwd + numpy Adam)" % (n - 1, wl["B"])}


# ------------------------------------------------------------------------------------------------- helpers of the GPU arm
class Timer:
    """Times blocks of exactly `steps` calls with CUDA events on the library's stream, barrier + synchronize on both sides, max over ranks."""

    def __init__(self, torch, dist, stream, dev, world, ctx):
        self.torch, self.dist, self.stream, self.dev, self.world, self.ctx = torch, dist, stream, dev, world, ctx

    def block(self, fn, steps):
        torch, dist = self.torch, self.dist
        with torch.cuda.stream(self.stream):
            if self.world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = self.ctx.launches
            e0.record(self.stream)
            for _ in range(steps):
                fn()
            e1.record(self.stream)
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier()
        ms = e0.elapsed_time(e1)
        if self.world > 1:
            t = torch.tensor([ms], device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, self.ctx.launches - l0

    def median(self, fn, steps, repeats=REPEATS):
        runs = [self.block(fn, steps) for _ in range(repeats)]
        ms = sorted(r[0] for r in runs)
        return ms[len(ms) // 2], runs[0][1], [round(m, 4) for m in ms]


def measure_tf32_roof(torch, dev, seconds=1.0):
    """TF32 tensor roof of this GPU: torch.matmul (cuBLAS, a plain library GEMM used as the yardstick only) on 8192^3 FP32
    operands with TF32 allowed; best single launch (burst) and back-to-back for `seconds` (sustained)."""
    n = 8192
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn((n, n), device=dev); b = torch.randn((n, n), device=dev)
        c = torch.empty((n, n), device=dev)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); torch.matmul(a, b, out=c); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(4, int(seconds * 1e3 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record(); torch.cuda.synchronize()
        sus = e0.elapsed_time(e1) / reps
        fl = 2.0 * n ** 3
        return {"tf32_tflops_burst": fl / best / 1e9, "tf32_tflops_sustained": fl / sus / 1e9,
                "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS): best of 8 / %d back to back" % reps}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
        del a, b, c
        torch.cuda.empty_cache()


def kernel_profile(pkg, lib, ctx, fn, nsteps, rank, every_rank_runs=True):
    """Library per-launch CUDA-event profile of `nsteps` calls of fn (kernels serialised: no side stream, no PDL overlap)."""
    if rank == 0:
        pkg.lib.check(lib.dincae_profile_begin(ctx.handle))
    for _ in range(nsteps):
        fn()
    if rank != 0:
        return None
    buf = ctypes.create_string_buffer(1 << 16)
    pkg.lib.check(lib.dincae_profile_end(ctx.handle, buf, len(buf)))
    tab = json.loads(buf.value.decode())
    tot = sum(k["ms"] for k in tab) or 1.0
    for k in tab:
        k["share"] = k["ms"] / tot
        k["ms_per_launch"] = k["ms"] / k["launches"]
        k["launches_per_step"] = k["launches"] / nsteps
    tab.sort(key=lambda k: -k["ms"])
    return tab


def roofline_of(k, peaks, traffic):
    """Roofline entry of one kernel of the profile table: bound = whichever of its algorithmic bytes / flops takes longer at the
    measured peak; achieved = algorithmic work of all its launches / their CUDA-event time."""
    pk_b, pk_t = peaks["hbm_gbs"], peaks["tf32_tflops"]
    t_s = k["ms"] * 1e-3
    t_flops = k["flops"] / (pk_t * 1e12)
    t_bytes = k["bytes"] / (pk_b * 1e9)
    tr = traffic.get(k["kernel"]) if traffic else None
    e = {"kernel": k["kernel"], "share_of_step": round(k["share"], 4), "launches_per_step": k["launches_per_step"],
         "avg_launch_us": round(1e3 * k["ms_per_launch"], 3),
         "algorithmic_bytes_per_launch": k["bytes"] / k["launches"], "algorithmic_flops_per_launch": k["flops"] / k["launches"],
         # mean DRAM read+write bytes per launch over ALL launches of this kernel in one step of the same command (ncu --set full)
         "traffic": tr["dram_bytes_per_launch"] if tr else None}
    if k["bytes"] == 0 and k["flops"] == 0:
        e.update(bound=None, achieved=None, peak=None, unit=None, frac=None)
    elif t_bytes >= t_flops:
        ach = k["bytes"] / t_s / 1e9
        e.update(bound="hbm", achieved=ach, peak=pk_b, unit="GB/s", frac=ach / pk_b)
    else:
        ach = k["flops"] / t_s / 1e12
        e.update(bound="tensor", achieved=ach, peak=pk_t, unit="TFLOP/s", frac=ach / pk_t)
    return e


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="dincae_b200", choices=["dincae_b200", "reference"])
    ap.add_argument("--precision", default=os.environ.get("DINCAE_PRECISION", "tf32"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the short runs of the other BASELINE configs")
    ap.add_argument("--recon-steps", type=int, default=512)
    ap.add_argument("--repeats", type=int, default=REPEATS)
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
    dev = torch.device("cuda", local)
    stream = torch.cuda.Stream()
    ctx = pkg.Context(local, torch_stream=stream)
    lib = pkg.lib.load()
    comm_info = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)          # plumbing only: barriers, max-over-ranks of the timings, the id broadcast
        box = [pkg.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ctx.comm_init(world, rank, box[0])                      # the gradient all-reduce itself lives inside the library
        comm_info = ctx.comm_info()
    timer = Timer(torch, dist, stream, dev, world, ctx)
    wl = dict(WORKLOAD)
    B = wl["B"]
    lon, lat, days, data, sea = synth_cube(wl["W"], wl["H"], wl["N"])
    ds = pkg.GriddedDataset(ctx, lon, lat, days, data, train=True, ntime_win=wl["T"], mask=sea)
    cfg = pkg.ModelConfig(W=wl["W"], H=wl["H"], ninput=10, enc_nfilter_internal=wl["F"], loss_weights_refine=wl["refine"],
                          regularization_L2_beta=wl["l2"], max_batch=B, precision=args.precision)
    model = pkg.Model(ctx, cfg, seed=1)          # same seed on every rank: replicas start (and stay) identical
    counts = ds.valid_counts(0)
    npix = wl["W"] * wl["H"]
    xin_d = torch.empty(B * 10 * npix, dtype=torch.float32, device=dev)
    xt_d = torch.empty(B * 2 * npix, dtype=torch.float32, device=dev)
    rng = np.random.default_rng(1234)            # identical on all ranks: every rank knows the whole global batch

    def draw():
        return rng.integers(0, wl["N"], B * world), rng.integers(0, wl["N"], B * world)

    step_no = [0]

    def step_device(want_loss=False):
        """device-resident step: assemble this rank's shard on the GPU, train, (all-reduce overlapped with backward), Adam."""
        idx, imask = draw()
        lo = rank * B
        ctr0 = step_no[0] * B * world + lo
        step_no[0] += 1
        # assembly (getxy! of this rank's samples) + training step in one library call = one CUDA graph per step
        if world == 1:
            return model.train_step_gridded(ds, idx[lo:lo + B], imask[lo:lo + B], wl["lr"], seed=7, sample_ctr0=ctr0, want_loss=want_loss)
        nglob = np.array([counts[idx].sum()], np.int64)       # n_noncloud of the GLOBAL batch from static per-step counts: no collective
        return model.train_step_gridded(ds, idx[lo:lo + B], imask[lo:lo + B], wl["lr"], seed=7, sample_ctr0=ctr0, n_valid_global=nglob,
                                        want_loss=want_loss)

    # ---- device-resident arm -------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms, launches, ms_all = timer.median(step_device, args.steps, args.repeats)
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
    nglob_e2e = np.array([counts[idx].sum()], np.int64)

    def step_e2e():
        # one-batch-ahead prefetch like the reference's PrefetchDataIter: the H2D copy of the NEXT batch (19.3 MB from pinned
        # memory) overlaps this step's kernels; every step still pays its own H2D copy and the D2H read of its loss
        s_ = e2e_slot[0]
        model.upload(s_ ^ 1, xin_np, xt_np)
        e2e_slot[0] = s_ ^ 1
        if world == 1:
            return model.train_step_slot(s_, B, wl["lr"], want_loss=True)
        return model.train_step_dp_slot(s_, B, wl["lr"], nglob_e2e, want_loss=True)
    for _ in range(3):
        step_e2e()
    ms_e2e, _, ms_e2e_all = timer.median(step_e2e, args.steps, args.repeats)
    clk = clocks.stop() if rank == 0 else None
    h2d = int(xin_h.numel() * 4 + xt_h.numel() * 4)
    e2e = {"value": args.steps * B * world / (ms_e2e / 1e3), "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
           "ms_per_step": ms_e2e / args.steps, "h2d_gbs_per_rank": h2d / (ms_e2e / args.steps * 1e-3) / 1e9, "ms_blocks": ms_e2e_all}
    divergence = model.replica_divergence() if world > 1 else None

    # ---- measured roofs ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    roofs = {"hbm_gbs": peaks.get("hbm_gbs", 6650.0), "hbm_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"}

    # ---- per-kernel CUDA-event profile of the same step (separate pass, so the timed region stays unperturbed) ----
    nprof = 5
    kern_table = kernel_profile(pkg, lib, ctx, step_device, nprof, rank)

    # ---- reconstruction throughput (time-sharded, no communication) ----
    nrec = min(args.recon_steps, wl["N"])
    per = nrec // world
    t0r, t1r = rank * per, (rank + 1) * per
    m_out = torch.empty(per * npix, dtype=torch.float32, device=dev); s_out = torch.empty_like(m_out)

    def rec():
        pkg.api.reconstruct_range(model, ds, t0r, t1r, 1, -9999.0, m_out, s_out)
    rec()
    ms_rec, _, _ = timer.median(rec, 2, 3)
    fl = train_flops_per_sample(wl["W"], wl["H"], 10, 1, wl["F"], 2)
    recon = {"value": 2 * per * world * npix / (ms_rec / 1e3) / 1e6, "unit": "Mpixel/s", "steps": per * world, "ms": ms_rec / 2}
    if rank == 0:
        recon["achieved_tflops"] = recon["value"] * 1e6 / npix * fl["fwd"] / 1e12
        rtab = kernel_profile(pkg, lib, ctx, rec, 1, 0)
        recon["kernels"] = [{"kernel": k["kernel"], "share": round(k["share"], 4), "launches": k["launches"], "ms": round(k["ms"], 4)} for k in rtab[:8]]
    elif world > 1:
        rec()           # keep the ranks in step with rank 0's profile pass
    model.close(); ds.close()
    del xin_d, xt_d, m_out, s_out
    torch.cuda.empty_cache()

    # ---- the other BASELINE configs (short runs) ----
    configs = None
    if not args.no_configs:
        import bench_configs
        configs = bench_configs.run_all(pkg, ctx, timer, torch, dev, rank, world, args.precision,
                                        cpu=(world == 1 and not args.no_cpu_baseline))

    # ---- the TF32 tensor roof, measured LAST: a second of back-to-back 8192^3 GEMMs heats the GPU into its power cap, and
    #      whatever is timed right after it runs at reduced clocks
    roof, rooflines = None, None
    if rank == 0:
        tf32 = measure_tf32_roof(torch, dev)
        roofs.update(tf32)
        roofs["tf32_tflops"] = tf32["tf32_tflops_sustained"]        # kernels timed inside a long step: the sustained figure
        roofs["tensor_source"] = "measured in this run: " + tf32["how"]
        traffic = {}
        try:
            import glob
            traffic = json.load(open(sorted(glob.glob(os.path.join(ROOT, "profiles", "r02*_ncu_traffic.json")))[-1]))      # latest capture of this round
        except Exception:
            pass
        rooflines = [roofline_of(k, roofs, traffic) for k in kern_table if k["share"] >= 0.03]
        roof = dict(rooflines[0], peak_source=roofs["hbm_source"] if rooflines[0]["bound"] == "hbm" else roofs["tensor_source"],
                    note="dominant kernel of the per-launch CUDA-event profile (%d steps, kernels serialised); traffic = mean DRAM bytes per launch "
                         "of the same kernel over one whole training step in profiles/r02*_ncu_traffic.json (ncu --set full of the same command)" % nprof)
        recon["frac_of_tf32_roof"] = recon["achieved_tflops"] / world / roofs["tf32_tflops"]
        if configs:
            bench_configs.finalize(configs, roofs, world)

    if rank == 0:
        line = {"metric": "train_samples_per_s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {"tf32": "tf32", "tf32x1": "tf32", "fp32": "f32"}[args.precision], "precision_mode": args.precision, "data": "synthetic",
                "config": workload_config(wl, world, "hbm"), "e2e": e2e, "gpu_launches": int(launches),
                "repeats": args.repeats, "ms_blocks": ms_all, "timing": "median of %d blocks of exactly %d steps (CUDA events, max over ranks)" % (args.repeats, args.steps),
                "clocks": clk, "roofline": roof, "rooflines": rooflines, "roofs": roofs, "recon": recon, "final_loss": float(final_loss),
                "train_gflop_per_sample": fl["train"] / 1e9, "achieved_tflops_whole_step": value * fl["train"] / 1e12,
                "frac_of_tf32_roof_whole_step": value * fl["train"] / 1e12 / world / roofs["tf32_tflops"],
                "kernels": [{kk: (round(v, 6) if isinstance(v, float) else v) for kk, v in k.items()} for k in (kern_table or [])[:14]],
                "configs": configs}
        if world > 1:
            line["comm"] = dict(comm_info, where="inside libdincae_b200 (dincae_train_step_dp: one all-reduce of the 6.2 MB gradient buffer after the backward pass; "
                                      "from 4 ranks on the library's own two-shot kernel over NVLink peer memory, else ncclAllReduce)", replica_max_abs_diff=divergence)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def emit(line):
    """The JSON line is the only thing written to the real stdout (libraries such as NCCL print banners on fd 1)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    main()
