#!/usr/bin/env python
"""What a kernel really costs inside the overlapped step (PDL + side stream), where per-kernel event timings over-count:
run the headline bench with one kernel family dropped at a time and print the step-time difference.

Needs a measurement build:  DINCAE_NVCC_EXTRA=-DDINCAE_ABLATE python -c "import __graft_entry__ as g; g.build_library(force=True)"
(rebuild without the flag afterwards: results of an ablated run are garbage by construction).
    python scripts/ablate.py [out.json]"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GROUPS = ["", "conv3x3_persist_bn64", "conv3x3_persist_bn32", "conv3x3_pair_bn64", "conv3x3_pair_bn128", "wgradm", "wgrad9_n64", "wgrad9_n32",
          "poolbwd_kernel", "upbwd_kernel", "up_nearest_kernel", "head_bwd_data", "head_up_conv", "head_wgrad", "reduce_batched_kernel",
          "pack_planar_kernel", "loss_grid_kernel", "adam_kernel", "assemble_kernel",
          "poolbwd_kernel,upbwd_kernel,up_nearest_kernel", "wgradm,wgrad9_n64,wgrad9_n32,head_wgrad,reduce_batched_kernel"]
res = {}
for g in GROUPS:
    env = dict(os.environ, DINCAE_ABLATE=g)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-configs", "--no-cpu-baseline", "--recon-steps", "8", "--steps", "60",
                        "--warmup", "5", "--repeats", "3"], env=env, capture_output=True, text=True, cwd=ROOT)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        res[g or "(nothing dropped)"] = d["ms_per_step"] * 1e3
    except Exception as e:                     # noqa: BLE001
        res[g] = None
        print("failed:", g, r.stderr[-300:], file=sys.stderr)
base = res["(nothing dropped)"]
for g, v in res.items():
    print("%-70s %8.1f us/step   saves %7.1f us" % (g, v or float("nan"), (base - v) if v else float("nan")))
if len(sys.argv) > 1:
    json.dump({"us_per_step_with_kernel_dropped": res}, open(sys.argv[1], "w"), indent=1)
