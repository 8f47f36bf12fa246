"""Worker of tests/test_gpu_nccl.py: one data-parallel rank on its own GPU.  The collective is the library's own
(dincae_comm_* / dincae_train_step_dp, NCCL bound inside libdincae_b200); the rendezvous id travels through a file, so no
torch.distributed is involved.  usage: _nccl_worker.py <rank> <world> <dir>"""
import json
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge


def main():
    rank, world, d = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    pkg = ge.load_package()
    ctx = pkg.Context(rank)
    idf = os.path.join(d, "nccl_id.bin")
    if rank == 0:
        uid = pkg.Context.comm_unique_id()
        with open(idf + ".tmp", "wb") as f:
            f.write(uid)
        os.rename(idf + ".tmp", idf)
    else:
        t0 = time.time()
        while not os.path.exists(idf):
            if time.time() - t0 > 120:
                raise SystemExit("rank %d: no rendezvous id" % rank)
            time.sleep(0.05)
        uid = open(idf, "rb").read()
    ctx.comm_init(world, rank, uid)
    info = ctx.comm_info()
    assert info["nranks"] == world and info["rank"] == rank and info["nccl_version"] > 20000, info

    from test_gpu_parity import _problem
    out = {"rank": rank, "info": info, "cases": []}
    for prec, F in (("fp32", (6, 8)), ("tf32", (32, 64))):
        rng = np.random.default_rng(8)                                  # same stream on every rank
        W, H, Cin, Bg = 24, 20, 10, 3 * world + 1                       # uneven shards
        cfg = pkg.ModelConfig(W=W, H=H, ninput=Cin, enc_nfilter_internal=F, loss_weights_refine=(0.3, 0.7),
                              regularization_L2_beta=1e-3, max_batch=Bg, precision=prec)
        model = pkg.Model(ctx, cfg, seed=2)                             # replicas start identical
        ref = pkg.Model(ctx, cfg, seed=2) if rank == 0 else None        # the whole mini-batch on one GPU
        bounds = np.linspace(0, Bg, world + 1).astype(int)               # uneven shards (3 + 4 for two ranks)
        lo, hi = bounds[rank], bounds[rank + 1]
        losses, losses_ref = [], []
        for step in range(3):
            xin, xt = _problem(rng, W, H, Cin, 1, Bg)
            nv = (xt[:, :, 1::2] != 0).sum(axis=(0, 1, 3)).astype(np.int64)        # n_noncloud of the GLOBAL mini-batch
            losses.append(model.train_step_dp(np.asfortranarray(xin[..., lo:hi]), np.asfortranarray(xt[..., lo:hi]), 1e-3, nv))
            if ref is not None:
                losses_ref.append(ref.train_step(xin, xt, 1e-3))
        div = model.replica_divergence()
        case = {"precision": prec, "losses": losses, "divergence": div}
        if ref is not None:
            case["losses_ref"] = losses_ref
            case["max_param_diff"] = float(np.abs(model.get_params() - ref.get_params()).max())
        out["cases"].append(case)
    # plain collective on a caller buffer
    import torch
    t = torch.full((1000,), float(rank + 1), device="cuda:%d" % rank)
    ctx.allreduce(t, t.numel()); ctx.sync()
    out["allreduce"] = float(t[0].item())
    with open(os.path.join(d, "rank%d.json" % rank), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()
