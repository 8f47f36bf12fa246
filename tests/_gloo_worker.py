"""Worker of tests/test_dp_gloo.py: one data-parallel rank on the CPU (gloo), oracle as the compute engine.
Checks the host-side logic of the N>1 path: identical draws on every rank, contiguous shards, global n_noncloud
from static per-step counts (no collective), one SUM all-reduce of [gradient sums, loss term]."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
from oracle import data as OD, model as OM


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    pkg = ge.load_package()
    rng = np.random.default_rng(0)                       # same stream on every rank
    W, H, N, T, B = 12, 10, 9, 3, 5                      # B = 5 over 2 ranks: uneven shards (3 + 2)
    data = (rng.standard_normal((W, H, 1, N)) + 10).astype(np.float32)
    data[rng.random((W, H, 1, N)) < 0.4] = np.nan
    od = OD.NCData(np.linspace(0, 1, W), np.linspace(0, 1, H), 11000.0 + np.arange(N), data, train=True, ntime_win=T)
    counts = (~np.isnan(data[:, :, 0, :])).sum(axis=(0, 1))          # what dincae_gridded_valid_counts returns

    class DS:                                            # the only thing DataBatches needs from a dataset here
        train = True
        def __len__(self):
            return N
    batches = pkg.api.DataBatches(DS(), B, seed=3, rank=rank, world_size=world)
    ocfg = OM.ModelConfig(W=W, H=H, ninput=10, nfilter=[4, 6], loss_weights_refine=(0.3, 0.7))
    flat = OM.init_params(ocfg, 2)
    ok = True
    for idx, imask, ctr in batches.batches():
        sidx, smask, lo = batches.shard(idx, imask)
        n_global = counts[idx].sum()
        n_local = counts[sidx].sum() if len(sidx) else 0
        if len(sidx):
            xin, xt = od.batch(sidx, smask, 3, ctr + lo)
            loss, g = OM.loss_and_grad(ocfg, flat, xin, xt)
            scale = n_local / n_global                  # local cost is normalised by n_local; the library uses n_global directly
            buf = torch.tensor(np.concatenate([g * scale, [loss * scale]]))
        else:
            buf = torch.zeros(flat.size + 1, dtype=torch.float64)
        dist.all_reduce(buf)                             # SUM
        if rank == 0:
            xin, xt = od.batch(idx, imask, 3, ctr)
            loss_full, g_full = OM.loss_and_grad(ocfg, flat, xin, xt)
            ok &= np.allclose(buf[:-1].numpy(), g_full, rtol=1e-9, atol=1e-12)
            ok &= abs(buf[-1].item() - loss_full) < 1e-9 * abs(loss_full)
    flag = torch.tensor([1.0 if ok else 0.0])
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
