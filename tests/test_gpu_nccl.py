"""Data-parallel training through the library's own collective on >= 2 GPUs (skipped on a one-GPU box): shards + the gradient
all-reduce (NCCL, and the peer-memory kernel) reproduce single-GPU training on the whole mini-batch, and the replicas stay
bit-identical."""
import json
import os
import subprocess
import sys
import tempfile
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("p2p", ["0", "1"])          # ncclAllReduce / the library's all-reduce kernel over NVLink peer memory
def test_train_step_dp_equals_full_batch_and_replicas_identical(p2p):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    world = 2
    with tempfile.TemporaryDirectory() as d:
        procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "_nccl_worker.py"), str(r), str(world), d],
                                  stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=dict(os.environ, DINCAE_P2P=p2p))
                 for r in range(world)]
        outs = [p.communicate(timeout=600)[0] for p in procs]
        assert all(p.returncode == 0 for p in procs), "\n".join(o[-3000:] for o in outs)
        res = [json.load(open(os.path.join(d, "rank%d.json" % r))) for r in range(world)]
    assert res[0]["allreduce"] == res[1]["allreduce"] == 3.0
    for c0, c1 in zip(res[0]["cases"], res[1]["cases"]):
        assert c0["losses"] == c1["losses"]                         # every rank reports the global loss
        assert c0["divergence"] == 0.0 and c1["divergence"] == 0.0   # replicas bit-identical after 3 steps
        tol = 1e-5 if c0["precision"] == "fp32" else 1e-4
        assert c0["losses"] == pytest.approx(c0["losses_ref"], rel=tol)
        assert c0["max_param_diff"] < (5e-6 if c0["precision"] == "fp32" else 2e-3)    # Adam's first steps move every weight by ~lr: sign noise on tiny gradients
