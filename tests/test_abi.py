"""CPU-side checks of the drop-in boundary: the library loads and exports every symbol that
include/dincae_b200.h declares (no compute without a GPU), and the host mirror validates its inputs."""
import ctypes
import numpy as np
import pytest


def test_library_exports_every_header_symbol(pkg):
    lib = pkg.lib.load()
    names = pkg.lib.header_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), n
    # every declared function also has a ctypes signature in the host binding
    assert set(names) <= set(pkg.lib.SIGNATURES), set(names) - set(pkg.lib.SIGNATURES)
    assert lib.dincae_version() >= 100


def test_config_struct_matches_header(pkg):
    # sizeof(dincae_config) from the C declaration: 5 + 8 ints, 1 uint, 2 ints, 4 floats, ... = 35 x 4 bytes
    assert ctypes.sizeof(pkg.lib.Config) == 4 * (5 + 8 + 1 + 1 + 1 + 4 + 1 + 1 + 2 + 2 + 1 + 3 + 1 + 1 + 1)


def test_no_gpu_fails_loudly(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.lib.DincaeError) as e:
        pkg.Context(0)
    assert "no CPU path" in str(e.value) or "CUDA" in str(e.value)


def test_model_config_to_struct(pkg):
    cfg = pkg.ModelConfig(W=112, H=112, ninput=10, enc_nfilter_internal=(32, 64, 128, 256), loss_weights_refine=(0.3, 0.7),
                          regularization_L2_beta=1e-3)
    c = cfg.to_struct()
    assert c.nlevels == 4 and list(c.nfilter)[:4] == [32, 64, 128, 256]
    assert c.skip_mask == 0b1110          # default skipconnections = 2:(L+1); level L+1 is the identity
    assert c.nchains == 2 and abs(c.gamma - 10.0) < 1e-5     # gamma = log(min_std_err^-2) = 10 for exp(-5)
    with pytest.raises(ValueError):
        pkg.ModelConfig(W=8, H=8, ninput=4, upsampling_method="cubic").to_struct()


def test_reconstruct_rejects_empty_save_epochs(pkg):
    # test/test_DINCAE_SST.jl:190-195, src/model.jl:475-477
    with pytest.raises(ValueError, match="No output will be saved"):
        pkg.reconstruct(None, [[]], epochs=3, save_epochs=[])
    with pytest.raises(ValueError, match="No output will be saved"):
        pkg.reconstruct_points(None, {}, None, epochs=3, save_epochs=[])   # test/test_reconstruct_points.jl:106-110


def test_databatches_sharding_and_order(pkg):
    class FakeDS:
        train = True
        def __len__(self):
            return 10
    d = pkg.api.DataBatches(FakeDS(), 4, seed=1, rank=1, world_size=2)
    seen = []
    for idx, imask, ctr in d.batches():
        assert len(imask) == len(idx) and imask.max() < 10
        seen += list(idx)
        lo_idx, lo_mask, off = d.shard(idx, imask)
        assert list(lo_idx) == list(idx[off:off + len(lo_idx)])
    assert sorted(seen) == list(range(10)) and len(d) == 3      # partial last batch kept (src/data.jl:803)
