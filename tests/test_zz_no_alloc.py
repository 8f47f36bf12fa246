"""dincae_interp_adjoint_noalloc: the adjoint scatter-add for hosts that cannot allocate GPU memory
(device memory owned by another pool, e.g. CUDA.jl's allocator around a ccall).  It must give exactly
what dincae_interp_adjoint gives, on repeated calls that reuse the library's built-in scratch array.
Collected last on purpose (file name), so a failure here cannot hide the other GPU tests under -x."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(rng, W, H, P):
    pos = np.stack([rng.uniform(1, W, P), rng.uniform(1, H, P)], 1).astype(np.float32)
    pos = np.minimum(pos, np.nextafter(np.array([W, H], np.float32), 0)).astype(np.float32)
    pos[:P // 8] = pos[0]                       # many points in one cell: long segments
    return pos, rng.standard_normal(P).astype(np.float32)


def test_noalloc_matches_allocating_entry_point(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ctx = pkg.Context(0)
    rng = np.random.default_rng(42)
    dev = torch.device("cuda", 0)
    # two calls in a row, the second smaller than the first so stale scratch content would show
    for W, H, P in ((177, 69, 40000), (64, 48, 1234), (177, 69, 65536)):
        pos, val = _case(rng, W, H, P)
        ref = pkg.interp_adjn(pos, val, (W, H), ctx)                      # existing entry point (allocates)
        d_pos = torch.from_numpy(pos).to(dev); d_val = torch.from_numpy(val).to(dev)
        d_out = torch.full((H, W), 7.0, dtype=torch.float32, device=dev)  # garbage: the call must clear it
        torch.cuda.synchronize()
        for rep in range(2):
            pkg.api.interp_adjn_noalloc(d_pos, d_val, P, (W, H), d_out, ctx)
            ctx.sync()
            out = d_out.cpu().numpy().T                                   # (H,W) C-order == (W,H) column-major
            assert np.array_equal(out, ref), "no-alloc result differs (W=%d H=%d P=%d rep=%d)" % (W, H, P, rep)
    # empty input clears the grid
    d_out = torch.full((8, 9), 3.0, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    pkg.api.interp_adjn_noalloc(None, None, 0, (9, 8), d_out, ctx)
    ctx.sync()
    assert not d_out.cpu().numpy().any()


def test_noalloc_rejects_oversize(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ctx = pkg.Context(0)
    dev = torch.device("cuda", 0)
    P = 65537
    d_pos = torch.ones((P, 2), dtype=torch.float32, device=dev); d_val = torch.ones(P, dtype=torch.float32, device=dev)
    d_out = torch.zeros((8, 8), dtype=torch.float32, device=dev)
    with pytest.raises(pkg.lib.DincaeError) as e:
        pkg.api.interp_adjn_noalloc(d_pos, d_val, P, (8, 8), d_out, ctx)
    assert e.value.code == pkg.lib.ESIZE and "65536" in str(e.value)


def test_noalloc_leaves_device_memory_untouched(pkg):
    """The property the entry point exists for: after the one-time address lookup, calls take no device memory at all.
    Two FRESH contexts (they have never run the allocating path, so no context scratch exists that a fallback could borrow)
    call it while the free-memory reading of the device is watched; the second context also shows the cached address is reused."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(7)
    W, H, P = 120, 90, 30000
    pos, val = _case(rng, W, H, P)
    ref = pkg.interp_adjn(pos, val, (W, H), pkg.Context(0))               # reference result from a separate context
    d_pos = torch.from_numpy(pos).to(dev); d_val = torch.from_numpy(val).to(dev)
    d_out = torch.empty((H, W), dtype=torch.float32, device=dev)
    a, b = pkg.Context(0), pkg.Context(0)
    pkg.api.interp_adjn_noalloc(d_pos, d_val, P, (W, H), d_out, a)        # warm-up: symbol address lookup / module load
    a.sync()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for ctx in (a, b, a, b):
        d_out.fill_(5.0)
        torch.cuda.synchronize()
        pkg.api.interp_adjn_noalloc(d_pos, d_val, P, (W, H), d_out, ctx)
        ctx.sync()
        assert np.array_equal(d_out.cpu().numpy().T, ref)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0                          # nothing was allocated (or freed) on the device
