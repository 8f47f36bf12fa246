"""2-D vector cost (src/vector2.jl, output_ndims = 2).  CPU part: the oracle against an independent scalar loop; GPU part: the
library's dincae_vector2_loss against the oracle, in the shape of the reference's own test (test/test_vector2.jl:10-42: sz = (100,120),
batch 3, nsites = 3, directionobs 360 * rand with a batch axis, cost isa Number, device == CPU to rtol 1e-4)."""
import math
import numpy as np
import pytest

from oracle import vector2 as ov


def _case(rng, W, H, B, nsites, dir_batch=True, holes=0.0, conditioned=False):
    xrec = rng.standard_normal((W, H, 5, B)).astype(np.float32)
    if conditioned:        # diagonal of L away from zero: p' P p >= 0.2, so neither the 1e-4 clamp nor Float32 cancellation decides a gradient
        xrec[:, :, 2, :] = rng.uniform(0.5, 1.5, (W, H, B)); xrec[:, :, 4, :] = rng.uniform(0.5, 1.5, (W, H, B))
    xtrue = rng.standard_normal((W, H, 2 * nsites, B)).astype(np.float32)
    if holes > 0:
        inv = xtrue[:, :, 1::2, :]
        inv[rng.uniform(size=inv.shape) < holes] = 0.0
        xtrue[:, :, 0::2, :][inv == 0] = 0.0
    shape = (W, H, nsites, B) if dir_batch else (W, H, nsites)
    direction = (360 * rng.uniform(size=shape)).astype(np.float32)
    return xrec, xtrue, direction


def _scalar_loop(xrec, xtrue, truth_uncertain, direction):
    """the same cost written pixel by pixel in double precision, straight from the formulas in the docstring of src/vector2.jl:1-12"""
    W, H, _, B = xrec.shape
    nsites = direction.shape[2]
    tot, n = 0.0, 0
    for b in range(B):
        for j in range(H):
            for i in range(W):
                x1, x2, l11, l21, l22 = (float(v) for v in xrec[i, j, :, b])
                L = np.array([[l11, 0.0], [l21, l22]])
                P = L @ L.T
                uv = P @ np.array([x1, x2])
                for s in range(nsites):
                    inv = float(xtrue[i, j, 2 * s + 1, b])
                    if inv == 0.0:
                        continue
                    d = float(direction[i, j, s, b] if direction.ndim == 4 else direction[i, j, s])
                    t = d * float(np.float32(np.pi)) / 180.0
                    pv = np.array([math.sin(t), math.cos(t)])
                    m = pv @ uv
                    s2 = max(pv @ P @ pv, 1e-4)
                    s2t = 1.0 / max(inv, 1e-3)
                    mt = float(xtrue[i, j, 2 * s, b]) * s2t
                    if truth_uncertain:
                        tot += math.log(s2 / s2t) + ((m - mt) ** 2 + s2t) / s2
                    else:
                        tot += math.log(s2) + (m - mt) ** 2 / s2
                    n += 1
    return tot / n


@pytest.mark.parametrize("truth_uncertain", [False, True])
@pytest.mark.parametrize("dir_batch", [True, False])
def test_oracle_matches_scalar_loop(truth_uncertain, dir_batch):
    rng = np.random.default_rng(5)
    xrec, xtrue, direction = _case(rng, 7, 6, 2, 3, dir_batch, holes=0.3)
    ref = _scalar_loop(xrec, xtrue, truth_uncertain, direction)
    got = ov.vector2_costfun(xrec, xtrue, truth_uncertain, direction)
    assert isinstance(got, float)
    assert abs(got - ref) <= 2e-5 * abs(ref)
    g64, _ = ov.vector2_costfun_grad(xrec, xtrue, truth_uncertain, direction)
    assert abs(g64 - ref) <= 1e-9 * abs(ref)


def test_oracle_gradient_is_the_derivative():
    rng = np.random.default_rng(6)
    xrec, xtrue, direction = _case(rng, 5, 4, 2, 2, True, holes=0.2)
    c0, g = ov.vector2_costfun_grad(xrec, xtrue, False, direction)
    x64 = xrec.astype(np.float64)
    for idx in [(0, 0, 0, 0), (2, 1, 2, 1), (4, 3, 3, 0), (1, 2, 4, 1), (3, 0, 1, 1)]:
        h = 1e-6
        xp = x64.copy(); xp[idx] += h
        xm = x64.copy(); xm[idx] -= h
        cp, _ = ov.vector2_costfun_grad(xp, xtrue, False, direction)
        cm, _ = ov.vector2_costfun_grad(xm, xtrue, False, direction)
        assert abs((cp - cm) / (2 * h) - g[idx]) <= 1e-4 * max(1.0, abs(g[idx])) + 1e-15 * abs(c0) / h     # the cost is ~1e7: cancellation


@pytest.mark.gpu
@pytest.mark.parametrize("truth_uncertain", [False, True])
@pytest.mark.parametrize("W,H,B,nsites,dir_batch,holes", [(100, 120, 3, 3, True, 0.0),      # test/test_vector2.jl:10-22
                                                          (100, 120, 3, 3, False, 0.4),
                                                          (177, 69, 5, 1, False, 0.9),
                                                          (33, 17, 1, 4, True, 0.5)])
def test_device_cost_and_gradient_match_the_oracle(pkg, truth_uncertain, W, H, B, nsites, dir_batch, holes):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ctx = pkg.Context(0)
    rng = np.random.default_rng(100 + W + nsites)
    xrec, xtrue, direction = _case(rng, W, H, B, nsites, dir_batch, holes)
    ref = ov.vector2_costfun(xrec, xtrue, truth_uncertain, direction)
    cost = pkg.api.vector2_costfun(xrec, xtrue, truth_uncertain, direction, ctx)
    assert isinstance(cost, float) and math.isfinite(cost)               # "cost isa Number"
    assert abs(cost - ref) <= 1e-4 * abs(ref)                            # the reference's own device-vs-CPU tolerance
    cost2, grad = pkg.api.vector2_costfun(xrec, xtrue, truth_uncertain, direction, ctx, with_grad=True)
    assert cost2 == cost and np.all(np.isfinite(grad))
    # gradient on a well-conditioned field (random L puts p' P p on the 1e-4 clamp, where Float32 and Float64 take different branches)
    xrec, xtrue, direction = _case(rng, W, H, B, nsites, dir_batch, holes, conditioned=True)
    c64, g64 = ov.vector2_costfun_grad(xrec, xtrue, truth_uncertain, direction)
    cost3, grad = pkg.api.vector2_costfun(xrec, xtrue, truth_uncertain, direction, ctx, with_grad=True)
    assert abs(cost3 - c64) <= 1e-5 * abs(c64)
    # Float32 device gradient against the Float64 derivative: entry by entry 1e-3 relative with a floor of 1e-5 of the largest entry
    scale = np.abs(g64).max()
    assert np.all(np.abs(grad - g64) <= 1e-3 * np.abs(g64) + 1e-5 * scale)


@pytest.mark.gpu
def test_device_pointers_and_nan_is_an_error(pkg):
    import ctypes as C
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ctx = pkg.Context(0)
    rng = np.random.default_rng(9)
    W, H, B, nsites = 24, 20, 2, 2
    xrec, xtrue, direction = _case(rng, W, H, B, nsites, False, 0.3)
    ref, gref = pkg.api.vector2_costfun(xrec, xtrue, False, direction, ctx, with_grad=True)
    dev = torch.device("cuda", 0)
    f = lambda a: torch.from_numpy(np.asfortranarray(a).ravel(order="F").copy()).to(dev)
    d_rec, d_true, d_dir = f(xrec), f(xtrue), f(direction)
    d_grad = torch.empty_like(d_rec)
    cost = C.c_double()
    L = pkg.lib
    L.check(L.load().dincae_vector2_loss(ctx.handle, L.ptr(d_rec), L.ptr(d_true), L.ptr(d_dir), W, H, nsites, B, 0, 0, L.DEVICE,
                                         C.byref(cost), L.ptr(d_grad)))
    assert cost.value == ref
    assert np.array_equal(d_grad.cpu().numpy(), gref.ravel(order="F"))
    bad = xrec.copy(); bad[3, 4, 2, 1] = np.nan
    with pytest.raises(L.DincaeError) as e:
        pkg.api.vector2_costfun(bad, xtrue, False, direction, ctx)
    assert "NaN" in str(e.value)
