"""Oracle for the 2-D vector cost (src/vector2.jl).  TEST INFRASTRUCTURE ONLY: nothing in the product imports this.

Parity: unpinned.  The reference's own test (test/test_vector2.jl:10-42) holds no values -- it asserts that the cost is a Number and
that a device array gives the CPU result to rtol 1e-4 -- and there is no Julia runtime here to generate any.  The restatement below
follows the source operation by operation in Float32; tests/test_vector2.py checks it against an independent scalar loop.

Arrays are indexed like Julia: xrec[i, j, c, b] (W,H,5,B), xtrue (W,H,2*nsites,B), directionobs (W,H,nsites[,B]).
"""
import numpy as np

F = np.float32


def sinv(x, minx=F(1e-3)):
    """src/model.jl:7-10"""
    return (F(1) / np.maximum(x, minx)).astype(F)


def vector2_covariance(xrec, obsoper=lambda x: x):
    """src/vector2.jl:13-24: P = L L' with L = [L11 0; L21 L22]"""
    L11 = obsoper(xrec[:, :, 2, :]); L21 = obsoper(xrec[:, :, 3, :]); L22 = obsoper(xrec[:, :, 4, :])
    return (L11 ** 2).astype(F), (L11 * L21).astype(F), (L21 ** 2 + L22 ** 2).astype(F)


def vector2_mean(xrec, P, obsoper=lambda x: x):
    """src/vector2.jl:46-50"""
    P11, P12, P22 = P
    a, b = obsoper(xrec[:, :, 0, :]), obsoper(xrec[:, :, 1, :])
    return (a * P11 + b * P12).astype(F), (a * P12 + b * P22).astype(F)


def vector2_projection(uv, P, directionobs):
    """src/vector2.jl:27-44: directionobs * pi / 180 in Float32 (Float32 array times the irrational pi, then / 180)"""
    P11, P12, P22 = P
    rad = (directionobs.astype(F) * F(np.pi)) / F(180)
    s, c = np.sin(rad).astype(F), np.cos(rad).astype(F)
    m = (s * uv[0] + c * uv[1]).astype(F)
    s2 = (s ** 2 * P11 + F(2) * s * c * P12 + c ** 2 * P22).astype(F)
    return m, s2


def costfun_single(m_rec, s2_rec, m_true, s2_true, mask, truth_uncertain):
    """src/model.jl:54-100 (Laplacian penalties at their default 0); sums in Float64 like the device code's block sums"""
    n = mask.sum()
    maskf = mask.astype(F)
    if truth_uncertain:
        ratio = (s2_rec / s2_true) * maskf + (F(1) - maskf)
        diff2 = ((m_rec - m_true) ** 2 + s2_true) * maskf
        return (np.log(ratio).astype(np.float64).sum() + (diff2 / s2_rec).astype(np.float64).sum()) / n
    nc = s2_rec * maskf + (F(1) - maskf)
    diff2 = (m_rec - m_true) ** 2 * maskf
    return (np.log(nc).astype(np.float64).sum() + (diff2 / s2_rec).astype(np.float64).sum()) / n


def vector2_costfun(xrec, xtrue, truth_uncertain, directionobs):
    """src/vector2.jl:53-100"""
    xrec = np.asarray(xrec, F); xtrue = np.asarray(xtrue, F); directionobs = np.asarray(directionobs, F)
    nsites = directionobs.shape[2]

    def obsoper(x):                                   # (W,H,B) -> (W,H,nsites,B), :58-62
        return np.repeat(x[:, :, None, :], nsites, axis=2)

    P = vector2_covariance(xrec, obsoper)
    uv = vector2_mean(xrec, P, obsoper)
    d = directionobs if directionobs.ndim == 4 else directionobs[:, :, :, None]
    m_rec, s2_rec = vector2_projection(uv, P, d)
    s2_rec = np.maximum(s2_rec, F(1e-4))              # :78
    inv_true = xtrue[:, :, 1:2 * nsites:2, :]
    s2_true = sinv(inv_true)                          # :85
    m_true = (xtrue[:, :, 0:2 * nsites:2, :] * s2_true).astype(F)
    mask = inv_true != 0                              # :90
    cost = costfun_single(m_rec, s2_rec, m_true, s2_true, mask, truth_uncertain)
    if np.isnan(cost):
        raise ValueError("cost function value is NaN: %r" % cost)     # :95-97
    return float(cost)


def vector2_costfun_grad(xrec, xtrue, truth_uncertain, directionobs):
    """(cost, d cost / d xrec) in Float64 through torch autograd of the same expression (what Zygote differentiates in the reference)."""
    import torch
    x = torch.tensor(np.asarray(xrec, np.float64), requires_grad=True)
    xt = torch.tensor(np.asarray(xtrue, np.float64))
    d = torch.tensor(np.asarray(directionobs, np.float64))
    nsites = d.shape[2]
    if d.ndim == 3:
        d = d[:, :, :, None]
    rep = lambda a: a[:, :, None, :].expand(-1, -1, nsites, -1)
    L11, L21, L22 = rep(x[:, :, 2, :]), rep(x[:, :, 3, :]), rep(x[:, :, 4, :])
    P11, P12, P22 = L11 ** 2, L11 * L21, L21 ** 2 + L22 ** 2
    a, b = rep(x[:, :, 0, :]), rep(x[:, :, 1, :])
    u, v = a * P11 + b * P12, a * P12 + b * P22
    rad = d * float(F(np.pi)) / 180.0
    s, c = torch.sin(rad), torch.cos(rad)
    m = s * u + c * v
    s2 = torch.clamp(s ** 2 * P11 + 2 * s * c * P12 + c ** 2 * P22, min=1e-4)
    inv_true = xt[:, :, 1:2 * nsites:2, :]
    s2t = 1.0 / torch.clamp(inv_true, min=1e-3)
    mt = xt[:, :, 0:2 * nsites:2, :] * s2t
    mask = (inv_true != 0).double()
    n = mask.sum()
    if truth_uncertain:
        cost = (torch.log((s2 / s2t) * mask + (1 - mask)).sum() + ((((m - mt) ** 2 + s2t) * mask) / s2).sum()) / n
    else:
        cost = (torch.log(s2 * mask + (1 - mask)).sum() + (((m - mt) ** 2 * mask) / s2).sum()) / n
    cost.backward()
    return float(cost.detach()), x.grad.numpy()
