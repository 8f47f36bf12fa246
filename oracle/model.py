"""Oracle for the model, loss and optimiser (src/model.jl, src/flux.jl).

torch CPU ops restate the third-party Flux/NNlib/Optimisers semantics
(PARITY UNPINNED, see oracle/__init__.py); everything DINCAE defines itself
(head, cost function, Laplacian penalty, refinement wiring, regularisers,
learning-rate schedule, epoch mean) follows the cited lines literally.

torch layout: x[b, c, j, i] with j = lat (Julia dim 2), i = lon (Julia dim 1),
i.e. a Julia WHCN array is `np.transpose(a, (3,2,1,0))` here.
"""
import math
from dataclasses import dataclass, field
from typing import List, Tuple
import numpy as np
import torch
import torch.nn.functional as F


@dataclass
class ModelConfig:
    W: int
    H: int
    ninput: int
    noutput: int = 1
    nfilter: List[int] = field(default_factory=lambda: [16, 24, 36, 54])   # enc_nfilter_internal
    skipconnections: Tuple[int, ...] = None        # 1-based levels; default 2:(L+1)  (src/model.jl:289)
    upsampling_method: str = "nearest"
    loss_weights_refine: Tuple[float, ...] = (1.0,)
    min_std_err: float = 0.006737946999085467
    truth_uncertain: bool = False
    regularization_L1_beta: float = 0.0
    regularization_L2_beta: float = 0.0
    laplacian_penalty: float = 0.0
    laplacian_error_penalty: float = None
    clip_grad: float = 5.0

    def __post_init__(self):
        if self.skipconnections is None:
            self.skipconnections = tuple(range(2, len(self.nfilter) + 2))
        if self.laplacian_error_penalty is None:
            self.laplacian_error_penalty = self.laplacian_penalty

    @property
    def gamma(self):                       # src/model.jl:318
        return math.log(self.min_std_err ** (-2))

    @property
    def nchains(self):                     # src/model.jl:322-340
        return len(self.loss_weights_refine)

    def enc_dec(self, chain):
        enc = [self.ninput] + list(self.nfilter)
        dec = [2 * self.noutput] + list(self.nfilter)
        if chain > 0:
            enc[0] += dec[0]               # src/model.jl:333
        return enc, dec


def param_shapes(cfg):
    """Flux.trainables order: per chain enc1.W, enc1.b, ...,
    encL.W, encL.b, decL.W, decL.b, ..., dec1.W, dec1.b; W is (kw,kh,cin,cout)."""
    shapes = []
    L = len(cfg.nfilter)
    for ch in range(cfg.nchains):
        enc, dec = cfg.enc_dec(ch)
        for l in range(L):
            shapes.append((3, 3, enc[l], enc[l + 1])); shapes.append((enc[l + 1],))
        for l in reversed(range(L)):
            shapes.append((3, 3, dec[l + 1], dec[l])); shapes.append((dec[l],))
    return shapes


def num_params(cfg):
    return sum(int(np.prod(s)) for s in param_shapes(cfg))


def init_params(cfg, seed=1):
    """glorot_uniform weights (U(-s,s), s = sqrt(6/(9cin+9cout))), zero biases;
    returned flat in Flux order, each W flattened column-major."""
    rng = np.random.default_rng(seed)
    out = []
    for s in param_shapes(cfg):
        if len(s) == 4:
            lim = math.sqrt(6.0 / (9 * s[2] + 9 * s[3]))
            out.append(rng.uniform(-lim, lim, size=int(np.prod(s))).astype(np.float32))
        else:
            out.append(np.zeros(s[0], np.float32))
    return np.concatenate(out)


def unpack(cfg, flat, dtype=torch.float64, requires_grad=False):
    """flat Flux-order vector -> list (per chain) of dicts of torch conv weights.

    Flux Conv = NNlib true convolution (kernel flipped):
    torch cross-correlation weight wt[co,ci,kh,kw] = W[2-kw, 2-kh, ci, co].
    The leaf tensor is the flat vector so gradients come back in Flux order."""
    theta = torch.as_tensor(np.asarray(flat), dtype=dtype).clone().requires_grad_(requires_grad)
    chains, pos = [], 0
    L = len(cfg.nfilter)
    it = iter(param_shapes(cfg))

    def take():
        nonlocal pos
        s = next(it)
        n = int(np.prod(s))
        v = theta[pos:pos + n]
        pos += n
        if len(s) == 4:
            w = v.reshape(s[3], s[2], s[1], s[0])          # column-major (kw,kh,ci,co) -> [co,ci,kh,kw]
            return torch.flip(w, dims=(2, 3))
        return v
    for ch in range(cfg.nchains):
        enc = [(take(), take()) for _ in range(L)]
        dec_rev = [(take(), take()) for _ in range(L)]
        chains.append({"enc": enc, "dec": list(reversed(dec_rev))})
    return theta, chains


def _level(x, p, cfg, l):
    """recmodel4, src/model.jl:242-284 (l is 0-based here; reference l = l+1)."""
    L = len(cfg.nfilter)
    if l == L:
        return x                                         # :263-264 identity
    Hc, Wc = x.shape[2], x.shape[3]
    odd = (Hc % 2, Wc % 2)
    w, b = p["enc"][l]
    y = F.relu(F.conv2d(x, w, b, padding=1))                                        # :267
    y = F.avg_pool2d(y, 2, 2, padding=odd, count_include_pad=True)                  # :268 MeanPool(pad=odd)
    y = _level(y, p, cfg, l + 1)                                                     # :269
    if cfg.upsampling_method == "nearest":                                           # :270, src/flux.jl:4
        y = F.interpolate(y, scale_factor=2, mode="nearest")
    else:
        y = F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=True)
    y = y[:, :, :Hc, :Wc]                                                            # :271 croppadding (drops LAST)
    w, b = p["dec"][l]
    y = F.conv2d(y, w, b, padding=1)                                                 # :272
    if l > 0:
        y = F.relu(y)                                                                # :251-256
    if (l + 1) in cfg.skipconnections:
        y = y + x                                                                    # :277-279 SumSkip
    return y


def sinv(x, minx=1e-3):                                   # src/model.jl:7-10
    return 1.0 / torch.clamp(x, min=minx)


def transform_msigma2(x, gamma):
    """src/model.jl:16-41.  x: (B, 2*nout, H, W) raw -> (m, sigma2) interleaved."""
    outs = []
    for iv in range(x.shape[1] // 2):
        loginv = x[:, 2 * iv + 1:2 * iv + 2]
        inv = torch.exp(torch.clamp(loginv, max=gamma))
        s2 = sinv(inv)
        outs += [x[:, 2 * iv:2 * iv + 1] * s2, s2]
    return torch.cat(outs, dim=1)


def sum_laplacian_penalty(penalty, x):                    # src/model.jl:44-52 (dims 1,2 = W,H)
    if penalty == 0:
        return 0.0
    di = x[..., 2:] - 2 * x[..., 1:-1] + x[..., :-2]
    dj = x[..., 2:, :] - 2 * x[..., 1:-1, :] + x[..., :-2, :]
    return penalty * ((di ** 2).sum() + (dj ** 2).sum())


def costfun_single(m_rec, s2_rec, xt1, xt2, truth_uncertain, lap=0.0, lap_err=0.0):
    """src/model.jl:54-129."""
    s2_true = sinv(xt2)                                   # :109
    m_true = xt1 * s2_true                                # :110
    M = (xt2 != 0).to(m_rec.dtype)                        # :114
    n = M.sum().detach()                                  # :61
    if truth_uncertain:                                   # :79-86
        ratio = (s2_rec / s2_true) * M + (1 - M)
        d2 = ((m_rec - m_true) ** 2 + s2_true) * M
        cost = (torch.log(ratio).sum() + (d2 / s2_rec).sum()) / n
    else:                                                 # :89-92
        s2nc = s2_rec * M + (1 - M)
        d2 = (m_rec - m_true) ** 2 * M
        cost = (torch.log(s2nc).sum() + (d2 / s2_rec).sum()) / n
    return cost + sum_laplacian_penalty(lap, m_rec) + sum_laplacian_penalty(lap_err, s2_rec)


def costfun(xrec, xtrue, cfg):                            # src/model.jl:132-148
    tot = 0.0
    for iv in range(xtrue.shape[1] // 2):
        tot = tot + costfun_single(xrec[:, 2 * iv:2 * iv + 1], xrec[:, 2 * iv + 1:2 * iv + 2],
                                   xtrue[:, 2 * iv:2 * iv + 1], xtrue[:, 2 * iv + 1:2 * iv + 2],
                                   cfg.truth_uncertain, cfg.laplacian_penalty, cfg.laplacian_error_penalty)
    return tot


def raw_outputs(chains, xin, cfg):
    """Raw 2*nout-channel output of every chain (src/model.jl:182-187, :196-203)."""
    outs = []
    xout = _level(xin, chains[0], cfg, 0)
    outs.append(xout)
    for p in chains[1:]:
        xout = _level(torch.cat([xout, xin], dim=1), p, cfg, 0)     # cat(xout,xin,dims=3): xout FIRST (T9)
        outs.append(xout)
    return outs


def forward(cfg, flat, xin, dtype=torch.float64):
    """StepModel call (src/model.jl:182-189): (m, sigma2) of the LAST chain.  xin WHCN numpy."""
    _, chains = unpack(cfg, flat, dtype)
    x = torch.as_tensor(np.transpose(np.asarray(xin), (3, 2, 1, 0)).copy(), dtype=dtype)
    with torch.no_grad():
        y = transform_msigma2(raw_outputs(chains, x, cfg)[-1], cfg.gamma)
    return np.asfortranarray(np.transpose(y.numpy(), (3, 2, 1, 0)))


def reg_terms(cfg, theta):
    """src/model.jl:210-219: over ALL trainables incl. biases (T5)."""
    r = 0.0
    if cfg.regularization_L1_beta != 0:
        r = r + cfg.regularization_L1_beta * theta.abs().sum()
    if cfg.regularization_L2_beta != 0:
        r = r + cfg.regularization_L2_beta * (theta ** 2).sum()
    return r


def loss_and_grad(cfg, flat, xin, xtrue, dtype=torch.float64, point_truth=None):
    """loss_function (src/model.jl:192-222) and its gradient in Flux order.

    point_truth: list (per sample) of (pos (P,2), x (P,2)) => the point-cloud cost
    (src/points.jl:354-426) instead of the gridded one."""
    theta, chains = unpack(cfg, flat, dtype, requires_grad=True)
    x = torch.as_tensor(np.transpose(np.asarray(xin), (3, 2, 1, 0)).copy(), dtype=dtype)
    if point_truth is None:
        xt = torch.as_tensor(np.transpose(np.asarray(xtrue), (3, 2, 1, 0)).copy(), dtype=dtype)
    loss = 0.0
    for wgt, raw in zip(cfg.loss_weights_refine, raw_outputs(chains, x, cfg)):
        xrec = transform_msigma2(raw, cfg.gamma)
        if point_truth is None:
            loss = loss + wgt * costfun(xrec, xt, cfg)
        else:
            loss = loss + wgt * point_costfun(xrec, point_truth, cfg)
    loss = loss + reg_terms(cfg, theta)
    loss.backward()
    return float(loss.detach()), theta.grad.numpy().copy()


def interp_torch(pos, A):
    """Differentiable interpnd (src/points.jl:6-43): A is (H, W) torch [j,i]; pos 1-based (i,j)."""
    pos = torch.as_tensor(np.asarray(pos, np.float32))
    ind = torch.floor(pos).long()
    c = (pos - ind.to(pos.dtype)).to(A.dtype)
    out = 0.0
    for oj in (0, 1):
        for oi in (0, 1):
            cc = (c[:, 0] if oi else 1 - c[:, 0]) * (c[:, 1] if oj else 1 - c[:, 1])
            out = out + A[ind[:, 1] - 1 + oj, ind[:, 0] - 1 + oi] * cc
    return out


def point_costfun(xrec, point_truth, cfg):
    """src/points.jl:354-426: gather (m, sigma2) at the observation positions of every
    sample, concatenate over the batch, generic NLL (n = all valid obs of the batch),
    Laplacian penalties on the gridded fields."""
    ms, ss, t1, t2 = [], [], [], []
    for b, (pos, xx) in enumerate(point_truth):
        ms.append(interp_torch(pos, xrec[b, 0])); ss.append(interp_torch(pos, xrec[b, 1]))
        xx = torch.as_tensor(np.asarray(xx), dtype=xrec.dtype)
        t1.append(xx[:, 0]); t2.append(xx[:, 1])
    cost = costfun_single(torch.cat(ms), torch.cat(ss), torch.cat(t1), torch.cat(t2), cfg.truth_uncertain)
    return (cost + sum_laplacian_penalty(cfg.laplacian_penalty, xrec[:, 0:1])
            + sum_laplacian_penalty(cfg.laplacian_error_penalty, xrec[:, 1:2]))


class Adam:
    """OptimiserChain(ClipGrad(delta), Adam(eta)) of src/flux.jl:7-21 (Optimisers.jl
    rules restated).  State arithmetic in float32 like Flux."""

    def __init__(self, n, clip_grad=5.0, beta=(0.9, 0.999), eps=1e-8, dtype=np.float32):
        self.m = np.zeros(n, dtype); self.v = np.zeros(n, dtype)
        self.bt = [1.0, 1.0]; self.beta = beta; self.eps = eps; self.clip = clip_grad; self.dtype = dtype

    def step(self, w, g, lr):
        dt = self.dtype
        g = np.asarray(g, dt)
        if self.clip is not None:
            g = np.clip(g, dt(-self.clip), dt(self.clip))             # ClipGrad: element-wise clamp (T4)
        b1, b2 = self.beta
        self.bt[0] *= b1; self.bt[1] *= b2
        self.m = (dt(b1) * self.m + dt(1 - b1) * g).astype(dt)
        self.v = (dt(b2) * self.v + dt(1 - b2) * g * g).astype(dt)
        mhat = self.m / dt(1 - self.bt[0])
        vhat = self.v / dt(1 - self.bt[1])
        return (np.asarray(w, dt) - dt(lr) * mhat / (np.sqrt(vhat) + dt(self.eps))).astype(np.float32)


def train_step(cfg, flat, opt, xin, xtrue, lr, dtype=torch.float64, point_truth=None):
    """One mini-batch of train_epoch! (src/flux.jl:28-36).  Returns (loss, new flat)."""
    loss, g = loss_and_grad(cfg, flat, xin, xtrue, dtype, point_truth)
    return loss, opt.step(flat, g, lr)


def learning_rate(lr0, epoch, decay_epoch):               # src/model.jl:590 (epoch 1-based)
    return lr0 * 0.5 ** (epoch / decay_epoch)
