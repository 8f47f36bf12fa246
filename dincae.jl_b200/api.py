"""Host-side mirror of the reference's operator interface over the C-ABI (no compute here).

Reference surface mirrored (file:line in gher-uliege/DINCAE.jl):
  interpnd / interp_adjn                     src/points.jl:29-34, :73-77
  NCData + getxy! + DataBatches              src/data.jl:250-369, :391-517, :771-868
  genmodel / StepModel / loss_function       src/model.jl:286-363, :151-222
  train_init / train_epoch!                  src/flux.jl:7-40
  reconstruct / reconstruct_points           src/model.jl:446-679, src/points.jl:509-696
Arrays are numpy, indexed like Julia ([i,j,...] = [lon,lat,...]) and Fortran-ordered, so the
pointer handed to the library is the memory a Julia caller would hand over.
"""
import ctypes as C
import math
import os
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import lib as L


def _f32(a, order="F"):
    return np.asarray(a, dtype=np.float32, order=order)


def _host_nccl_first():
    """PyTorch bundles its own libnccl.so.2; the loader keeps one object per soname, so let torch load its copy before the
    library dlopens NCCL (otherwise a later `import torch` would be handed the system copy and fail on missing symbols)."""
    try:
        import torch  # noqa: F401
    except Exception:
        pass


class Context:
    """One GPU (replaces `_to_device(Atype)`, src/flux.jl:47-49)."""

    def __init__(self, device: int = 0, torch_stream=None):
        self._lib = L.load()
        h = C.c_void_p()
        L.check(self._lib.dincae_ctx_create(device, C.byref(h)))
        self.handle = h
        self.device = device
        self._stream = None
        if torch_stream is not None:
            self.set_stream(torch_stream)

    def set_stream(self, torch_stream):
        """Run on a torch.cuda.Stream so torch events / NCCL collectives order with the kernels."""
        self._stream = torch_stream
        L.check(self._lib.dincae_ctx_set_stream(self.handle, C.c_void_p(torch_stream.cuda_stream)))

    def sync(self):
        L.check(self._lib.dincae_sync(self.handle))

    # -- data-parallel communicator (one process per GPU; the collective lives in the library, NCCL over NVLink)
    @staticmethod
    def comm_unique_id() -> bytes:
        """Rank 0: the 128-byte NCCL rendezvous id, to be handed to every rank by the host's own means."""
        _host_nccl_first()
        buf = C.create_string_buffer(L.COMM_ID_BYTES)
        L.check(L.load().dincae_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, nranks: int, rank: int, unique_id: bytes):
        if len(unique_id) != L.COMM_ID_BYTES:
            raise ValueError("unique_id must be %d bytes" % L.COMM_ID_BYTES)
        _host_nccl_first()
        L.check(self._lib.dincae_comm_init_rank(self.handle, nranks, rank, C.c_char_p(unique_id)))

    def comm_info(self):
        n, r, v = C.c_int32(), C.c_int32(), C.c_int32()
        L.check(self._lib.dincae_comm_info(self.handle, C.byref(n), C.byref(r), C.byref(v)))
        return {"nranks": n.value, "rank": r.value, "nccl_version": v.value}

    def allreduce(self, dev_buf, n):
        L.check(self._lib.dincae_allreduce(self.handle, L.ptr(dev_buf), int(n)))

    @property
    def launches(self) -> int:
        return int(self._lib.dincae_launch_count(self.handle))

    def close(self):
        if self.handle:
            self._lib.dincae_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx = None


def default_context() -> Context:
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(0)
    return _default_ctx


# ---- K4 ------------------------------------------------------------------------------------
def interpnd(pos, A, ctx: Optional[Context] = None):
    """interpnd(pos, A) (src/points.jl:29-34): pos (P,2) 1-based, A (W,H) -> (P,)"""
    ctx = ctx or default_context()
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 2)
    A = _f32(A)
    out = np.zeros(len(pos), np.float32)
    L.check(ctx._lib.dincae_interp(ctx.handle, L.ptr(pos), len(pos), L.ptr(A), A.shape[0], A.shape[1], L.ptr(out), L.HOST))
    return out


def interp_adjn(pos, values, sz, ctx: Optional[Context] = None):
    """interp_adjn(pos, values, sz) (src/points.jl:73-77) -> (W,H)"""
    ctx = ctx or default_context()
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 2)
    values = np.ascontiguousarray(values, np.float32)
    if len(values) != len(pos):
        raise ValueError("interp_adjn: %d positions but %d values" % (len(pos), len(values)))
    out = np.zeros(sz, np.float32, order="F")
    L.check(ctx._lib.dincae_interp_adjoint(ctx.handle, L.ptr(pos), L.ptr(values), len(pos), sz[0], sz[1], L.ptr(out), L.HOST))
    return out


def vector2_costfun(xrec, xtrue, truth_uncertain, directionobs, ctx: Optional[Context] = None, with_grad=False):
    """vector2_costfun(xrec, xtrue, truth_uncertain, directionobs) (src/vector2.jl:53-100): xrec (W,H,5,B), xtrue (W,H,2*nsites,B),
    directionobs (W,H,nsites) or (W,H,nsites,B) in degrees -> cost, or (cost, d cost / d xrec) with with_grad=True."""
    ctx = ctx or default_context()
    xrec = np.asfortranarray(xrec, np.float32); xtrue = np.asfortranarray(xtrue, np.float32)
    directionobs = np.asfortranarray(directionobs, np.float32)
    W, H, five, B = xrec.shape
    nsites = directionobs.shape[2]
    if five != 5 or xtrue.shape != (W, H, 2 * nsites, B) or directionobs.shape[:2] != (W, H) or directionobs.ndim not in (3, 4):
        raise ValueError("vector2_costfun: xrec %s, xtrue %s, directionobs %s do not fit together" % (xrec.shape, xtrue.shape, directionobs.shape))
    if directionobs.ndim == 4 and directionobs.shape[3] != B:
        raise ValueError("vector2_costfun: directionobs has %d samples, xrec %d" % (directionobs.shape[3], B))
    cost = C.c_double()
    grad = np.zeros(xrec.shape, np.float32, order="F") if with_grad else None
    L.check(ctx._lib.dincae_vector2_loss(ctx.handle, L.ptr(xrec), L.ptr(xtrue), L.ptr(directionobs), W, H, nsites, B,
                                         int(directionobs.ndim == 4), int(bool(truth_uncertain)), L.HOST, C.byref(cost),
                                         L.ptr(grad) if with_grad else None))
    return (cost.value, grad) if with_grad else cost.value


def interp_adjn_noalloc(pos_dev, values_dev, P, sz, out_dev, ctx: Optional[Context] = None):
    """interp_adjn for callers that may not allocate GPU memory (their device memory belongs to another
    pool, as under CUDA.jl's allocator around a ccall).  All three buffers are device memory the caller
    owns (torch tensors or raw addresses): pos (P,2), values (P,), out (W,H) which is cleared and filled.
    The scratch is a fixed array inside the library: at most 65536 points per call (more raises
    DincaeError/ESIZE), shared by every call on the device -- do not let two calls overlap on the GPU.
    The first call per device resolves the array's address and should happen before any stream capture.
    Asynchronous on the context's stream; bit-identical to interp_adjn."""
    ctx = ctx or default_context()
    L.check(ctx._lib.dincae_interp_adjoint_noalloc(ctx.handle, L.ptr(pos_dev), L.ptr(values_dev), int(P), sz[0], sz[1], L.ptr(out_dev)))
    return out_dev


def interp_cells(pos, ctx: Optional[Context] = None):
    ctx = ctx or default_context()
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 2)
    out = np.zeros((len(pos), 2), np.int32)
    L.check(ctx._lib.dincae_interp_cells(ctx.handle, L.ptr(pos), len(pos), L.ptr(out), L.HOST))
    return out


def posindices(grid, lon, lat, ctx: Optional[Context] = None):
    """posindices! for one time step (src/points.jl:98-134): returns (pos (P,2), inside (P,) bool)."""
    ctx = ctx or default_context()
    lon = np.ascontiguousarray(lon, np.float32); lat = np.ascontiguousarray(lat, np.float32)
    P = len(lon)
    pos = np.zeros((P, 2), np.float32); inside = np.zeros(P, np.uint8)
    gx, gy = np.asarray(grid[0], np.float64), np.asarray(grid[1], np.float64)
    L.check(ctx._lib.dincae_posindices(ctx.handle, L.ptr(lon), L.ptr(lat), P, gx[0], gx[-1], len(gx), gy[0], gy[-1], len(gy),
                                       L.ptr(pos), L.ptr(inside), L.HOST))
    return pos, inside.astype(bool)


# ---- model ------------------------------------------------------------------------------------
@dataclass
class ModelConfig:
    """Keyword surface of genmodel / reconstruct that shapes the model (src/model.jl:286-299)."""
    W: int
    H: int
    ninput: int
    noutput: int = 1
    enc_nfilter_internal: Sequence[int] = (16, 24, 36, 54)
    skipconnections: Optional[Sequence[int]] = None           # 1-based levels, default 2:(L+1)
    upsampling_method: str = "nearest"
    loss_weights_refine: Tuple[float, ...] = (1.0,)
    min_std_err: float = 0.006737946999085467
    truth_uncertain: bool = False
    regularization_L1_beta: float = 0.0
    regularization_L2_beta: float = 0.0
    laplacian_penalty: float = 0.0
    laplacian_error_penalty: Optional[float] = None
    clip_grad: Optional[float] = 5.0
    adam_beta: Tuple[float, float] = (0.9, 0.999)
    adam_eps: float = 1e-8
    leaky_slope: float = 0.0
    precision: str = "tf32"
    max_batch: int = 50

    def to_struct(self) -> L.Config:
        nf = list(self.enc_nfilter_internal)
        nl = len(nf)
        if nl > L.MAX_LEVELS:
            raise ValueError("at most %d levels" % L.MAX_LEVELS)
        if len(self.loss_weights_refine) > L.MAX_CHAINS:
            raise ValueError("at most %d refinement chains" % L.MAX_CHAINS)
        skips = range(2, nl + 2) if self.skipconnections is None else self.skipconnections
        mask = 0
        for s in skips:
            if 1 <= s <= nl:            # level L+1 is the identity (src/model.jl:263-264): no effect
                mask |= 1 << (s - 1)
        if self.upsampling_method not in ("nearest", "bilinear"):
            raise ValueError("upsampling_method must be :nearest or :bilinear")
        c = L.Config()
        c.W, c.H, c.ninput, c.noutput, c.nlevels = self.W, self.H, self.ninput, self.noutput, nl
        for i, f in enumerate(nf):
            c.nfilter[i] = int(f)
        c.skip_mask = mask
        c.upsampling = L.UP_NEAREST if self.upsampling_method == "nearest" else L.UP_BILINEAR
        c.nchains = len(self.loss_weights_refine)
        for i, w in enumerate(self.loss_weights_refine):
            c.loss_weights[i] = float(w)
        c.gamma = math.log(self.min_std_err ** (-2))                 # src/model.jl:318
        c.truth_uncertain = int(bool(self.truth_uncertain))
        c.l1_beta = float(self.regularization_L1_beta)
        c.l2_beta = float(self.regularization_L2_beta)
        c.laplacian_penalty = float(self.laplacian_penalty)
        lep = self.laplacian_penalty if self.laplacian_error_penalty is None else self.laplacian_error_penalty
        c.laplacian_error_penalty = float(lep)
        c.clip_grad = float(self.clip_grad) if self.clip_grad is not None else 0.0
        c.adam_beta1, c.adam_beta2, c.adam_eps = self.adam_beta[0], self.adam_beta[1], self.adam_eps
        c.leaky_slope = float(self.leaky_slope)
        c.precision = {"fp32": L.PREC_FP32, "tf32": L.PREC_TF32, "tf32x1": L.PREC_TF32X1}[self.precision]
        c.max_batch = int(self.max_batch)
        return c


class Model:
    """genmodel(...) |> device + train_init (src/model.jl:546-559, src/flux.jl:7-21)."""

    def __init__(self, ctx: Context, cfg: ModelConfig, seed: Optional[int] = 0):
        self.ctx, self.cfg, self._lib = ctx, cfg, ctx._lib
        self._cstruct = cfg.to_struct()
        h = C.c_void_p()
        L.check(self._lib.dincae_model_create(ctx.handle, C.byref(self._cstruct), C.byref(h)))
        self.handle = h
        n = C.c_int64()
        L.check(self._lib.dincae_model_num_params(h, C.byref(n)))
        self.num_params = n.value
        if seed is not None:
            L.check(self._lib.dincae_model_init_glorot(h, seed))

    # -- parameters in Flux.trainables order
    def get_params(self):
        out = np.zeros(self.num_params, np.float32)
        L.check(self._lib.dincae_model_get_params(self.handle, L.ptr(out), out.size))
        return out

    def set_params(self, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        L.check(self._lib.dincae_model_set_params(self.handle, L.ptr(flat), flat.size))

    def get_opt_state(self):
        m = np.zeros(self.num_params, np.float32); v = np.zeros_like(m); t = C.c_int64()
        L.check(self._lib.dincae_model_get_opt_state(self.handle, L.ptr(m), L.ptr(v), m.size, C.byref(t)))
        return m, v, t.value

    def set_opt_state(self, m, v, t):
        m = np.ascontiguousarray(m, np.float32); v = np.ascontiguousarray(v, np.float32)
        L.check(self._lib.dincae_model_set_opt_state(self.handle, L.ptr(m), L.ptr(v), m.size, int(t)))

    # -- checkpoint: weights in Flux.trainables order as Flux.state/loadmodel! walk them (src/model.jl:366-399,
    #    :599-620), plus the Adam state the reference does not save (so a run can really be resumed)
    def flux_layout(self):
        """[(name, shape, offset)] of the flat parameter vector: per chain enc1..encL then decL..dec1, W as (3,3,cin,cout)
        column-major (kw fastest, Flux's true-convolution kernel) followed by b (cout)."""
        nf = list(self.cfg.enc_nfilter_internal)
        out, off = [], 0
        for c in range(len(self.cfg.loss_weights_refine)):
            cin0 = self.cfg.ninput + (2 * self.cfg.noutput if c > 0 else 0)
            enc = [cin0] + nf
            dec = [2 * self.cfg.noutput] + nf
            layers = [("chain%d.enc%d" % (c + 1, l + 1), enc[l], enc[l + 1]) for l in range(len(nf))]
            layers += [("chain%d.dec%d" % (c + 1, l + 1), dec[l + 1], dec[l]) for l in reversed(range(len(nf)))]
            for name, ci, co in layers:
                out.append((name + ".weight", (3, 3, ci, co), off)); off += 9 * ci * co
                out.append((name + ".bias", (co,), off)); off += co
        assert off == self.num_params, (off, self.num_params)
        return out

    def save_checkpoint(self, path, **extra):
        m, v, t = self.get_opt_state()
        cfg = {k: (list(val) if isinstance(val, (tuple, list)) else val) for k, val in self.cfg.__dict__.items()}
        np.savez(path, params=self.get_params(), adam_m=m, adam_v=v, adam_t=np.int64(t),
                 config=np.array(repr(cfg)), layout=np.array(repr(self.flux_layout())), **extra)

    def load_checkpoint(self, path, load_optimizer=True):
        with np.load(path, allow_pickle=False) as z:
            if z["params"].size != self.num_params:
                raise ValueError("checkpoint holds %d parameters, the model has %d" % (z["params"].size, self.num_params))
            self.set_params(z["params"])
            if load_optimizer and "adam_m" in z:
                self.set_opt_state(z["adam_m"], z["adam_v"], int(z["adam_t"]))

    # -- calls; x may be numpy (host) or an int / torch tensor (device)
    @staticmethod
    def _loc(a):
        return L.HOST if isinstance(a, np.ndarray) else L.DEVICE

    def _batch(self, xin, B):
        if isinstance(xin, np.ndarray):
            xin = _f32(xin)
            return xin, xin.shape[3]
        if B is None:
            raise ValueError("B is required with device pointers")
        return xin, B

    def __call__(self, xin, B=None, out=None):
        """model(xin) -> (W,H,2*noutput,B) = (mean, sigma^2) (src/model.jl:182-189)."""
        xin, B = self._batch(xin, B)
        loc = self._loc(xin)
        if out is None:
            if loc != L.HOST:
                raise ValueError("pass `out` (device) with device inputs")
            out = np.zeros((self.cfg.W, self.cfg.H, 2 * self.cfg.noutput, B), np.float32, order="F")
        L.check(self._lib.dincae_forward(self.handle, L.ptr(xin), B, L.ptr(out), loc))
        return out

    def loss(self, xin, xtrue, B=None):
        xin, B = self._batch(xin, B)
        if isinstance(xtrue, np.ndarray):
            xtrue = _f32(xtrue)
        out = C.c_double()
        L.check(self._lib.dincae_loss(self.handle, L.ptr(xin), L.ptr(xtrue), B, self._loc(xin), C.byref(out)))
        return out.value

    def train_step(self, xin, xtrue, lr, B=None, want_loss=True):
        xin, B = self._batch(xin, B)
        if isinstance(xtrue, np.ndarray):
            xtrue = _f32(xtrue)
        out = C.c_float()
        L.check(self._lib.dincae_train_step(self.handle, L.ptr(xin), L.ptr(xtrue), B, lr, self._loc(xin),
                                            C.byref(out) if want_loss else None))
        return out.value if want_loss else None

    def train_step_gridded(self, ds, time_idx, imask, lr, seed=0, sample_ctr0=0, n_valid_global=None, want_loss=True):
        """One mini-batch of train_epoch! straight from the device-resident dataset `ds`: batch assembly (getxy!) + training step as
        one call (dincae_train_step_gridded); with n_valid_global the data-parallel form on this rank's shard."""
        time_idx = np.ascontiguousarray(time_idx, np.int64)
        im = None if imask is None else np.ascontiguousarray(imask, np.int64)
        out = C.c_float()
        po = C.byref(out) if want_loss else None
        if n_valid_global is None:
            L.check(self._lib.dincae_train_step_gridded(self.handle, ds.handle, L.ptr(time_idx), L.ptr(im), seed, sample_ctr0,
                                                        len(time_idx), lr, po))
        else:
            nv = np.ascontiguousarray(n_valid_global, np.int64)
            L.check(self._lib.dincae_train_step_dp_gridded(self.handle, ds.handle, L.ptr(time_idx), L.ptr(im), seed, sample_ctr0,
                                                           len(time_idx), L.ptr(nv), lr, po))
        return out.value if want_loss else None

    def step_graph(self, enable=None):
        """Switch the CUDA-graph replay of the gridded gradient step on/off for this model (None: only query);
        returns the number of captured graphs (dincae_model_step_graph)."""
        n = C.c_int32()
        L.check(self._lib.dincae_model_step_graph(self.handle, -1 if enable is None else int(bool(enable)), C.byref(n)))
        return n.value

    def upload(self, slot, xin, xtrue):
        """Start the host->device copy of a mini-batch into prefetch slot 0/1 (PrefetchDataIter, src/data.jl:872-904)."""
        xin = _f32(xin); xtrue = _f32(xtrue)
        L.check(self._lib.dincae_batch_upload(self.handle, slot, L.ptr(xin), L.ptr(xtrue), xin.shape[3]))
        return xin, xtrue            # keep these alive until train_step_slot returns

    def train_step_slot(self, slot, B, lr, want_loss=True):
        out = C.c_float()
        L.check(self._lib.dincae_train_step_slot(self.handle, slot, B, lr, C.byref(out) if want_loss else None))
        return out.value if want_loss else None

    def grad_step_slot(self, slot, B, n_valid_global):
        nv = np.ascontiguousarray(n_valid_global, np.int64)
        L.check(self._lib.dincae_grad_step_slot(self.handle, slot, B, L.ptr(nv)))

    def grad_step(self, xin, xtrue, B=None, n_valid_global=None):
        xin, B = self._batch(xin, B)
        if isinstance(xtrue, np.ndarray):
            xtrue = _f32(xtrue)
        nv = None if n_valid_global is None else np.ascontiguousarray(n_valid_global, np.int64)
        L.check(self._lib.dincae_grad_step(self.handle, L.ptr(xin), L.ptr(xtrue), B, L.ptr(nv), self._loc(xin)))

    def allreduce_grads(self):
        """SUM the flat gradient buffer (+ loss slot) over all ranks of the context's communicator."""
        L.check(self._lib.dincae_allreduce_grads(self.handle))

    def train_step_dp(self, xin, xtrue, lr, n_valid_global, B=None, want_loss=True):
        """One data-parallel mini-batch on this rank's shard: backward with the per-chain gradient buckets all-reduced behind it,
        then the (replica-identical) L2 + ClipGrad + Adam update."""
        xin, B = self._batch(xin, B)
        if isinstance(xtrue, np.ndarray):
            xtrue = _f32(xtrue)
        nv = np.ascontiguousarray(n_valid_global, np.int64)
        out = C.c_float()
        L.check(self._lib.dincae_train_step_dp(self.handle, L.ptr(xin), L.ptr(xtrue), B, L.ptr(nv), lr, self._loc(xin),
                                               C.byref(out) if want_loss else None))
        return out.value if want_loss else None

    def train_step_dp_slot(self, slot, B, lr, n_valid_global, want_loss=True):
        nv = np.ascontiguousarray(n_valid_global, np.int64)
        out = C.c_float()
        L.check(self._lib.dincae_train_step_dp_slot(self.handle, slot, B, L.ptr(nv), lr, C.byref(out) if want_loss else None))
        return out.value if want_loss else None

    def replica_divergence(self):
        """max |w_rank - w_rank0| over all ranks and parameters (collective); 0.0 = bit-identical replicas."""
        out = C.c_float()
        L.check(self._lib.dincae_replica_divergence(self.handle, C.byref(out)))
        return out.value

    def grad_buffer(self):
        p = C.c_void_p(); n = C.c_int64()
        L.check(self._lib.dincae_grad_buffer(self.handle, C.byref(p), C.byref(n)))
        return p.value, n.value

    def grad_tensor(self):
        """Zero-copy torch view of the flat gradient buffer (for torch.distributed.all_reduce)."""
        import torch
        p, n = self.grad_buffer()

        class _View:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (p, False), "version": 2}
        return torch.as_tensor(_View(), device=torch.device("cuda", self.ctx.device))

    def apply_step(self, lr, want_loss=True):
        out = C.c_float()
        L.check(self._lib.dincae_apply_step(self.handle, lr, C.byref(out) if want_loss else None))
        return out.value if want_loss else None

    def count_valid(self, xtrue, B=None):
        if isinstance(xtrue, np.ndarray):
            xtrue = _f32(xtrue); B = xtrue.shape[3]
        out = np.zeros(self.cfg.noutput, np.int64)
        L.check(self._lib.dincae_count_valid(self.handle, L.ptr(xtrue), B, self._loc(xtrue), L.ptr(out)))
        return out

    # -- point-cloud truth: list of (pos (P,2), x (P,2)) per sample (src/points.jl:310)
    @staticmethod
    def _pack_points(point_truth):
        offs = np.zeros(len(point_truth) + 1, np.int64)
        for b, (p, _) in enumerate(point_truth):
            offs[b + 1] = offs[b] + len(p)
        pos = np.concatenate([np.asarray(p, np.float32).reshape(-1, 2) for p, _ in point_truth]) if offs[-1] else np.zeros((0, 2), np.float32)
        x = np.concatenate([np.asarray(xx, np.float32).reshape(-1, 2) for _, xx in point_truth]) if offs[-1] else np.zeros((0, 2), np.float32)
        return np.ascontiguousarray(pos), np.asfortranarray(x), offs

    def loss_points(self, xin, point_truth):
        xin, B = self._batch(xin, None)
        pos, x, offs = self._pack_points(point_truth)
        out = C.c_double()
        L.check(self._lib.dincae_loss_points(self.handle, L.ptr(xin), L.ptr(pos), L.ptr(x), L.ptr(offs), B, L.HOST, C.byref(out)))
        return out.value

    def train_step_points(self, xin, point_truth, lr, want_loss=True):
        """xin: host array, or a device buffer (e.g. filled by PointDataset.assemble(..., xin=buf)); the point truth stays on the host."""
        xin, B = self._batch(xin, len(point_truth))
        pos, x, offs = self._pack_points(point_truth)
        out = C.c_float()
        loc = L.HOST if isinstance(xin, np.ndarray) else L.DEVICE_XIN
        L.check(self._lib.dincae_train_step_points(self.handle, L.ptr(xin), L.ptr(pos), L.ptr(x), L.ptr(offs), B, lr, loc,
                                                   C.byref(out) if want_loss else None))
        return out.value if want_loss else None

    def grad_step_points(self, xin, point_truth, n_valid_global=-1):
        xin, B = self._batch(xin, len(point_truth))
        pos, x, offs = self._pack_points(point_truth)
        loc = L.HOST if isinstance(xin, np.ndarray) else L.DEVICE_XIN
        L.check(self._lib.dincae_grad_step_points(self.handle, L.ptr(xin), L.ptr(pos), L.ptr(x), L.ptr(offs), B, int(n_valid_global), loc))

    def close(self):
        if self.handle:
            self._lib.dincae_model_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- gridded data -------------------------------------------------------------------------------
def time_cos_sin(days_since_1970, cycle_period=365.25):
    """time_cos/time_sin of NCData (src/data.jl:289-297): the reference's `t` is in HOURS."""
    t = np.asarray(days_since_1970, np.float64) * 24.0
    return (np.cos(2 * np.pi * t / cycle_period).astype(np.float32), np.sin(2 * np.pi * t / cycle_period).astype(np.float32))


def scale_pm1(v):
    v = np.asarray(v, np.float64)
    return (2 * (v - v.min()) / (v.max() - v.min()) - 1).astype(np.float32)      # src/data.jl:342-343


class GriddedDataset:
    """NCData (src/data.jl:148-170, ctor :250-369) resident on the device."""

    def __init__(self, ctx: Context, lon, lat, days, data_full, train=False, obs_err_std=None, jitter_std=None,
                 ntime_win=3, isoutput=None, remove_mean=True, mean_data=None, mask=None, cycle_periods=(365.25,)):
        if len(cycle_periods) != 1:
            raise NotImplementedError("exactly one cycle period is supported (4 meta channels, src/data.jl:199)")
        self.ctx, self._lib = ctx, ctx._lib
        if hasattr(data_full, "data_ptr"):
            # a cube that already lives on the device: a contiguous float32 torch tensor of shape (N,V,H,W), i.e. the memory of
            # a Julia (W,H,V,N) array
            if data_full.dim() != 4 or not data_full.is_contiguous() or str(data_full.dtype) != "torch.float32" or not data_full.is_cuda:
                raise ValueError("device cube must be a contiguous float32 CUDA tensor of shape (N,V,H,W)")
            N, V, H, W = data_full.shape
            cube_loc = L.DEVICE
        else:
            data_full = np.asfortranarray(data_full, np.float32)
            if data_full.ndim != 4:
                raise ValueError("data_full must be (W,H,V,N)")
            W, H, V, N = data_full.shape
            cube_loc = L.HOST
        self.shape = (W, H, V, N)
        self.train, self.ntime_win = train, ntime_win
        self.isoutput = (np.arange(V) == 0) if isoutput is None else np.asarray(isoutput, bool)
        self.noutput = int(self.isoutput.sum())
        err_field = None
        if obs_err_std is not None and np.ndim(obs_err_std) == 4:         # per-pixel error field (errvarname, src/data.jl:323-326)
            err_field = np.asfortranarray(obs_err_std, np.float32)
            if err_field.shape != (W, H, V, N):
                raise ValueError("obs_err_std field must be (W,H,V,N)")
            obs_err_std = None
        oes = np.ones(V, np.float32) if obs_err_std is None else np.ascontiguousarray(obs_err_std, np.float32)
        js = np.full(V, 0.05, np.float32) if jitter_std is None else np.ascontiguousarray(jitter_std, np.float32)
        iso = np.ascontiguousarray(self.isoutput, np.uint8)
        tc, ts = time_cos_sin(days, cycle_periods[0])
        lm = None if mask is None else np.asfortranarray(mask, np.uint8)
        md = None if mean_data is None else _f32(mean_data)
        self.lon, self.lat = np.asarray(lon, np.float64), np.asarray(lat, np.float64)
        self.lon_scaled, self.lat_scaled = scale_pm1(lon), scale_pm1(lat)
        h = C.c_void_p()
        L.check(self._lib.dincae_gridded_create(ctx.handle, L.ptr(data_full), W, H, V, N, L.ptr(oes), L.ptr(js), L.ptr(iso),
                                                int(remove_mean), L.ptr(md), L.ptr(lm), L.ptr(self.lon_scaled), L.ptr(self.lat_scaled),
                                                L.ptr(tc), L.ptr(ts), ntime_win, cube_loc, C.byref(h)))
        self.handle = h
        self.mask = mask
        if err_field is not None:
            L.check(self._lib.dincae_gridded_set_error(h, L.ptr(err_field), L.HOST))

    def __len__(self):
        return self.shape[3]

    def sizex(self):
        W, H, V, N = self.shape
        return (W, H, 4 + 2 * V * self.ntime_win)

    def sizey(self):
        W, H, V, N = self.shape
        return (W, H, 2 * self.noutput)

    @property
    def meandata(self):
        W, H, V, N = self.shape
        out = np.zeros((W, H, V), np.float32, order="F")
        L.check(self._lib.dincae_gridded_get_mean(self.handle, L.ptr(out)))
        return out

    def valid_counts(self, v=0):
        out = np.zeros(self.shape[3], np.int64)
        L.check(self._lib.dincae_gridded_valid_counts(self.handle, v, L.ptr(out)))
        return out

    def assemble(self, time_idx, imask=None, seed=0, sample_ctr0=0, xin=None, xtrue=None):
        """One mini-batch (getxy! for every sample of DataBatches.iterate); host arrays unless device buffers are given."""
        time_idx = np.ascontiguousarray(time_idx, np.int64)
        B = len(time_idx)
        im = None if imask is None else np.ascontiguousarray(imask, np.int64)
        if xin is None:
            xin = np.zeros(self.sizex() + (B,), np.float32, order="F")
            xtrue = np.zeros(self.sizey() + (B,), np.float32, order="F")
            loc = L.HOST
        else:
            loc = L.DEVICE
        L.check(self._lib.dincae_gridded_assemble(self.handle, L.ptr(time_idx), L.ptr(im), seed, sample_ctr0, B, L.ptr(xin), L.ptr(xtrue), loc))
        return xin, xtrue

    def close(self):
        if self.handle:
            self._lib.dincae_gridded_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DataBatches:
    """DataBatches (src/data.jl:771-868): mini-batches in `perm` order, partial last batch kept.

    The integer draws (perm, imask) come from a host numpy Generator; the float jitter comes from the
    library's Philox stream keyed by (seed, global sample counter)."""

    def __init__(self, ds: GriddedDataset, batch_size, rng: Optional[np.random.Generator] = None, seed=0,
                 device_buffers=None, rank=0, world_size=1):
        self.ds, self.batch_size = ds, batch_size
        self.rng = rng or np.random.default_rng(seed)
        self.seed = seed
        self.shuffle = ds.train
        n = len(ds)
        self.perm = self.rng.permutation(n) if self.shuffle else np.arange(n)
        self.device_buffers = device_buffers
        self.rank, self.world_size = rank, world_size
        self.sample_ctr = 0

    def __len__(self):
        return -(-len(self.ds) // self.batch_size)        # partial = true (src/data.jl:803)

    def shuffle_(self):                                   # Random.shuffle!(d), src/data.jl:796-799
        if self.shuffle:
            self.perm = self.rng.permutation(len(self.ds))

    def batches(self):
        """Yield (time_idx, imask, sample_ctr0) of every mini-batch: the integer work of one epoch."""
        n = len(self.ds)
        for start in range(0, n, self.batch_size):
            idx = self.perm[start:start + self.batch_size]
            imask = self.rng.integers(0, n, size=len(idx)) if self.ds.train else None      # rand(1:ntime), src/data.jl:471
            ctr = self.sample_ctr
            self.sample_ctr += len(idx)
            yield idx, imask, ctr

    def shard(self, idx, imask):
        """This rank's contiguous share of a mini-batch (data-parallel training)."""
        if self.world_size == 1:
            return idx, imask, 0
        per = -(-len(idx) // self.world_size)
        lo, hi = min(self.rank * per, len(idx)), min((self.rank + 1) * per, len(idx))
        return idx[lo:hi], (None if imask is None else imask[lo:hi]), lo

    def __iter__(self):
        for idx, imask, ctr in self.batches():
            if self.device_buffers is None:
                yield self.ds.assemble(idx, imask, self.seed, ctr) + (len(idx),)
            else:
                xin, xtrue = self.device_buffers
                self.ds.assemble(idx, imask, self.seed, ctr, xin, xtrue)
                yield xin, xtrue, len(idx)


def learning_rate(lr0, epoch, decay_epoch):
    """lr of epoch e (1-based), src/model.jl:590."""
    return lr0 * 0.5 ** (epoch / decay_epoch)


def train_epoch(model: Model, batches, lr):
    """train_epoch! (src/flux.jl:24-40): mean of the per-batch losses, partial batch included (T17)."""
    loss_sum, n = 0.0, 0
    if isinstance(batches, DataBatches) and batches.device_buffers is not None and batches.world_size == 1 and batches.ds.ctx is model.ctx:
        # device-resident dataset: assembly + step of every mini-batch in one library call (one CUDA graph per step);
        # bit-identical to assembling into the device buffers and calling train_step on them
        for idx, imask, ctr in batches.batches():
            loss_sum += model.train_step_gridded(batches.ds, idx, imask, lr, seed=batches.seed, sample_ctr0=ctr)
            n += 1
        return loss_sum / n
    for xin, xtrue, B in batches:
        loss_sum += model.train_step(xin, xtrue, lr, B=B)
        n += 1
    return loss_sum / n


def train_epoch_prefetch(model: Model, host_batches, lr):
    """train_epoch!(MO, PrefetchDataIter(train), lr) (src/model.jl:592): host mini-batches, the next one is uploaded while
    the current one trains."""
    it = iter(host_batches)
    cur = next(it, None)
    if cur is None:
        raise ValueError("empty data iterator")
    keep = model.upload(0, cur[0], cur[1])
    slot, loss_sum, n = 0, 0.0, 0
    while cur is not None:
        nxt = next(it, None)
        if nxt is not None:
            keep_next = model.upload(slot ^ 1, nxt[0], nxt[1])
        loss_sum += model.train_step_slot(slot, cur[0].shape[3], lr)
        n += 1
        cur, slot = nxt, slot ^ 1
        if nxt is not None:
            keep = keep_next
    return loss_sum / n


def reconstruct_range(model: Model, ds: GriddedDataset, t0, t1, count=1, fill_value=-9999.0, m_out=None, sigma_out=None):
    """Save loop of reconstruct (src/model.jl:624-644) + savesample arithmetic (src/data.jl:615-675) for steps [t0,t1)."""
    W, H, V, N = ds.shape
    if m_out is None:
        m_out = np.zeros((W, H, t1 - t0, ds.noutput), np.float32, order="F")
        sigma_out = np.zeros_like(m_out, order="F")
    loc = L.HOST if isinstance(m_out, np.ndarray) else L.DEVICE
    L.check(model._lib.dincae_reconstruct_range(model.handle, ds.handle, int(t0), int(t1), int(count), float(fill_value),
                                                L.ptr(m_out), L.ptr(sigma_out), loc))
    return m_out, sigma_out


def _device_batch_buffers(ds, batch_size):
    import torch
    dev = torch.device("cuda", ds.ctx.device)
    W, H, C_ = ds.sizex()
    xin = torch.empty(batch_size * C_ * H * W, dtype=torch.float32, device=dev)
    xtrue = torch.empty(batch_size * 2 * ds.noutput * H * W, dtype=torch.float32, device=dev)
    return xin, xtrue


def _load_entry(d):
    """One variable of data_all: either in-memory arrays or the reference's (filename, varname, errvarname) form."""
    if "data" in d:
        return d
    from . import ncio
    lon, lat, time, data, err, _, mask = ncio.load_gridded_nc(d["filename"], d["varname"], d.get("errvarname"),
                                                              obs_err_std=d.get("obs_err_std", 1.0))
    out = dict(d)
    out.update(lon=lon, lat=lat, time=time, data=data, mask=mask)
    if d.get("errvarname") is not None:
        out["err"] = err                       # per-pixel error standard deviation (W,H,N), Inf where missing (src/data.jl:66-67)
    return out


def reconstruct(ctx, data_all, fnames_rec=None, *,
                epochs=1000, batch_size=50, truth_uncertain=False, enc_nfilter_internal=(16, 24, 36, 54),
                skipconnections=None, clip_grad=5.0, regularization_L1_beta=0, regularization_L2_beta=0,
                save_epochs=None, is3D=False, upsampling_method="nearest", ntime_win=3, learning_rate=0.001,
                learning_rate_decay_epoch=math.inf, min_std_err=0.006737946999085467, loss_weights_refine=(1.0,),
                cycle_periods=(365.25,), output_ndims=1, direction_obs=None, remove_mean=True, paramfile=None,
                laplacian_penalty=0, laplacian_error_penalty=None, modeldir=None,
                seed=0, precision="tf32", init_params=None, progress=None):
    """DINCAE.reconstruct (src/model.jl:446-679) with its keyword surface kept verbatim.

    `ctx` takes the place of `Atype`.  File I/O is out of scope: every entry of `data_all`
    is a list of dicts with in-memory fields {lon, lat, time (days since 1970), data (W,H,N) with NaN,
    mask (W,H) optional, obs_err_std, jitter_std, isoutput}; data_all[0] trains, data_all[1:] are
    reconstructed at save_epochs.  Returns (losses, results) where results[k] = dict(mean, sigma, count)
    holds what savesample would have written to fnames_rec[k]; if fnames_rec is given they are saved as .npz.
    """
    if save_epochs is None:
        save_epochs = range(min(epochs, 200), epochs + 1, 10)
    save_epochs = list(save_epochs)
    if len(save_epochs) == 0 or epochs < min(save_epochs):
        raise ValueError("No output will be saved. Consider to adjust save_epochs (currently %s) or epochs (currently %d)."
                         % (save_epochs, epochs))                     # src/model.jl:475-477
    if is3D or output_ndims != 1 or direction_obs is not None:
        raise L.DincaeError(L.EUNSUPPORTED, "is3D=true / output_ndims=2 / direction_obs are outside the hot path (SURVEY section 8f N4)")

    def make(dlist, train, mean_data=None, mask=None):
        dlist = [_load_entry(d) for d in dlist]                                          # {filename, varname} -> arrays (src/data.jl:88-110)
        d0 = dlist[0]
        data = np.stack([np.asarray(d["data"], np.float32) for d in dlist], axis=2)      # (W,H,V,N)
        oes = [d.get("obs_err_std", 1.0) for d in dlist]
        if any("err" in d for d in dlist):       # error_full of load_gridded_nc(data) (src/data.jl:88-110): a field for every variable
            oes = np.stack([np.asarray(d["err"], np.float32) if "err" in d else np.full(data.shape[:2] + data.shape[3:], d.get("obs_err_std", 1.0), np.float32)
                            for d in dlist], axis=2)
        return GriddedDataset(ctx, d0["lon"], d0["lat"], d0["time"], data, train=train,
                              obs_err_std=oes,
                              jitter_std=[d.get("jitter_std", 0.05) for d in dlist],
                              ntime_win=ntime_win, isoutput=[d.get("isoutput", i == 0) for i, d in enumerate(dlist)],
                              remove_mean=remove_mean, mean_data=mean_data, mask=d0.get("mask") if mask is None else mask,
                              cycle_periods=cycle_periods)

    train_data = make(data_all[0], True)
    train_mean = train_data.meandata
    # use common mean (src/model.jl:503-504); savesample masks with the TRAINING data's mask (mask = train_data.mask, :641)
    tests = [make(d, False, train_mean, mask=train_data.mask) for d in data_all[1:]]
    W, H, Cin = train_data.sizex()
    cfg = ModelConfig(W=W, H=H, ninput=Cin, noutput=train_data.noutput, enc_nfilter_internal=tuple(enc_nfilter_internal),
                      skipconnections=skipconnections, upsampling_method=str(upsampling_method).lstrip(":"),
                      loss_weights_refine=tuple(loss_weights_refine), min_std_err=min_std_err, truth_uncertain=truth_uncertain,
                      regularization_L1_beta=regularization_L1_beta, regularization_L2_beta=regularization_L2_beta,
                      laplacian_penalty=laplacian_penalty, laplacian_error_penalty=laplacian_error_penalty,
                      clip_grad=clip_grad, precision=precision, max_batch=batch_size)
    model = Model(ctx, cfg, seed=seed)
    if init_params is not None:
        model.set_params(init_params)
    bufs = _device_batch_buffers(train_data, batch_size)
    train = DataBatches(train_data, batch_size, seed=seed, device_buffers=bufs)
    losses = []
    results = [dict(mean=None, sigma=None, count=0) for _ in tests]
    for e in range(1, epochs + 1):
        lr = learning_rate * 0.5 ** (e / learning_rate_decay_epoch)    # src/model.jl:590
        losses.append(train_epoch(model, train, lr))
        if progress:
            progress("epoch: %5d loss %5.4f" % (e, losses[-1]))        # src/model.jl:596
        if e in save_epochs:
            if modeldir is not None:                                   # model-checkpoint-%05d at every save epoch, src/model.jl:599-620
                os.makedirs(modeldir, exist_ok=True)
                model.save_checkpoint(os.path.join(modeldir, "model-checkpoint-%05d" % e), e=e, ntime_win=ntime_win, is3D=int(is3D),
                                      cycle_periods=np.asarray(cycle_periods, np.float64), remove_mean=int(remove_mean),
                                      train_mean_data=np.asarray(train_mean), noutput=train_data.noutput,
                                      input_size=np.asarray([W, H, Cin]), output_ndims=output_ndims)
            for ds_, res in zip(tests, results):
                res["count"] += 1                                      # ds.attrib["count"] += 1, src/data.jl:621-623
                res["mean"], res["sigma"] = reconstruct_range(model, ds_, 0, len(ds_), res["count"], -9999.0, res["mean"], res["sigma"])
        train.shuffle_()                                               # src/model.jl:650
    if fnames_rec is not None:
        for fn, res, dl, ds_ in zip(fnames_rec, results, data_all[1:], tests):
            if str(fn).endswith(".nc"):                                # the reference's output file (src/data.jl:712-769)
                from . import ncio
                names = [d.get("varname", "var%d" % (i + 1)) for i, d in enumerate(dl) if d.get("isoutput", i == 0)]
                nv = res["mean"].shape[3]                               # results are (W, H, N, noutput)
                ncio.write_reconstruction(fn, names[:nv], ds_.lon, ds_.lat, np.asarray(train_mean)[:, :, np.flatnonzero(train_data.isoutput)[:nv]],
                                          [res["mean"][:, :, :, i] for i in range(nv)], [res["sigma"][:, :, :, i] for i in range(nv)], res["count"])
            else:
                np.savez(fn, **{k: v for k, v in res.items() if v is not None})
    if paramfile is not None:                                          # src/model.jl:655-676, attribute for attribute
        from . import ncio
        L_ = len(tuple(enc_nfilter_internal))
        ncio.write_paramfile(paramfile, losses, dict(
            epochs=epochs, batch_size=batch_size, truth_uncertain=bool(truth_uncertain), enc_nfilter_internal=list(enc_nfilter_internal),
            skipconnections=list(range(2, L_ + 2) if skipconnections is None else skipconnections), clip_grad=float(clip_grad or 0.0),
            regularization_L1_beta=float(regularization_L1_beta), regularization_L2_beta=float(regularization_L2_beta),
            save_epochs=list(save_epochs), is3D=bool(is3D), upsampling_method=str(upsampling_method).lstrip(":"), ntime_win=ntime_win,
            learning_rate=float(learning_rate), learning_rate_decay_epoch=float(learning_rate_decay_epoch), min_std_err=float(min_std_err),
            loss_weights_refine=[float(x) for x in loss_weights_refine], cycle_periods=[float(x) for x in cycle_periods]))
    reconstruct.last_model = model
    return losses, results


# ---- point clouds ---------------------------------------------------------------------------------
class PointDataset:
    """PointCloud (src/points.jl:137-227) on the device; construction filters with posindices!."""

    def __init__(self, ctx: Context, grid, lon, lat, values, ids, dayofyear, ntime_win=9, probability_skip_for_training=0.2,
                 jitter_std_pos=(5.0, 5.0), auxdata=None, train=True):
        self.ctx, self._lib = ctx, ctx._lib
        self.grid = grid
        self.sz = (len(grid[0]), len(grid[1]))
        N = len(values)
        self.pos, self.x, self.id = [], [], []
        for n in range(N):
            p, inside = posindices(grid, lon[n], lat[n], ctx)
            v = np.asarray(values[n], np.float32)[inside]
            xx = np.zeros((len(v), 2), np.float32)
            xx[:, 0] = np.where(np.isnan(v), 0, v)              # src/points.jl:188
            xx[:, 1] = 1 - np.isnan(v).astype(np.float32)       # src/points.jl:189
            self.pos.append(p[inside]); self.x.append(xx); self.id.append(np.asarray(ids[n], np.int64)[inside])
        self.offsets = np.zeros(N + 1, np.int64)
        self.offsets[1:] = np.cumsum([len(p) for p in self.pos])
        pos_all = np.ascontiguousarray(np.concatenate(self.pos)) if self.offsets[-1] else np.zeros((0, 2), np.float32)
        x_all = np.asfortranarray(np.concatenate(self.x)) if self.offsets[-1] else np.zeros((0, 2), np.float32, order="F")
        self.ntime_win, self.train, self.pskip = ntime_win, train, probability_skip_for_training
        d = np.asarray(dayofyear, np.float64)
        dcos = np.cos(2 * np.pi * d / 365.25).astype(np.float32)          # src/points.jl:199-202
        dsin = np.sin(2 * np.pi * d / 365.25).astype(np.float32)
        W, H = self.sz
        self.naux = 0 if auxdata is None else auxdata.shape[2]
        aux = None if auxdata is None else _f32(auxdata)
        lon_s, lat_s = scale_pm1(grid[0]), scale_pm1(grid[1])             # src/points.jl:195-196
        h = C.c_void_p()
        L.check(self._lib.dincae_points_create(ctx.handle, W, H, N, L.ptr(pos_all), L.ptr(x_all), L.ptr(self.offsets), L.ptr(aux), self.naux,
                                               L.ptr(lon_s), L.ptr(lat_s), L.ptr(dcos), L.ptr(dsin), ntime_win,
                                               float(jitter_std_pos[0]), float(jitter_std_pos[1]), C.byref(h)))
        self.handle = h

    def __len__(self):
        return len(self.pos)

    def sizex(self):
        return (self.sz[0], self.sz[1], 4 + 2 * self.ntime_win + self.naux)       # src/points.jl:345-352

    def draw_track_keep(self, ind, rng):
        """Track dropout of the centre step (src/points.jl:284-300): integer work drawn on the host."""
        ids = self.id[ind]
        sel = np.ones(len(ids), bool)
        uniq = list(dict.fromkeys(ids.tolist()))                                   # unique(d.id[ind]) keeps first-seen order
        for uid in uniq:
            if rng.random() < self.pskip:
                sel[ids == uid] = False
        if sel.sum() == 0 and len(uniq) > 0:
            sel[ids == uniq[rng.integers(0, len(uniq))]] = True                    # at least one track
        return sel

    def assemble(self, time_idx, seed=0, sample_ctr0=0, keep=None, perturb=None, xin=None):
        time_idx = np.ascontiguousarray(time_idx, np.int64)
        B = len(time_idx)
        perturb = self.train if perturb is None else perturb
        kp = ko = None
        if keep is not None:
            ko = np.zeros(B + 1, np.int64)
            ko[1:] = np.cumsum([len(k) for k in keep])
            kp = np.ascontiguousarray(np.concatenate(keep) if ko[-1] else np.zeros(0), np.uint8)
        loc = L.HOST if xin is None else L.DEVICE                                # device buffer: the batch never visits the host
        if xin is None:
            xin = np.zeros(self.sizex() + (B,), np.float32, order="F")
        L.check(self._lib.dincae_points_assemble(self.handle, L.ptr(time_idx), B, int(perturb), seed, sample_ctr0,
                                                 L.ptr(kp), L.ptr(ko), L.ptr(xin), loc))
        truth = [(self.pos[i], self.x[i]) for i in time_idx]                       # xtrue = (pos[ind], x[ind]), src/points.jl:310
        return xin, truth

    def close(self):
        if self.handle:
            self._lib.dincae_points_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def reconstruct_points(ctx, data, grid, fnames_rec=None, *,
                       epochs=60, batch_size=32, truth_uncertain=False, enc_nfilter_internal=(10, 20, 30, 40, 50),
                       skipconnections=None, clip_grad=5.0, regularization_L1_beta=0, regularization_L2_beta=0,
                       save_epochs=None, upsampling_method="nearest", probability_skip_for_training=0.2,
                       jitter_std_pos=(5.0, 5.0), ntime_win=9, learning_rate=0.001, learning_rate_decay_epoch=math.inf,
                       min_std_err=0.006737946999085467, loss_weights_refine=(1.0,), auxdata_files=(), paramfile=None,
                       savesnapshot=False, laplacian_penalty=0, laplacian_error_penalty=None,
                       varname=None, auxdata=None, seed=0, precision="tf32", init_params=None, progress=None):
    """DINCAE.reconstruct_points (src/points.jl:509-696), keyword surface verbatim.

    `data` is the reference's `filename` (a CF contiguous-ragged NetCDF-3 file read by ncio.load_ragged_points, with
    `varname`), or the same content in memory: dict(lon, lat, values, id: lists of per-time-step vectors; dayofyear: (N,)).
    `auxdata_files` = [dict(filename, varname, errvarname)] as in the reference (src/points.jl:543, src/data.jl:124-146);
    `auxdata` (W,H,2*naux,N) passes the already normalised array instead."""
    if save_epochs is None:
        save_epochs = range(min(epochs, 60), epochs + 1, 10)
    save_epochs = list(save_epochs)
    if len(save_epochs) == 0 or epochs < min(save_epochs):
        raise ValueError("No output will be saved. Consider to adjust save_epochs (currently %s) or epochs (currently %d)."
                         % (save_epochs, epochs))                     # src/points.jl:535-537
    if isinstance(data, (str, bytes, os.PathLike)):
        if varname is None:
            raise ValueError("reconstruct_points: varname is required with a file name")
        from . import ncio
        values, (plon, plat), _dtime, pid, _dates, doy = ncio.load_ragged_points(os.fspath(data), varname)     # DINCAE.loaddata, src/points.jl:541
        data = dict(lon=plon, lat=plat, values=values, id=pid, dayofyear=doy)
    if len(auxdata_files) > 0:
        if auxdata is not None:
            raise ValueError("pass either auxdata_files or auxdata")
        from . import ncio
        auxdata = ncio.load_aux_data((len(grid[0]), len(grid[1]), len(data["values"])), list(auxdata_files))    # src/points.jl:545
    kw = dict(ntime_win=ntime_win, probability_skip_for_training=probability_skip_for_training,
              jitter_std_pos=jitter_std_pos, auxdata=auxdata)
    train = PointDataset(ctx, grid, data["lon"], data["lat"], data["values"], data["id"], data["dayofyear"], train=True, **kw)
    test = PointDataset(ctx, grid, data["lon"], data["lat"], data["values"], data["id"], data["dayofyear"], train=False, **kw)
    W, H, nvar = train.sizex()
    cfg = ModelConfig(W=W, H=H, ninput=nvar, noutput=1, enc_nfilter_internal=tuple(enc_nfilter_internal),
                      skipconnections=skipconnections, upsampling_method=str(upsampling_method).lstrip(":"),
                      loss_weights_refine=tuple(loss_weights_refine), min_std_err=min_std_err, truth_uncertain=truth_uncertain,
                      regularization_L1_beta=regularization_L1_beta, regularization_L2_beta=regularization_L2_beta,
                      laplacian_penalty=laplacian_penalty, laplacian_error_penalty=laplacian_error_penalty,
                      clip_grad=clip_grad, precision=precision, max_batch=batch_size)
    model = Model(ctx, cfg, seed=seed)
    if init_params is not None:
        model.set_params(init_params)
    rng = np.random.default_rng(seed)
    N = len(train)
    perm = rng.permutation(N)                                          # randperm(ntime), src/points.jl:204-209 (never reshuffled)
    losses = []
    result = dict(mean=None, sigma=None, count=0)
    ctr = 0
    import torch                                                       # device memory for the gridded mini-batch: it never visits the host
    xbuf = torch.empty(int(np.prod(train.sizex())) * batch_size, dtype=torch.float32, device="cuda:%d" % ctx.device)
    for e in range(1, epochs + 1):
        lr = learning_rate * 0.5 ** (e / learning_rate_decay_epoch)    # src/points.jl:622
        s, nb = 0.0, 0
        for start in range(0, N, batch_size):
            idx = perm[start:start + batch_size]
            keep = [train.draw_track_keep(int(i), rng) for i in idx]
            xin, truth = train.assemble(idx, seed, ctr, keep, xin=xbuf)
            ctr += len(idx)
            s += model.train_step_points(xin, truth, lr)
            nb += 1
        losses.append(s / nb)
        if progress:
            progress("epoch: %5d loss %5.4f" % (e, losses[-1]))
        if e in save_epochs:
            result["count"] += 1
            if result["mean"] is None:
                result["mean"] = np.zeros((W, H, N), np.float32, order="F"); result["sigma"] = np.zeros((W, H, N), np.float32, order="F")
            c = result["count"]
            for start in range(0, N, batch_size):
                idx = np.arange(start, min(start + batch_size, N))
                xin, _ = test.assemble(idx, perturb=False)
                xrec = model(xin)
                with np.errstate(invalid="ignore"):
                    m, sg = xrec[:, :, 0, :], np.sqrt(xrec[:, :, 1, :])          # meandata = zeros, src/points.jl:613
                for name, new in (("mean", m), ("sigma", sg)):                     # accumulate!, src/data.jl:615-619
                    old = result[name][:, :, idx]
                    v = (np.float32(c - 1) * old + new) / np.float32(c)
                    result[name][:, :, idx] = np.where(np.isnan(v), np.float32(-9999.0), v)
    if fnames_rec is not None:
        np.savez(fnames_rec[0], **{k: v for k, v in result.items() if v is not None})
    if paramfile is not None:                                          # src/points.jl:658-692, attribute for attribute
        from . import ncio
        L_ = len(tuple(enc_nfilter_internal))
        ncio.write_paramfile(paramfile, losses, dict(
            epochs=epochs, batch_size=batch_size, truth_uncertain=bool(truth_uncertain), enc_nfilter_internal=list(enc_nfilter_internal),
            skipconnections=list(range(2, L_ + 2) if skipconnections is None else skipconnections), clip_grad=float(clip_grad or 0.0),
            regularization_L1_beta=float(regularization_L1_beta), regularization_L2_beta=float(regularization_L2_beta),
            save_epochs=list(save_epochs), upsampling_method=str(upsampling_method).lstrip(":"),
            probability_skip_for_training=float(probability_skip_for_training), jitter_std_pos=[float(x) for x in jitter_std_pos],
            ntime_win=ntime_win, learning_rate=float(learning_rate), learning_rate_decay_epoch=float(learning_rate_decay_epoch),
            min_std_err=float(min_std_err), loss_weights_refine=[float(x) for x in loss_weights_refine],
            auxdata_filenames=[f["filename"] for f in auxdata_files] if len(auxdata_files) else "",
            auxdata_varnames=[f["varname"] for f in auxdata_files] if len(auxdata_files) else "",
            auxdata_errvarname=[str(f.get("errvarname")) for f in auxdata_files] if len(auxdata_files) else "",
            savesnapshot=bool(savesnapshot), laplacian_penalty=float(laplacian_penalty),
            laplacian_error_penalty=float(laplacian_penalty if laplacian_error_penalty is None else laplacian_error_penalty)))
    reconstruct_points.last_model = model
    return losses, result
