"""ctypes binding of oracle/c/liboracle.so (plain-C restatement; test infrastructure)."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "c")
_LIB = None


def build():
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_costfun.restype = C.c_double
    return _LIB


def _f(a):
    return np.asfortranarray(a, dtype=np.float32)


def interpnd(pos, A):
    pos = np.ascontiguousarray(pos, np.float32); A = _f(A)
    out = np.zeros(len(pos), np.float32)
    lib().orc_interpnd(pos.ctypes, C.c_int64(len(pos)), A.ctypes, A.shape[0], A.shape[1], out.ctypes)
    return out


def interp_adjn(pos, values, sz):
    pos = np.ascontiguousarray(pos, np.float32); values = np.ascontiguousarray(values, np.float32)
    out = np.zeros(sz, np.float32, order="F")
    lib().orc_interp_adjn(pos.ctypes, values.ctypes, C.c_int64(len(pos)), sz[0], sz[1], out.ctypes)
    return out


def conv3x3(x, Wt, bias, act):
    x = _f(x); Wt = _f(Wt); bias = np.ascontiguousarray(bias, np.float32)
    W, H, Cin, B = x.shape
    Cout = Wt.shape[3]
    out = np.zeros((W, H, Cout, B), np.float32, order="F")
    lib().orc_conv3x3(x.ctypes, W, H, Cin, B, Wt.ctypes, bias.ctypes, Cout, int(act), out.ctypes)
    return out


def meanpool2(x):
    x = _f(x); W, H, Cc, B = x.shape
    out = np.zeros((W // 2 + W % 2, H // 2 + H % 2, Cc, B), np.float32, order="F")
    lib().orc_meanpool2(x.ctypes, W, H, Cc, B, out.ctypes)
    return out


def upsample2_crop(x, method, Wo, Ho):
    x = _f(x); W, H, Cc, B = x.shape
    out = np.zeros((Wo, Ho, Cc, B), np.float32, order="F")
    lib().orc_upsample2_crop(x.ctypes, W, H, Cc, B, int(method), Wo, Ho, out.ctypes)
    return out


def head(x, gamma):
    x = _f(x); W, H, C2, B = x.shape
    out = np.zeros_like(x, order="F")
    lib().orc_head(x.ctypes, C.c_int64(W * H), C2 // 2, B, C.c_float(gamma), out.ctypes)
    return out


def costfun(xrec, xtrue, truth_uncertain=False):
    xrec = _f(xrec); xtrue = _f(xtrue); W, H, C2, B = xrec.shape
    return lib().orc_costfun(xrec.ctypes, xtrue.ctypes, C.c_int64(W * H), C2 // 2, B, int(truth_uncertain))


def adam(w, g, m, v, lr, clip, b1, b2, eps, bt1, bt2):
    for a in (w, g, m, v):
        assert a.dtype == np.float32 and a.flags.c_contiguous
    lib().orc_adam(w.ctypes, g.ctypes, m.ctypes, v.ctypes, C.c_int64(w.size), C.c_float(lr), C.c_float(clip),
                   C.c_float(b1), C.c_float(b2), C.c_float(eps), C.c_float(bt1), C.c_float(bt2))


def chain_forward(cfg, flat, xin):
    """Whole StepModel forward by the plain-C loops (independent of torch)."""
    from . import model as M
    L = len(cfg.nfilter)
    shapes = M.param_shapes(cfg)
    arrs, pos = [], 0
    for s in shapes:
        n = int(np.prod(s))
        arrs.append(np.asarray(flat[pos:pos + n], np.float32).reshape(s, order="F"))
        pos += n

    def level(x, P, l):
        if l == L:
            return x
        W, H = x.shape[:2]
        y = conv3x3(x, P["enc"][l][0], P["enc"][l][1], 1)
        y = meanpool2(y)
        y = level(y, P, l + 1)
        y = upsample2_crop(y, 0 if cfg.upsampling_method == "nearest" else 1, W, H)
        y = conv3x3(y, P["dec"][l][0], P["dec"][l][1], 1 if l > 0 else 0)
        if (l + 1) in cfg.skipconnections:
            y = y + x
        return y
    k = 0
    xout = None
    for ch in range(cfg.nchains):
        enc = [(arrs[k + 2 * i], arrs[k + 2 * i + 1]) for i in range(L)]; k += 2 * L
        dec = [(arrs[k + 2 * i], arrs[k + 2 * i + 1]) for i in range(L)][::-1]; k += 2 * L
        inp = xin if ch == 0 else np.concatenate([xout, xin], axis=2)
        xout = level(_f(inp), {"enc": enc, "dec": dec}, 0)
    return head(xout, cfg.gamma)
