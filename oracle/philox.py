"""Philox4x32-10 counter-based RNG + Box-Muller normals (numpy restatement).

The reference draws `rand/randn` from Julia's task-local Xoshiro streams whose
order depends on task scheduling (src/data.jl:471,494,842,892), so its random
sequence cannot be reproduced bit for bit.  The boundary
therefore takes the integer draws (perm, imask, track keep flags) from the
caller and generates the float jitter from this documented counter-based
generator, which the CUDA kernels and this oracle both implement.
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  All inputs broadcastable uint32 arrays."""
    c0, c1, c2, c3 = [np.asarray(c, dtype=np.uint64) & MASK for c in np.broadcast_arrays(c0, c1, c2, c3)]
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return [c.astype(np.uint32) for c in (c0, c1, c2, c3)]


def u32_to_unit(x):
    """uint32 -> float32 in (0,1): ((x >> 8) + 0.5) * 2^-24."""
    return ((x >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -24)


def normals4(c0, c1, c2, c3, seed):
    """Four standard normals (float32) per counter: Box-Muller on (x0,x1), (x2,x3)."""
    k0 = seed & 0xFFFFFFFF
    k1 = (seed >> 32) & 0xFFFFFFFF
    x0, x1, x2, x3 = philox4x32(c0, c1, c2, c3, k0, k1)
    out = []
    for a, b in ((x0, x1), (x2, x3)):
        u1 = u32_to_unit(a)
        u2 = u32_to_unit(b)
        r = np.sqrt(np.float32(-2.0) * np.log(u1)).astype(np.float32)
        th = (np.float32(6.283185307179586) * u2).astype(np.float32)
        out.append((r * np.cos(th)).astype(np.float32))
        out.append((r * np.sin(th)).astype(np.float32))
    return out


def normal_field(n, c1, c2, c3, seed):
    """n normals indexed e=0..n-1: element e uses counter (e>>2, c1, c2, c3), lane e&3."""
    e = np.arange(n, dtype=np.uint64)
    z = normals4((e >> np.uint64(2)), c1, c2, c3, seed)
    lane = (e & np.uint64(3)).astype(np.int64)
    out = np.choose(lane, z).astype(np.float32)
    return out
