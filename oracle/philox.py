"""Philox4x32-10 counter RNG and the dropout-mask stream of the CUDA kernels, in numpy.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  The reference draws dropout masks
from torch's generator (``baselineModel.py:21-22,81``: ``torch.nn.Dropout``); the CUDA
path draws them from Philox counters so that backward can address them again and so
that a data-parallel run sees the masks of the single-process global batch.  This file
is the bit-exact CPU twin of ``csrc/philox.cuh``; ``oracle/make_golden.py`` feeds these
masks into the *reference* module (dropout patched) to pin training parity.

Stream definition.  One call yields 128 bits, cut into lanes of k bits, one lane per keep
decision; k is the smallest of {1, 8, 16} that represents p exactly (p == 0.5: k = 1;
p a multiple of 1/256: k = 8; else k = 16, p rounded to 1/65536).  A call covers
R = 16 / k consecutive rows of 8 consecutive columns:
    counter = (row // R, col // 8, site, step)      key = (seed_lo, seed_hi)
    r0..r3  = Philox4x32-10(counter, key)           bits = r0 | r1 << 32 | r2 << 64 | r3 << 96
    lane    = (row % R) * 8 + col % 8               value = (bits >> (lane * k)) & (2^k - 1)
    keep    = value >= floor(float32(p) * 2^k + 0.5)
"""
import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  Inputs broadcastable uint32 arrays; returns 4 uint32 arrays."""
    c0, c1, c2, c3 = [np.asarray(c, dtype=np.uint64) & _MASK32 for c in (c0, c1, c2, c3)]
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def dropout_lanes(p):
    """(k, R, threshold) for dropout probability p: lane width in bits, rows per call, keep <=> value >= threshold
    (same expressions as ``dropout_lanes`` in csrc/philox.cuh)."""
    p = float(np.float32(p))
    if p == 0.5:
        k = 1
    elif p * 256.0 == np.floor(p * 256.0):
        k = 8
    else:
        k = 16
    t = int(min(max(int(np.floor(p * float(1 << k) + 0.5)), 0), 1 << k))
    return k, 16 // k, t


def dropout_keep_mask(rows, cols, site, step, seed, p, row_offset=0):
    """Boolean keep-mask (rows, cols) of dropout site ``site`` at optimiser step ``step``.

    ``row_offset`` is the global index of row 0 (data-parallel shards pass their offset).
    """
    assert cols % 8 == 0
    k, R, thr = dropout_lanes(p)
    grow = np.arange(rows, dtype=np.uint64) + np.uint64(row_offset)
    r = (grow // np.uint64(R))[:, None]
    g = np.arange(cols // 8, dtype=np.uint64)[None, :]
    out = philox4x32_10(r, g, np.uint64(site), np.uint64(step & 0xFFFFFFFF),
                        seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    words = np.stack([o.astype(np.uint64) for o in out], axis=-1)          # (rows, cols/8, 4)
    keep = np.empty((rows, cols // 8, 8), dtype=bool)
    sub = (grow % np.uint64(R)).astype(np.int64)[:, None]                   # row inside the call
    for e in range(8):
        bit = (sub * 8 + e) * k                                            # first bit of the lane, (rows, 1)
        w = np.take_along_axis(words, np.broadcast_to((bit // 32)[..., None], (rows, cols // 8, 1)), axis=-1)[..., 0]
        val = (w >> (bit % 32).astype(np.uint64)) & np.uint64((1 << k) - 1)
        keep[:, :, e] = val >= np.uint64(thr)
    return keep.reshape(rows, cols)
