"""Philox4x32-10 counter RNG and the dropout-mask stream of the CUDA kernels, in numpy.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  The reference draws dropout masks
from torch's generator (``baselineModel.py:21-22,81``: ``torch.nn.Dropout``); the CUDA
path draws them from Philox counters so that backward can address them again and so
that a data-parallel run sees the masks of the single-process global batch.  This file
is the bit-exact CPU twin of ``csrc/philox.cuh``; ``oracle/make_golden.py`` feeds these
masks into the *reference* module (dropout patched) to pin training parity.

Stream definition (one call yields eight 16-bit lanes = eight consecutive columns):
    counter = (row, col // 8, site, step)      key = (seed_lo, seed_hi)
    r0..r3  = Philox4x32-10(counter, key)
    u16[e]  = (r[e // 2] >> (16 * (e % 2))) & 0xFFFF           e = col % 8
    keep    = u16 >= floor(float32(p) * 65536 + 0.5)
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


def dropout_threshold(p):
    """round-half-up of float32(p) * 65536, clipped to [0, 65536] (same expression as csrc/philox.cuh)."""
    t = float(np.float32(p)) * 65536.0
    return int(min(max(int(np.floor(t + 0.5)), 0), 65536))


def dropout_keep_mask(rows, cols, site, step, seed, p, row_offset=0):
    """Boolean keep-mask (rows, cols) of dropout site ``site`` at optimiser step ``step``.

    ``row_offset`` is the global index of row 0 (data-parallel shards pass their offset).
    """
    assert cols % 8 == 0
    r = (np.arange(rows, dtype=np.uint64) + np.uint64(row_offset))[:, None]
    g = np.arange(cols // 8, dtype=np.uint64)[None, :]
    out = philox4x32_10(r, g, np.uint64(site), np.uint64(step & 0xFFFFFFFF),
                        seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    u16 = np.empty((rows, cols // 8, 8), dtype=np.uint32)
    for e in range(8):
        u16[:, :, e] = (out[e // 2] >> np.uint32(16 * (e % 2))) & np.uint32(0xFFFF)
    return (u16.reshape(rows, cols) >= np.uint32(dropout_threshold(p)))
