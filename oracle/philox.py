"""Philox4x32-10 restated in numpy  --  TEST INFRASTRUCTURE (see oracle/ddsp_oracle.py header).

The reference draws its noise with ``torch.rand`` on the CPU generator
(model/ddsp/filtered_noise.py:44-48), which no GPU kernel can reproduce bit for bit.
The B200 noise kernel instead generates uniform noise in-kernel with the published
Philox4x32-10 counter RNG (Salmon et al., SC'11; same round constants as cuRAND / torch
CUDA).  This module is the independent CPU statement of that generator and of the
counter layout the kernel documents in ``include/ddsp_b200.h``:

    key      = (seed & 0xffffffff, seed >> 32)
    counter  = (j >> 2, t, b_global, offset)          # j: sample in frame, t: frame, b: GLOBAL voice
    word     = out[j & 3]
    uniform  = (word >> 8) * 2**-24  in [0,1)   ->   noise = 2*uniform - 1  in [-1,1)

so a given (seed, offset, voice, frame, sample) yields the same bits in forward and
backward and for any batch sharding across GPUs.
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32 with 10 rounds. Counters are uint32 arrays (broadcastable); returns 4 uint32 arrays."""
    c0, c1, c2, c3 = np.broadcast_arrays(*(np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3)))
    c0, c1, c2, c3 = c0.copy(), c1.copy(), c2.copy(), c3.copy()
    k0 &= 0xFFFFFFFF
    k1 &= 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)) & _MASK, lo1, (hi0 ^ c3 ^ np.uint64(k1)) & _MASK, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def noise_uniform(B: int, T: int, hop: int, seed: int, offset: int, voice_offset: int = 0) -> np.ndarray:
    """The [-1,1) noise block [B,T,hop] (float32) the CUDA noise kernel generates for these arguments."""
    j = np.arange(hop, dtype=np.uint64)
    t = np.arange(T, dtype=np.uint64)
    b = np.arange(B, dtype=np.uint64) + np.uint64(voice_offset)
    quad = (j >> np.uint64(2))[None, None, :]
    words = philox4x32_10(quad, t[None, :, None], b[:, None, None], np.uint64(offset & 0xFFFFFFFF),
                          seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    stacked = np.stack(words, axis=-1)                              # [B,T,hop,4]
    lane = (j & np.uint64(3)).astype(np.int64)
    w = np.take_along_axis(stacked, np.broadcast_to(lane[None, None, :, None], (B, T, hop, 1)), axis=-1)[..., 0]
    u = (w >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)
    return (u * np.float32(2.0) - np.float32(1.0)).astype(np.float32)
