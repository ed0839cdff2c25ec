"""ORACLE -- test infrastructure only (tests/ may import this; the product package must never do so).

numpy restatement of the dropout mask function of csrc/dropout.cuh, so that the masks the kernels draw are pinned by
something that is not CUDA. The reference draws its masks from torch's generator (nn.Dropout and MHA's dropout:
mingpt/model.py:140,150,183,215); no counter-based generator can reproduce those bits, so what the reference fixes is
the DISTRIBUTION -- every element kept independently with probability 1 - p and scaled by 1/(1 - p) -- and the function
below is this repository's own definition of a mask with that distribution. "parity unpinned" by the reference, pinned
by this file.

    keep(seed, site, i) for the i-th element of a dropout site:
        r  = Philox4x32, 7 rounds, key = (seed lo, seed hi), counter = (i >> 4 lo, i >> 4 hi, site, 0x6d677062)
        w  = r[(i & 15) >> 2]                       one of the four 32-bit outputs
        e  = i & 3
        v  = (byte e of w) << 8 | byte (e + 1) % 4 of w       exactly uniform on [0, 65536)
        keep = v >= round(p * 65536)

Element numbering: an elementwise site numbers the elements of its contiguous tensor; the attention-probability site
numbers (b, h, q, k) as ((b * H + h) * T + q) * 8 * ceil(T / 8) + k (rows padded to a multiple of 8 keys).
"""
import numpy as np

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
ROUNDS = 7
MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32(key0, key1, c0, c1, c2, c3, rounds=ROUNDS):
    """vectorised over the counters (uint64 arrays holding 32-bit values); key is scalar"""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint64(key0), np.uint64(key1)
    for _ in range(rounds):
        p0 = np.uint64(M0) * c0
        p1 = np.uint64(M1) * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0 = (k0 + np.uint64(W0)) & MASK32
        k1 = (k1 + np.uint64(W1)) & MASK32
    return c0, c1, c2, c3


def thresh16(p: float) -> int:
    t = np.float32(p) * np.float32(65536.0) + np.float32(0.5)
    return 0 if t <= 0 else (65535 if t >= 65535 else int(t))


def keep_elements(seed: int, site: int, idx, p: float) -> np.ndarray:
    """bool array: element idx[j] of the site is kept"""
    idx = np.asarray(idx, dtype=np.uint64)
    g = idx >> np.uint64(4)
    r = philox4x32(seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF, g & MASK32, g >> np.uint64(32),
                   np.full(idx.shape, site, dtype=np.uint64), np.full(idx.shape, 0x6D677062, dtype=np.uint64))
    word_sel = (idx & np.uint64(15)) >> np.uint64(2)
    w = np.choose(word_sel.astype(np.int64), r)
    e = idx & np.uint64(3)
    hi = (w >> (np.uint64(8) * e)) & np.uint64(0xFF)
    lo = (w >> (np.uint64(8) * ((e + np.uint64(1)) & np.uint64(3)))) & np.uint64(0xFF)
    return ((hi << np.uint64(8)) | lo) >= np.uint64(thresh16(p))


def elementwise_keep(seed: int, site: int, n: int, p: float) -> np.ndarray:
    return keep_elements(seed, site, np.arange(n, dtype=np.uint64), p)


def attention_keep(seed: int, site: int, B: int, H: int, T: int, p: float) -> np.ndarray:
    """keep[b, h, q, k] of the attention-probability dropout"""
    t8 = (T + 7) // 8
    row = np.arange(B * H * T, dtype=np.uint64).reshape(B, H, T, 1)
    k = np.arange(T, dtype=np.uint64).reshape(1, 1, 1, T)
    return keep_elements(seed, site, row * np.uint64(8 * t8) + k, p)
