"""Deterministic synthetic rectified pairs: a counter-based splitmix64 hash, so the
same pair can be regenerated anywhere (bench.py on every rank, tests, the C oracle) without
shipping data.  numpy only; independent of the oracle (tests check both generators agree)."""
import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(z):
    z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return z ^ (z >> np.uint64(31))


def _hN(seed, pair, y, x, c):
    with np.errstate(over="ignore"):
        k = (np.uint64(seed) ^ (np.uint64(pair) * np.uint64(0x9E3779B97F4A7C15))) ^ (y.astype(np.uint64) << np.uint64(40)) \
            ^ (x.astype(np.uint64) << np.uint64(8)) ^ c.astype(np.uint64)
        return (_splitmix64(k) & np.uint64(0xFF)).astype(np.int64)


def synth_pair(seed, pair, w, h, min_disp, max_disp, quant=False):
    """Returns (left, right) HxWx3 uint8 BGR.  Right image = low-frequency lattice + noise in
    [16,239] (never black); left image = right shifted by a piecewise-constant true disparity
    g(y,x) = min_disp + ((x>>6)*5 + (y>>5)*3) % (max_disp-min_disp-1), plus +-2 noise."""
    with np.errstate(over="ignore"):
        y = np.arange(h, dtype=np.int64)[:, None, None]
        x = np.arange(w, dtype=np.int64)[None, :, None]
        c = np.arange(3, dtype=np.int64)[None, None, :]
        yy, xx, cc = np.broadcast_arrays(y, x, c)
        coarse = _hN(seed, pair, yy >> 2, xx >> 2, cc)
        fine = _hN(seed, pair, yy, xx, cc)
        R = 16 + ((((3 * coarse + fine) >> 2) * 7) >> 3)
        if quant:
            R = 16 + (R >> 6) * 64
        L = 1 | _hN(seed + 1, pair, yy, xx, cc)
        span = max(max_disp - min_disp - 1, 1)
        g = min_disp + (((np.arange(w) >> 6) * 5)[None, :] + ((np.arange(h) >> 5) * 3)[:, None]) % span
        tx = np.arange(w)[None, :] + g                      # target column of each right pixel
        # "for x ascending: L[y, x+g] = R[y, x]": later x overwrite earlier ones
        for row in range(h):
            ok = tx[row] < w
            L[row, tx[row][ok]] = R[row, np.nonzero(ok)[0]]  # numpy assigns in index order: last write wins
        if not quant:
            noise = _hN(seed + 2, pair, yy, xx, cc) % 5 - 2
            L = np.clip(L + noise, 1, 255)
    return L.astype(np.uint8), R.astype(np.uint8)


def effective_evals(w1, w2, rows, block, min_disp, max_disp):
    """Reference-semantics evaluation count for `rows` rows of a pair without black pixels:
    candidates before the break x + d + r >= w1 (BlockSearch.cpp:147-149)."""
    half = (block - 1) // 2
    x = np.arange(w2)
    r = np.minimum(w2 - x - 1, half)
    lim = np.minimum(max_disp, w1 - x - r)
    return int(np.maximum(lim - min_disp, 0).sum()) * int(rows)
