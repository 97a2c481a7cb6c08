"""numpy restatement of the counter-based dropout RNG used by the CUDA kernels.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference uses torch's Philox dropout (``nn.Dropout`` at
``src/model/bsarec.py:12,85`` and ``src/model/_modules.py:35,94,98``) whose stream
cannot be reproduced from a custom kernel, so the kernels define their own
stateless generator and this file restates it bit-for-bit so train-mode parity
(p > 0) can be checked mask-for-mask:

    key(site)   = splitmix64(seed + GOLDEN * (site + 1))
    word(i)     = splitmix64(key + (i >> 2))            # one 64-bit word per 4 elements
    r16(i)      = (word >> (16 * (i & 3))) & 0xFFFF
    keep(i)     = r16(i) >= round(p * 65536)
    y           = keep ? x * (1 / (1 - p)) : 0          # inverted dropout, as torch

``i`` is the row-major linear index of the element in the site's logical tensor
([B,L,d] for hidden dropouts, [B,H,L,L] for attention-probability dropout).
Site ids: 0 = embedding; for layer l: 1+4l filter, 2+4l attention probs,
3+4l attention output, 4+4l feed-forward output.
"""
import numpy as np

GOLDEN = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(z):
    """Vectorised splitmix64 finaliser on uint64 arrays (wrap-around arithmetic)."""
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z + GOLDEN
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        z = z ^ (z >> np.uint64(31))
    return z


def site_key(seed: int, site: int) -> np.uint64:
    with np.errstate(over="ignore"):
        s = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + GOLDEN * np.uint64(site + 1)
    return splitmix64(s)


def dropout_threshold(p: float) -> int:
    return int(round(float(p) * 65536.0))


def dropout_keep_mask(seed: int, site: int, shape, p: float) -> np.ndarray:
    """Boolean keep-mask of ``shape`` for dropout probability ``p``."""
    n = int(np.prod(shape))
    if p <= 0.0:
        return np.ones(shape, dtype=bool)
    key = site_key(seed, site)
    idx = np.arange(n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        words = splitmix64(key + (idx >> np.uint64(2)))
    lane = (idx & np.uint64(3)) * np.uint64(16)
    r16 = (words >> lane) & np.uint64(0xFFFF)
    return (r16 >= np.uint64(dropout_threshold(p))).reshape(shape)
