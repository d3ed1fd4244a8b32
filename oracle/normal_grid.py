"""N(0,1) rounded to the bf16 grid (oracle; test infrastructure only).

The reference draws eps with ``randn_tensor(latents.shape, dtype=bf16)`` (train_sana.py:183): an fp32 standard normal
rounded to nearest-even bf16.  This module gives the exact probability mass of every bf16 value under that rounding
(float64 normal CDF at the rounding boundaries), so a sample of the device's counter-RNG noise can be tested on the
grid it lives on: discrete Kolmogorov-Smirnov distance, tail mass, moments of the rounded variable.
"""
import numpy as np
import torch


def bf16_values():
    """All finite bf16 values, ascending, as float64, and their 16-bit patterns."""
    bits = np.arange(65536, dtype=np.uint32)
    vals = (bits << 16).astype(np.uint32).view(np.float32).astype(np.float64)
    ok = np.isfinite(vals) & ~((bits == 0x8000))            # drop inf / nan and -0 (rounding of a normal never gives -0... kept out)
    bits, vals = bits[ok], vals[ok]
    order = np.argsort(vals, kind="stable")
    return vals[order], bits[order]


def _ndtr(x):
    return torch.special.ndtr(torch.from_numpy(np.asarray(x, dtype=np.float64))).numpy()


def bf16_normal_bins():
    """(values [n] ascending, bit patterns [n], mass [n], cdf [n]) of Z ~ N(0,1) rounded to nearest bf16.
    ``cdf[i]`` = P(rounded Z <= values[i]).  Round-to-nearest: the boundaries are the midpoints of adjacent values
    (ties have probability 0)."""
    vals, bits = bf16_values()
    mid = 0.5 * (vals[1:] + vals[:-1])
    # upper tail evaluated as a survival function so that the far tails keep relative accuracy
    cdf_mid = _ndtr(mid)
    sf_mid = _ndtr(-mid)
    cdf = np.concatenate([cdf_mid, [1.0]])
    lo = np.concatenate([[0.0], cdf_mid])
    mass = cdf - lo
    # for positive values use the survival form (1 - cdf loses digits beyond 8 sigma; irrelevant for tests, cheap to do right)
    sf_lo = np.concatenate([[1.0], sf_mid])
    sf_hi = np.concatenate([sf_mid, [0.0]])
    mass_sf = sf_lo - sf_hi
    mass = np.where(vals > 0, mass_sf, mass)
    return vals, bits, mass, cdf


def discrete_ks(counts_by_bits, n=None):
    """counts_by_bits: int64 [65536] histogram over bf16 bit patterns.  Returns (D, n, value at which D is reached)
    against the rounded normal."""
    vals, bits, mass, cdf = bf16_normal_bins()
    c = np.asarray(counts_by_bits, dtype=np.float64)
    # +0 and -0 are the same value: fold the -0 pattern into +0
    c = c.copy()
    c[0x0000] += c[0x8000]
    c[0x8000] = 0.0
    n = float(c.sum()) if n is None else float(n)
    emp = np.cumsum(c[bits]) / n
    d = np.abs(emp - cdf)
    i = int(np.argmax(d))
    return float(d[i]), n, float(vals[i])


def rounded_normal_moments():
    """mean, variance, skewness, excess kurtosis of the bf16-rounded standard normal (exact sums over the grid)."""
    vals, _, mass, _ = bf16_normal_bins()
    m1 = float((mass * vals).sum())
    mu2 = float((mass * (vals - m1) ** 2).sum())
    mu3 = float((mass * (vals - m1) ** 3).sum())
    mu4 = float((mass * (vals - m1) ** 4).sum())
    return m1, mu2, mu3 / mu2 ** 1.5, mu4 / mu2 ** 2 - 3.0
