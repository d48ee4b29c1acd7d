"""Fit of the branch-free erfc used by gelu_fast (csrc/common.cuh): erfc(t) ~= 2^-q(t) on [0, 4.2]."""
import numpy as np
from scipy.special import erf, erfc

T = 4.2
t = np.cos(np.pi * (np.arange(8000) + 0.5) / 8000) * T / 2 + T / 2
q = -np.log2(erfc(t))
co = np.polynomial.chebyshev.Chebyshev.fit(t, q, 7, domain=[0, T]).convert(kind=np.polynomial.Polynomial).coef
print("coefficients (c0..c7):", [float(np.float32(c)) for c in co])


def gelu_fast(x):
    x = x.astype(np.float32)
    ax = np.abs(x)
    tt = np.minimum(ax * np.float32(0.70710678), np.float32(T))
    acc = np.float32(co[7]) * np.ones_like(tt)
    for k in range(6, -1, -1):
        acc = (acc * tt + np.float32(co[k])).astype(np.float32)
    return np.maximum(x, 0) - np.float32(0.5) * ax * np.exp2(-acc).astype(np.float32)


x = np.linspace(-12, 12, 2_000_001)
print("max |gelu_fast - gelu|:", np.abs(gelu_fast(x) - 0.5 * x * (1 + erf(x / np.sqrt(2)))).max())
