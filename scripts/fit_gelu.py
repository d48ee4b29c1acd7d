"""Fit of the branch-free erfc used by gelu_fast (csrc/common.cuh): 0.5 erfc(a / sqrt 2) ~= 2^-q(a), a = |x|.

The fit is weighted by (a * 0.5 erfc(a / sqrt 2))^0.25: the error of gelu is a * 2^-q * ln2 * dq, so accuracy of q
matters most where that product is large. The polynomial is used without a clamp: check that q keeps growing
beyond the fitted range.
"""
import numpy as np
from scipy.special import erf, erfc

DEG = 7
T = 4.2 * np.sqrt(2)
a = np.cos(np.pi * (np.arange(8000) + 0.5) / 8000) * T / 2 + T / 2
q = 1.0 - np.log2(erfc(a / np.sqrt(2)))
w = (a * 0.5 * erfc(a / np.sqrt(2)) + 1e-4) ** 0.25
co = np.polynomial.chebyshev.Chebyshev.fit(a, q, DEG, domain=[0, T], w=w).convert(kind=np.polynomial.Polynomial).coef
co32 = [np.float32(c) for c in co]
print("coefficients (c0..c7):", [float(c) for c in co32])


def gelu_fast(x):
    x = x.astype(np.float32)
    ax = np.abs(x)
    acc = co32[DEG] * np.ones_like(ax)
    for k in range(DEG - 1, -1, -1):
        acc = (acc * ax + co32[k]).astype(np.float32)
    with np.errstate(all="ignore"):
        return np.maximum(x, 0) - ax * np.exp2(-acc).astype(np.float32), acc


x = np.linspace(-12, 12, 2_000_001)
y, _ = gelu_fast(x)
print("max |gelu_fast - gelu| on [-12, 12]:", np.abs(y - 0.5 * x * (1 + erf(x / np.sqrt(2)))).max())
x = np.linspace(-1e4, 1e4, 2_000_001)
y, acc = gelu_fast(x)
print("min q beyond the fitted range:", acc[np.abs(x) > T].min(), " max rel err on [-1e4, 1e4]:",
      (np.abs(y - 0.5 * x * (1 + erf(x / np.sqrt(2)))) / np.maximum(np.abs(x), 1)).max())
