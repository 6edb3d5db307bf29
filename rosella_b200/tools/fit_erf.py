#!/usr/bin/env python
"""Fits the polynomial used by erf_fast() in csrc/knn_depth.cu:
    erf(x) = sign(x) * (1 - 2^(-t g(t))),  t = min(|x|, 4),  g(t) ~ -log2(erfc(t)) / t.
Weighted (by d erf / d g) iteratively re-weighted least squares on Chebyshev nodes, then the
FP32 Horner evaluation is simulated to report the achieved absolute error."""
import numpy as np
from scipy.special import erf, erfc

T = 4.0


def fit(deg):
    t = np.cos(np.linspace(0, np.pi, 4001)) * T / 2 + T / 2
    t = t[t > 1e-9]
    g = -np.log2(erfc(t)) / t
    w = erfc(t) * np.log(2) * t
    W = w.copy()
    for _ in range(60):
        c = np.polynomial.polynomial.polyfit(t, g, deg, w=W)
        e = w * (np.polynomial.polynomial.polyval(t, c) - g)
        W = W * (1 + 0.5 * np.abs(e) / np.abs(e).max())
    return c


def simulate(c):
    c32 = c.astype(np.float32)
    tt = np.linspace(0, T, 200001).astype(np.float32)
    p = np.full_like(tt, c32[-1])
    for k in range(len(c) - 2, -1, -1):
        p = (p * tt + c32[k]).astype(np.float32)
    y = (p * tt).astype(np.float32)
    val = np.float32(1) - np.exp2(-y.astype(np.float64)).astype(np.float32)
    return np.abs(val.astype(np.float64) - erf(tt.astype(np.float64))).max()


if __name__ == "__main__":
    for deg in (5, 6, 7):
        c = fit(deg)
        print(deg, "max abs err (FP32 Horner)", simulate(c))
        print("   coefficients c0..c%d:" % deg, [float(x) for x in c])
