#!/usr/bin/env python
"""Fits the polynomial used by the bf16-mode GELU epilogue (csrc/gelu_fast.cuh).

gelu(x) = x * Phi(x),  Phi(x) = 0.5 + 0.5 * erf(x / sqrt 2).  For |x| <= C:  erf(x / sqrt 2) ~= x * P(u),
u = 2 x^2 / C^2 - 1 in [-1, 1], P of degree DEG (Chebyshev least squares weighted by x^2, i.e. by the error it causes
in gelu); outside, x is clamped to +-C first (Phi is 0 / 1 there to 3.4e-6).  Prints the coefficients (power basis in
u, ascending) and the maximum absolute error of the float32 evaluation against the exact erf GELU.
"""
import numpy as np
from numpy.polynomial import chebyshev as Ch
from scipy.special import erf

C, DEG = 4.5, 8


def fit():
    n = 8000
    u = np.cos(np.pi * (np.arange(n) + 0.5) / n)
    x = np.sqrt((u + 1) / 2) * C
    y = np.where(x > 1e-8, erf(x / np.sqrt(2)) / np.maximum(x, 1e-8), np.sqrt(2 / np.pi))
    coef = Ch.chebfit(u, y, DEG, w=0.5 * x * x + 1e-3)
    return Ch.cheb2poly(coef)


def eval_f32(co, x):
    f = np.float32
    x = x.astype(f)
    xc = np.clip(x, f(-C), f(C))
    u = (xc * f(2.0 / (C * C))) * xc - f(1)
    r = np.full_like(u, f(co[-1]))
    for a in co[-2::-1]:
        r = r * u + f(a)
    phi = (f(0.5) * xc) * r + f(0.5)
    return x * phi


if __name__ == "__main__":
    co = fit()
    xs = np.linspace(-12, 12, 2400001)
    ex = xs * 0.5 * (1 + erf(xs / np.sqrt(2)))
    err = np.abs(eval_f32(co, xs).astype(np.float64) - ex)
    print("// tools/fit_gelu.py: C = %g, degree %d, max |gelu_fast - gelu| = %.3e at x = %.3f" % (C, DEG, err.max(), xs[err.argmax()]))
    print("constexpr float GELU_C = %.9gf;" % C)
    print("constexpr float GELU_P[%d] = {%s};" % (DEG + 1, ", ".join("%.9gf" % a for a in co)))
