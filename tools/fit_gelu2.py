#!/usr/bin/env python
"""Fits the clamp-free GELU of the bf16 fc1 epilogue (csrc/gelu_fast.cuh, gelu_sat2).

gelu(x) = x * Phi(x);  Phi(x) ~= sat(0.5 + x * R(v)),  v = x^2,  R a degree-DEG polynomial in v (power basis, Horner),
sat = clamp to [0, 1] (the .sat modifier of the last FMA).  R is fitted on |x| <= C (Chebyshev least squares weighted
by the error it causes in gelu); beyond C the fit only has to keep |x * R(v)| >= 0.5 with the sign of x, which is
checked here up to |x| = 1e4 (and holds at infinity when the leading coefficient is positive).  Prints coefficients
and the maximum absolute error of the float32 evaluation against the exact erf GELU (torchvision convnext.py:56).
"""
import sys

import numpy as np
from numpy.polynomial import chebyshev as Ch
from numpy.polynomial import polynomial as Pl
from scipy.special import erf


def fit(C, DEG):
    n = 8000
    u = np.cos(np.pi * (np.arange(n) + 0.5) / n)
    x = np.sqrt((u + 1) / 2) * C
    y = 0.5 * np.where(x > 1e-8, erf(x / np.sqrt(2)) / np.maximum(x, 1e-8), np.sqrt(2 / np.pi))
    coef = Ch.chebfit(u, y, DEG, w=x * x + 1e-3)
    pu = Ch.cheb2poly(coef)                      # power basis in u = 2 v / C^2 - 1
    pv = np.zeros(1)
    lin = np.array([-1.0, 2.0 / (C * C)])
    for a in pu[::-1]:
        pv = Pl.polyadd(Pl.polymul(pv, lin), [a])
    return pv


def eval_f32(co, x):
    f = np.float32
    x = x.astype(f)
    v = x * x
    r = np.full_like(v, f(co[-1]))
    for a in co[-2::-1]:
        r = r * v + f(a)
    with np.errstate(over="ignore", invalid="ignore"):
        phi = np.clip(x * r + f(0.5), f(0), f(1))
    return x * phi


if __name__ == "__main__":
    best = None
    for DEG in (4, 5, 6, 7):
        for C in np.arange(3.0, 5.01, 0.125):
            co = fit(C, DEG)
            xs = np.concatenate([np.linspace(-12, 12, 480001), np.geomspace(12, 1e4, 20000), -np.geomspace(12, 1e4, 20000)])
            ex = xs * 0.5 * (1 + erf(xs / np.sqrt(2)))
            err = np.abs(eval_f32(co, xs).astype(np.float64) - ex)
            ok = co[-1] > 0
            print("deg %d C %.3f lead %+.2e max err %.3e at %.3f %s" % (DEG, C, co[-1], err.max(), xs[err.argmax()], "" if ok else "(negative lead)"))
            if ok and (best is None or err.max() < best[0]) and DEG == int(sys.argv[1] if len(sys.argv) > 1 else 5):
                best = (err.max(), C, DEG, co)
    e, C, DEG, co = best
    print("// tools/fit_gelu2.py: fit on |x| <= %g, degree %d in x^2, max |gelu_sat - gelu| = %.3e" % (C, DEG, e))
    print("constexpr float GELU_R[%d] = {%s};" % (DEG + 1, ", ".join("%.9gf" % a for a in co)))
