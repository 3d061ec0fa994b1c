#!/usr/bin/env python
"""Coefficients of the degree-8 polynomial of 2^f on |f| <= 1/2 used by exp10_small
(epik_b200/csrc/place_kernel.cuh) and its relative error bound. p(0) = 1 exactly: the fit is of
g(f) = (2^f - 1) / f at Chebyshev nodes (near-minimax), so p(f) = 1 + f * g(f).
usage: python tools/exp2_poly.py [degree]"""
import sys

import numpy as np
from numpy.polynomial import chebyshev as Ch


def fit(deg):
    n = 400
    x = np.cos(np.pi * (np.arange(n) + 0.5) / n) * 0.5
    xl = x.astype(np.longdouble)
    ln2 = np.log(np.longdouble(2))
    g = np.where(np.abs(xl) > 1e-6, np.expm1(xl * ln2) / xl, ln2 * (1 + xl * ln2 / 2))
    c = Ch.chebfit(np.asarray(x * 2, np.float64), np.asarray(g, np.float64), deg - 1)
    pc = Ch.cheb2poly(c)
    coef = np.concatenate([[1.0], [pc[i] * 2.0 ** i for i in range(len(pc))]])
    xs = np.linspace(-0.5, 0.5, 200001).astype(np.longdouble)
    val = np.zeros_like(xs)
    for cc in coef[::-1]:
        val = val * xs + np.longdouble(cc)
    return coef, float(np.max(np.abs(val / np.exp2(xs) - 1)))


if __name__ == "__main__":
    deg = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    coef, err = fit(deg)
    print("degree %d, max relative error %.3g" % (deg, err))
    print("c%d .. c1 (Horner order):" % deg)
    print(",\n".join("    %r" % float(v) for v in coef[:0:-1]))
