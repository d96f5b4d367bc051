"""Coefficients of K1's erfc (pair_kernel.cu: erfc_fast): erfc(x) = exp(-x^2) * P((x - 4) / (x + 4)) on 0 <= x <= 6.
exp(x^2) erfc(x) is interpolated at 17 Chebyshev nodes of the mapped variable in 50-digit arithmetic (mpmath), the Chebyshev
sum is converted to monomials in the same precision, and the rounded coefficients are checked against mpmath with a double
Horner evaluation: max relative error 3.5e-15 (degree 16; sum |a_k| = 1.0, so Horner is well conditioned).
usage: python profiles/k1_erfc_coeffs.py"""
import mpmath as mp
import numpy as np

mp.mp.dps = 50
C, XM, DEG = 4.0, 6.0, 16


def erfcx(x):
    return mp.erfc(x) * mp.exp(mp.mpf(x) ** 2)


def fit(c=C, deg=DEG, xm=XM):
    ymax = mp.mpf(xm - c) / (xm + c)
    n = deg + 1
    us = [mp.cos(mp.pi * (k + mp.mpf(1) / 2) / n) for k in range(n)]
    ys = [(u * (ymax + 1) / 2) + (ymax - 1) / 2 for u in us]
    fs = [erfcx(c * (1 + y) / (1 - y)) for y in ys]
    cj = [2 * sum(fs[k] * mp.cos(mp.pi * j * (k + mp.mpf(1) / 2) / n) for k in range(n)) / n for j in range(n)]
    cj[0] /= 2
    a, b = 2 / (ymax + 1), -(ymax - 1) / (ymax + 1)  # u = a y + b

    def padd(p, q):
        m = max(len(p), len(q))
        return [(p[i] if i < len(p) else 0) + (q[i] if i < len(q) else 0) for i in range(m)]

    def pmul(p, q):
        r = [mp.mpf(0)] * (len(p) + len(q) - 1)
        for i, pi in enumerate(p):
            for j, qj in enumerate(q):
                r[i + j] += pi * qj
        return r

    T = [[mp.mpf(1)], [b, a]]
    for j in range(2, n):
        T.append(padd(pmul([2 * b, 2 * a], T[j - 1]), [-t for t in T[j - 2]]))
    poly = [mp.mpf(0)]
    for j in range(n):
        poly = padd(poly, [cj[j] * t for t in T[j]])
    return [float(p) for p in poly]


if __name__ == "__main__":
    coef = fit()
    xt = np.linspace(0, XM, 6001)
    yt = (xt - C) / (xt + C)
    p = np.zeros_like(yt) + coef[-1]
    for a in coef[-2::-1]:
        p = p * yt + a
    ref = np.array([float(erfcx(mp.mpf(float(x)))) for x in xt])
    print("max relative error of P against exp(x^2) erfc(x): %.2e" % np.max(np.abs(p - ref) / ref))
    print(",\n".join("%.17g" % a for a in coef))
