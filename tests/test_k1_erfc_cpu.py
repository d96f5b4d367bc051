"""K1's erfc (mpmc_b200/csrc/pair_kernel.cu: erfc_fast) on the CPU: the 17 coefficients compiled into the kernel are the ones
profiles/k1_erfc_coeffs.py derives (50-digit Chebyshev interpolation of exp(x^2) erfc(x) in the mapped variable, converted to
monomials), and the kernel's arithmetic -- exp(-x^2) times a Horner sum in (x - 4) / (x + 4), restated here in double precision --
stays within 1e-14 of a 50-digit reference on the whole range it is used on (0 <= x <= 6).  The GPU parity tests then hold the
energies that use it to 1e-10 against the reference's fixtures."""
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _kernel_coefficients():
    src = open(os.path.join(ROOT, "mpmc_b200", "csrc", "pair_kernel.cu")).read()
    m = re.search(r"erfc_poly\[17\]\s*=\s*\{([^}]*)\}", src)
    assert m, "erfc_poly[17] not found in pair_kernel.cu"
    body = re.sub(r"//[^\n]*", "", m.group(1))
    vals = [float(x) for x in body.replace("\n", " ").split(",") if x.strip()]
    assert len(vals) == 17
    return vals


def test_kernel_coefficients_are_the_generated_ones():
    sys.path.insert(0, os.path.join(ROOT, "profiles"))
    import k1_erfc_coeffs as gen

    want = gen.fit()
    got = _kernel_coefficients()
    assert np.allclose(got, want, rtol=0.0, atol=1e-17), np.abs(np.array(got) - np.array(want)).max()


def test_erfc_fast_arithmetic_against_a_50_digit_reference():
    import mpmath as mp

    mp.mp.dps = 50
    coef = _kernel_coefficients()
    x = np.concatenate([np.linspace(0.0, 6.0, 3001), np.random.default_rng(7).uniform(0.0, 6.0, 2000)])
    y = (x - 4.0) * (1.0 / (x + 4.0))  # the kernel multiplies by a reciprocal refined to full precision
    p = np.zeros_like(y) + coef[16]
    for a in coef[15::-1]:
        p = p * y + a
    got = np.exp(-(x * x)) * p
    ref = np.array([float(mp.erfc(mp.mpf(float(v)))) for v in x])
    err = np.abs(got - ref) / ref
    assert err.max() <= 1e-14, err.max()
    assert abs(got[0] - 1.0) <= 1e-15  # erfc(0) = 1 to a few ulp
