"""Derives the degree-11 polynomial used by the branch-free exp() of the kernel-row kernels (csrc/common.cuh: exp_nb).

exp(x) = 2^k exp(r), k = rint(x log2 e), |r| <= ln2 / 2.  The polynomial interpolates exp at the 12 Chebyshev nodes of
[-a, a], a = 0.3466 (exact rational arithmetic on 60-digit values, then rounded to double); its truncation error is
~3e-18, so what remains is the rounding of the Horner evaluation (< 1 ulp).  Prints the coefficients as hex doubles and
the measured error of the double-precision evaluation against 60-digit exp on a dense grid.

    python tools/exp_poly.py
"""
import math
import struct
from decimal import Decimal, getcontext
from fractions import Fraction

getcontext().prec = 60
DEG = 11
A = Fraction(3466, 10000)


def dec_cos(x: Decimal) -> Decimal:
    s, term, k = Decimal(0), Decimal(1), 0
    while abs(term) > Decimal(10) ** -58:
        s += term
        k += 2
        term = -term * x * x / (k * (k - 1))
    return s


def coefficients():
    """The DEG + 1 monomial coefficients as doubles, c0 first."""
    pi = Decimal('3.14159265358979323846264338327950288419716939937510582097494')
    nodes = [A * Fraction(dec_cos(pi * (2 * i + 1) / (2 * (DEG + 1)))) for i in range(DEG + 1)]
    vals = [Fraction((Decimal(n.numerator) / Decimal(n.denominator)).exp()) for n in nodes]
    # exact solve of the Vandermonde system
    m = [[n ** j for j in range(DEG + 1)] + [v] for n, v in zip(nodes, vals)]
    for c in range(DEG + 1):
        piv = next(r for r in range(c, DEG + 1) if m[r][c] != 0)
        m[c], m[piv] = m[piv], m[c]
        m[c] = [v / m[c][c] for v in m[c]]
        for r in range(DEG + 1):
            if r != c and m[r][c] != 0:
                f = m[r][c]
                m[r] = [a - f * b for a, b in zip(m[r], m[c])]
    return [float(m[j][DEG + 1]) for j in range(DEG + 1)]


def main():
    coef = coefficients()
    for j, c in enumerate(coef):
        print(f'c{j:<2d} = {c!r:26s} 0x{struct.unpack("<Q", struct.pack("<d", c))[0]:016x}   (1/{j}! = {1 / math.factorial(j)!r})')
    worst = 0.0
    steps = 20001
    for i in range(steps):
        r = -0.34658 + 2 * 0.34658 * i / (steps - 1)
        p = coef[DEG]
        for j in range(DEG - 1, -1, -1):
            p = p * r + coef[j]  # (the kernel uses fma: one rounding fewer per step)
        exact = Decimal(r).exp()
        worst = max(worst, abs(float((Decimal(p) - exact) / exact)))
    print(f'max relative error of the double Horner evaluation on |r| <= 0.34658: {worst:.3e} ({worst / 2 ** -53:.2f} x 2^-53)')


if __name__ == '__main__':
    main()
