"""Accuracy of the hot-path exp (gpb_build.cu: exp_nonpos4) in exact rational arithmetic: every FMA is
evaluated exactly and rounded once, the result is compared with a 60-digit exp.  Prints the worst error in ulps."""
from fractions import Fraction as F
from decimal import Decimal, getcontext
import math, random, struct
getcontext().prec = 60
C = [2.5052097064908941E-008, 2.7626262793835868E-007, 2.7557414788000726E-006, 2.4801504602132958E-005,
     1.9841269707468915E-004, 1.3888888932258898E-003, 8.3333333333978320E-003, 4.1666666666573905E-002,
     1.6666666666666563E-001, 5.0000000000000056E-001, 1.0, 1.0]
L2E = 1.4426950408889634; MAGIC = 6755399441055744.0
LN2HI = 6.93147180559945286e-01; LN2LO = 2.31904681384629956e-17
def fma(a, b, c): return float(F(a) * F(b) + F(c))
def exp_neg(x):
    t = fma(x, L2E, MAGIC)
    k = struct.unpack('<i', struct.pack('<d', t)[:4])[0]
    t = t - MAGIC
    r = fma(t, -LN2HI, x); r = fma(t, -LN2LO, r)
    p = C[0]
    for c in C[1:]: p = fma(p, r, c)
    return math.ldexp(p, k), k
def ulp(x): return math.ulp(x)
random.seed(1)
worst = 0; worst_lib = 0
for i in range(30000):
    x = -random.random() * (700 if i % 2 else 30)
    if i % 7 == 0: x = -random.random() * 1e-3
    e, k = exp_neg(x)
    ex = Decimal(x).exp()
    err = abs((Decimal(e) - ex) / Decimal(ulp(e)))
    errl = abs((Decimal(math.exp(x)) - ex) / Decimal(ulp(e)))
    worst = max(worst, float(err)); worst_lib = max(worst_lib, float(errl))
print("worst ulp error (ours)", worst, " libm", worst_lib)
