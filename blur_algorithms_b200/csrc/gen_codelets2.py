#!/usr/bin/env python3
"""Generates dft_codelets2.cuh: straight-line small DFTs on packed complex values.

Each codelet `cdftR_f` / `cdftR_b` (forward e^{-i..} / backward e^{+i..}, unscaled) transforms R
complex values held as cplx x[R] (cplx2.cuh) in place, natural order in and out, using only the packed
complex primitives of cplx2.cuh (one FADD2 / FMUL2 / FFMA2 each).  They are the per-thread
butterflies of the overlap-save FFT passes in ols_core.cuh.

Construction (symbolic; emitted as SSA `const float2 tN = op(...);` statements):
  * base radices 2, 3, 4, 5 by hand;
  * R = P*Q, gcd(P,Q) == 1  -> Good-Thomas prime-factor map (no internal twiddles);
  * otherwise               -> Cooley-Tukey with constant twiddles.
Multiplications by i^k are never executed: a value carries a pending rotation `rot` (times i^rot)
that is folded into the next addition (cadd / caddi / csub / csubi, cfmar / cfmai).

Usage: python gen_codelets2.py > dft_codelets2.cuh   (the csrc Makefile does this)
"""
import math
import sys
from math import gcd

RADICES = [2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16, 18, 20, 24, 25, 27, 30, 32]

SPLIT = {
    6: (2, 3), 8: (2, 4), 9: (3, 3), 10: (2, 5), 12: (3, 4), 15: (3, 5), 16: (4, 4),
    18: (2, 9), 20: (4, 5), 24: (3, 8), 25: (5, 5), 27: (3, 9), 30: (5, 6), 32: (4, 8),
}


def lit(v: float) -> str:
    return f"{v:.9e}f"


class Gen:
    """Values are (name, rot): the complex number  name * i^rot."""

    def __init__(self, sign: int):
        self.s = sign          # -1 forward, +1 backward
        self.lines = []
        self.n = 0
        self.ops = 0

    def e(self, expr: str) -> str:
        self.n += 1
        v = f"t{self.n}"
        self.lines.append(f"    const cplx {v} = {expr};")
        self.ops += 1
        return v

    # A + B
    def add(self, A, B):
        (a, ra), (b, rb) = A, B
        if ra != 0 and rb == 0:
            (a, ra), (b, rb) = (b, rb), (a, ra)
        d = (rb - ra) % 4
        fn = ("cadd", "caddi", "csub", "csubi")[d]
        return (self.e(f"{fn}({a}, {b})"), ra)

    def sub(self, A, B):
        return self.add(A, (B[0], (B[1] + 2) % 4))

    # A + c*B  (c real python float)
    def fma(self, A, c, B):
        (a, ra), (b, rb) = A, B
        d = (rb - ra) % 4
        if d == 0:
            return (self.e(f"cfmar({a}, {lit(c)}, {b})"), ra)
        if d == 2:
            return (self.e(f"cfmar({a}, {lit(-c)}, {b})"), ra)
        if d == 1:
            return (self.e(f"cfmai({a}, {lit(c)}, {b})"), ra)
        return (self.e(f"cfmai({a}, {lit(-c)}, {b})"), ra)

    def scale(self, A, c):
        return (self.e(f"cscale({A[0]}, {lit(c)})"), A[1])

    def rot(self, A, k):
        return (A[0], (A[1] + k) % 4)

    def cmul_const(self, A, num, den):
        """A * exp(s*2*pi*i*num/den)"""
        num %= den
        if num == 0:
            return A
        g = gcd(num, den)
        num, den = num // g, den // g
        if den == 2:
            return self.rot(A, 2)
        if den == 4:
            ang_pos = (self.s * num) % 4      # exp(s*2*pi*i*num/4) = i^(s*num)
            return self.rot(A, ang_pos)
        ang = self.s * 2.0 * math.pi * num / den
        c, sn = math.cos(ang), math.sin(ang)
        # a*(c + i sn) = c*a + sn*(i a)
        t = self.scale(A, c)
        return self.fma(t, sn, self.rot(A, 1))

    def materialise(self, A):
        a, r = A
        if r == 0:
            return a
        if r == 2:
            return self.e(f"cscale({a}, -1.0f)")
        zero = "czero()"
        return self.e(f"{'caddi' if r == 1 else 'csubi'}({zero}, {a})")

    # ---- base DFTs
    def dft2(self, x):
        return [self.add(x[0], x[1]), self.sub(x[0], x[1])]

    def dft3(self, x):
        s = self.s
        t1 = self.add(x[1], x[2])
        d = self.sub(x[1], x[2])
        X0 = self.add(x[0], t1)
        m1 = self.fma(x[0], -0.5, t1)
        k = math.sqrt(3.0) / 2.0
        # X1 = m1 + s*i*k*d ; X2 = m1 - s*i*k*d
        X1 = self.fma(m1, s * k, self.rot(d, 1))
        X2 = self.fma(m1, -s * k, self.rot(d, 1))
        return [X0, X1, X2]

    def dft4(self, x):
        s = self.s
        s0 = self.add(x[0], x[2])
        s1 = self.sub(x[0], x[2])
        s2 = self.add(x[1], x[3])
        s3 = self.sub(x[1], x[3])
        X0 = self.add(s0, s2)
        X2 = self.sub(s0, s2)
        # X1 = s1 + s*i*s3 ; X3 = s1 - s*i*s3
        X1 = self.add(s1, self.rot(s3, s % 4))
        X3 = self.sub(s1, self.rot(s3, s % 4))
        return [X0, X1, X2, X3]

    def dft5(self, x):
        s = self.s
        c1, c2 = math.cos(2 * math.pi / 5), math.cos(4 * math.pi / 5)
        s1, s2 = math.sin(2 * math.pi / 5), math.sin(4 * math.pi / 5)
        a1 = self.add(x[1], x[4]); b1 = self.sub(x[1], x[4])
        a2 = self.add(x[2], x[3]); b2 = self.sub(x[2], x[3])
        X0 = self.add(self.add(x[0], a1), a2)
        p1 = self.fma(self.fma(x[0], c1, a1), c2, a2)
        p2 = self.fma(self.fma(x[0], c2, a1), c1, a2)
        q1 = self.fma(self.scale(b1, s1), s2, b2)       # s1*b1 + s2*b2
        q2 = self.fma(self.scale(b1, s2), -s1, b2)      # s2*b1 - s1*b2
        # X1 = p1 + s*i*q1, X4 = p1 - s*i*q1, X2 = p2 + s*i*q2, X3 = p2 - s*i*q2
        r = s % 4
        X1 = self.add(p1, self.rot(q1, r)); X4 = self.sub(p1, self.rot(q1, r))
        X2 = self.add(p2, self.rot(q2, r)); X3 = self.sub(p2, self.rot(q2, r))
        return [X0, X1, X2, X3, X4]

    def dft(self, x):
        R = len(x)
        if R == 1:
            return x
        if R == 2:
            return self.dft2(x)
        if R == 3:
            return self.dft3(x)
        if R == 4:
            return self.dft4(x)
        if R == 5:
            return self.dft5(x)
        P, Q = SPLIT[R]
        out = [None] * R
        if gcd(P, Q) == 1:
            qi = pow(Q, -1, P)
            pi = pow(P, -1, Q)
            A = [[None] * Q for _ in range(P)]
            for n2 in range(Q):
                col = self.dft([x[(Q * n1 + P * n2) % R] for n1 in range(P)])
                for k1 in range(P):
                    A[k1][n2] = col[k1]
            for k1 in range(P):
                row = self.dft(A[k1])
                for k2 in range(Q):
                    out[(Q * qi * k1 + P * pi * k2) % R] = row[k2]
            return out
        A = [[None] * Q for _ in range(P)]
        for j2 in range(Q):
            col = self.dft([x[j1 * Q + j2] for j1 in range(P)])
            for k1 in range(P):
                A[k1][j2] = self.cmul_const(col[k1], j2 * k1, R)
        for k1 in range(P):
            row = self.dft(A[k1])
            for k2 in range(Q):
                out[k1 + P * k2] = row[k2]
        return out


def codelet(R: int, sign: int) -> str:
    g = Gen(sign)
    xin = [(f"x[{j}]", 0) for j in range(R)]
    # inputs are read through SSA names so the in-place write-back below is safe
    xin = []
    for j in range(R):
        g.n += 1
        g.lines.append(f"    const cplx t{g.n} = x[{j}];")
        xin.append((f"t{g.n}", 0))
    y = g.dft(xin)
    outs = [g.materialise(v) for v in y]
    name = f"cdft{R}_{'f' if sign < 0 else 'b'}"
    body = "\n".join(g.lines)
    wb = "\n".join(f"    x[{k}] = {outs[k]};" for k in range(R))
    return (f"// radix {R} {'forward' if sign < 0 else 'backward'}: {g.ops} packed instructions\n"
            f"BLUR_HD void {name}(cplx* __restrict__ x)\n{{\n{body}\n{wb}\n}}\n")


def main() -> None:
    out = []
    out.append("// GENERATED by gen_codelets2.py - do not edit.\n#pragma once\n#include \"cplx2.cuh\"\n")
    out.append("namespace blurb200 {\n")
    for R in RADICES:
        out.append(codelet(R, -1))
        out.append(codelet(R, +1))
    out.append("template <int R, bool INV> struct CDft;\n")
    out.append("template <bool INV> struct CDft<1, INV> { static BLUR_HD void run(cplx*) {} };")
    for R in RADICES:
        out.append(f"template <> struct CDft<{R}, false> {{ static BLUR_HD void run(cplx* x) {{ cdft{R}_f(x); }} }};")
        out.append(f"template <> struct CDft<{R}, true>  {{ static BLUR_HD void run(cplx* x) {{ cdft{R}_b(x); }} }};")
    out.append("\n}  // namespace blurb200\n")
    sys.stdout.write("\n".join(out))


if __name__ == "__main__":
    main()
