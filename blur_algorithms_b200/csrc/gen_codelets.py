#!/usr/bin/env python3
"""Generates dft_codelets.cuh: straight-line, register-resident small DFTs.

Each codelet `dftR_f` / `dftR_b` (forward e^{-i..} / backward e^{+i..}, unscaled) transforms
R complex values held as two float arrays (xr[], xi[]) in place, natural order in and out.
They are the per-thread butterflies of the shared-memory FFT passes in fft_core.cuh.

Construction (symbolic, emitted as SSA `const float tN = ...;` statements so ptxas sees pure
register code with literal constants):
  * base radices 2, 3, 4, 5 written out by hand below;
  * R = P*Q with gcd(P,Q) == 1  -> Good-Thomas prime-factor map, no internal twiddles;
  * otherwise                   -> Cooley-Tukey (P-point DFTs, constant twiddles w_R^{j2*k1},
                                   Q-point DFTs); multiples of 90/45 degrees are special-cased.

Usage: python gen_codelets.py > dft_codelets.cuh   (build.py does this)
"""
import math
import sys
from math import gcd

RADICES = [2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16, 18, 20, 24, 25, 27, 30, 32]

# preferred split R -> (P, Q)
SPLIT = {
    6: (2, 3), 8: (2, 4), 9: (3, 3), 10: (2, 5), 12: (3, 4), 15: (3, 5), 16: (4, 4),
    18: (2, 9), 20: (4, 5), 24: (3, 8), 25: (5, 5), 27: (3, 9), 30: (5, 6), 32: (4, 8),
}


def lit(v: float) -> str:
    return f"{v:.9e}f"   # 9 significant digits round-trip float32


class Gen:
    def __init__(self, sign: int):
        self.s = sign          # -1 forward, +1 backward
        self.lines = []
        self.n = 0
        self.flops = 0

    def e(self, expr: str) -> str:
        self.n += 1
        v = f"t{self.n}"
        self.lines.append(f"    const float {v} = {expr};")
        self.flops += 1
        return v

    # complex helpers: values are (re, im) variable names
    def add(self, a, b):
        return (self.e(f"{a[0]} + {b[0]}"), self.e(f"{a[1]} + {b[1]}"))

    def sub(self, a, b):
        return (self.e(f"{a[0]} - {b[0]}"), self.e(f"{a[1]} - {b[1]}"))

    def mul_i(self, a, sgn):
        """a * (sgn*i): (re,im) -> (-sgn*im, sgn*re).  Returned lazily as negated names."""
        if sgn > 0:
            return (f"(-{a[1]})", a[0])
        return (a[1], f"(-{a[0]})")

    def cmul_const(self, a, num, den):
        """a * exp(s*2*pi*i*num/den)"""
        num %= den
        if num == 0:
            return a
        g = gcd(num, den)
        num, den = num // g, den // g
        ang = self.s * 2.0 * math.pi * num / den
        c, sn = math.cos(ang), math.sin(ang)
        if den == 2:                       # -1
            return (f"(-{a[0]})", f"(-{a[1]})")
        if den == 4:                       # +-i
            return self.mul_i(a, 1 if sn > 0 else -1)
        if den == 8:                       # (+-1 +- i)/sqrt2
            h = lit(math.sqrt(0.5))
            sc, ss = (1 if c > 0 else -1), (1 if sn > 0 else -1)
            # re = h*(sc*ar - ss*ai), im = h*(ss*ar + sc*ai)
            re_in = self.e(f"{'' if sc > 0 else '-'}{a[0]} {'-' if ss > 0 else '+'} {a[1]}")
            im_in = self.e(f"{'' if ss > 0 else '-'}{a[0]} {'+' if sc > 0 else '-'} {a[1]}")
            return (self.e(f"{h} * {re_in}"), self.e(f"{h} * {im_in}"))
        # general: re = c*ar - sn*ai ; im = sn*ar + c*ai   (mul + fma each)
        re = self.e(f"fmaf({lit(c)}, {a[0]}, {lit(-sn)} * {a[1]})")
        im = self.e(f"fmaf({lit(sn)}, {a[0]}, {lit(c)} * {a[1]})")
        self.flops += 2
        return (re, im)

    # ---- base DFTs: list of complex in natural order -> list of complex in natural order
    def dft2(self, x):
        return [self.add(x[0], x[1]), self.sub(x[0], x[1])]

    def dft3(self, x):
        s = self.s
        t1 = self.add(x[1], x[2])
        d = self.sub(x[1], x[2])
        m0 = self.add(x[0], t1)
        m1 = (self.e(f"fmaf(-0.5f, {t1[0]}, {x[0][0]})"), self.e(f"fmaf(-0.5f, {t1[1]}, {x[0][1]})"))
        k = lit(math.sqrt(3.0) / 2.0)
        # X1 = m1 + s*i*k*d ; X2 = m1 - s*i*k*d ;  i*d = (-d.im, d.re)
        if s < 0:
            X1 = (self.e(f"fmaf({k}, {d[1]}, {m1[0]})"), self.e(f"fmaf(-{k}, {d[0]}, {m1[1]})"))
            X2 = (self.e(f"fmaf(-{k}, {d[1]}, {m1[0]})"), self.e(f"fmaf({k}, {d[0]}, {m1[1]})"))
        else:
            X1 = (self.e(f"fmaf(-{k}, {d[1]}, {m1[0]})"), self.e(f"fmaf({k}, {d[0]}, {m1[1]})"))
            X2 = (self.e(f"fmaf({k}, {d[1]}, {m1[0]})"), self.e(f"fmaf(-{k}, {d[0]}, {m1[1]})"))
        return [m0, X1, X2]

    def dft4(self, x):
        s = self.s
        s0 = self.add(x[0], x[2])
        s1 = self.sub(x[0], x[2])
        s2 = self.add(x[1], x[3])
        s3 = self.sub(x[1], x[3])
        X0 = self.add(s0, s2)
        X2 = self.sub(s0, s2)
        # X1 = s1 + s*i*s3 ; X3 = s1 - s*i*s3 ; i*s3 = (-s3.im, s3.re)
        if s < 0:
            X1 = (self.e(f"{s1[0]} + {s3[1]}"), self.e(f"{s1[1]} - {s3[0]}"))
            X3 = (self.e(f"{s1[0]} - {s3[1]}"), self.e(f"{s1[1]} + {s3[0]}"))
        else:
            X1 = (self.e(f"{s1[0]} - {s3[1]}"), self.e(f"{s1[1]} + {s3[0]}"))
            X3 = (self.e(f"{s1[0]} + {s3[1]}"), self.e(f"{s1[1]} - {s3[0]}"))
        return [X0, X1, X2, X3]

    def dft5(self, x):
        s = self.s
        c1, c2 = math.cos(2 * math.pi / 5), math.cos(4 * math.pi / 5)
        s1, s2 = math.sin(2 * math.pi / 5), math.sin(4 * math.pi / 5)
        a1 = self.add(x[1], x[4]); b1 = self.sub(x[1], x[4])
        a2 = self.add(x[2], x[3]); b2 = self.sub(x[2], x[3])
        X0 = (self.e(f"{x[0][0]} + {a1[0]} + {a2[0]}"), self.e(f"{x[0][1]} + {a1[1]} + {a2[1]}"))
        self.flops += 2
        p1 = tuple(self.e(f"fmaf({lit(c2)}, {a2[q]}, fmaf({lit(c1)}, {a1[q]}, {x[0][q]}))") for q in (0, 1))
        p2 = tuple(self.e(f"fmaf({lit(c1)}, {a2[q]}, fmaf({lit(c2)}, {a1[q]}, {x[0][q]}))") for q in (0, 1))
        q1 = tuple(self.e(f"fmaf({lit(s2)}, {b2[q]}, {lit(s1)} * {b1[q]})") for q in (0, 1))
        q2 = tuple(self.e(f"fmaf({lit(-s1)}, {b2[q]}, {lit(s2)} * {b1[q]})") for q in (0, 1))
        self.flops += 8
        # X1 = p1 + s*i*q1, X4 = p1 - s*i*q1, X2 = p2 + s*i*q2, X3 = p2 - s*i*q2 ; i*q = (-q.im, q.re)
        def pm(p, q, sg):
            if sg > 0:
                return (self.e(f"{p[0]} - {q[1]}"), self.e(f"{p[1]} + {q[0]}"))
            return (self.e(f"{p[0]} + {q[1]}"), self.e(f"{p[1]} - {q[0]}"))
        X1 = pm(p1, q1, s); X4 = pm(p1, q1, -s)
        X2 = pm(p2, q2, s); X3 = pm(p2, q2, -s)
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
            # Good-Thomas: n = (Q*n1 + P*n2) % R ; k = (Q*qi*k1 + P*pi*k2) % R
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
        # Cooley-Tukey: j = j1*Q + j2 ; k = k1 + P*k2
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
    x = [(f"xr[{j}]", f"xi[{j}]") for j in range(R)]
    # read inputs into SSA temps first so in-place write-back is safe
    xin = [(g.e(a), g.e(b)) for a, b in x]
    g.flops = 0
    y = g.dft(xin)
    name = f"dft{R}_{'f' if sign < 0 else 'b'}"
    body = "\n".join(g.lines)
    wb = "\n".join(f"    xr[{k}] = {y[k][0]}; xi[{k}] = {y[k][1]};" for k in range(R))
    return (f"// radix {R} {'forward' if sign < 0 else 'backward'}: ~{g.flops} flop-instructions\n"
            f"BLUR_HD void {name}(float* __restrict__ xr, float* __restrict__ xi)\n{{\n{body}\n{wb}\n}}\n")


def main() -> None:
    out = []
    out.append("// GENERATED by gen_codelets.py - do not edit.\n#pragma once\n")
    out.append("#ifndef BLUR_HD\n#ifdef __CUDACC__\n#define BLUR_HD __host__ __device__ __forceinline__\n#else\n"
               "#include <cmath>\n#define BLUR_HD inline\n#endif\n#endif\n")
    out.append("namespace blurb200 {\n")
    for R in RADICES:
        out.append(codelet(R, -1))
        out.append(codelet(R, +1))
    out.append("template <int R, bool INV> struct Dft;\n")
    for R in RADICES:
        out.append(f"template <> struct Dft<{R}, false> {{ static BLUR_HD void run(float* xr, float* xi) {{ dft{R}_f(xr, xi); }} }};")
        out.append(f"template <> struct Dft<{R}, true>  {{ static BLUR_HD void run(float* xr, float* xi) {{ dft{R}_b(xr, xi); }} }};")
    out.append("\n}  // namespace blurb200\n")
    sys.stdout.write("\n".join(out))


if __name__ == "__main__":
    main()
