#!/usr/bin/env python3
"""Generates dft_codelets3.cuh: small DFTs on packed complex values with FUSED twiddles.

Same contract as gen_codelets2.py (cplx x[R] in place, natural order in and out, one packed
instruction per primitive), plus two things the two-factor overlap-save passes (ols2_core.cuh)
need:

  * a value may carry a PENDING twiddle w (a compile-time constant, or a run-time complex held in
    a register pair).  A radix-2 butterfly on (a, w*b) is then
        s = a + w*b   = fma(fma(a, w.re, b), w.im, i*b)      two packed FMAs, no separate multiply
        d = a - w*b   = 2a - s                               one packed FMA
    i.e. three instructions instead of four (multiply 2 + add + sub).  Internal twiddles of the
    Cooley-Tukey splits become pending values, so every codelet is shorter than its
    gen_codelets2 twin (radix 32: 216 -> see the generated comments).
  * `cdftR_f_tw(x, w)` / `cdftR_b_twc(x, w)`: forward DFT of x[j]*w[j] / backward DFT of
    x[j]*conj(w[j]) (j >= 1; w[0] is not read).  The inter-pass twiddles W_L^{jk} of a
    two-factor FFT ride into the first butterflies this way.

Usage: python gen_codelets3.py > dft_codelets3.cuh   (the csrc Makefile does this)
"""
import math
import sys
from math import gcd

RADICES = [16, 32, 64]
SPLIT = {8: (2, 4), 16: (4, 4), 32: (4, 8), 64: (8, 8)}


def lit(v: float) -> str:
    return f"{v:.9e}f"


class Val:
    """The complex number  name * i^rot * tw   (tw = None | ('c', cos, sin) | ('r', expr, conj))."""
    __slots__ = ("name", "rot", "tw")

    def __init__(self, name, rot=0, tw=None):
        self.name, self.rot, self.tw = name, rot % 4, tw


class Gen:
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

    # ---- twiddle helpers: (re, im) of a pending twiddle as C expressions + python sign of im
    @staticmethod
    def tw_parts(tw):
        if tw[0] == 'c':
            return lit(tw[1]), lit(abs(tw[2])), (1 if tw[2] >= 0 else -1)
        _, expr, conj = tw
        return f"cre({expr})", f"cim({expr})", (-1 if conj else 1)

    def plain_add(self, a, ra, b, rb):
        """name of (a*i^ra + b*i^rb) expressed with rotation ra"""
        d = (rb - ra) % 4
        fn = ("cadd", "caddi", "csub", "csubi")[d]
        return self.e(f"{fn}({a}, {b})")

    def fma_real(self, a, ra, c_expr, b, rb, neg=False):
        """a*i^ra + (+-c) * b*i^rb  with rotation ra; c_expr is a C float expression (>= 0 or any)."""
        d = (rb - ra) % 4
        if neg:
            d = (d + 2) % 4
        fn = ("cfmar", "cfmai", "cfmsr", "cfmsi")[d]
        return self.e(f"{fn}({a}, {c_expr}, {b})")

    def materialise_tw(self, A: Val) -> Val:
        """A with its pending twiddle executed (2 packed instructions)."""
        if A.tw is None:
            return A
        re, im, sg = self.tw_parts(A.tw)
        # w*a = re*a + im*(i a)
        t = self.e(f"cscale({A.name}, {re})")
        t2 = self.e(f"{'cfmai' if sg > 0 else 'cfmsi'}({t}, {im}, {A.name})")
        return Val(t2, A.rot, None)

    # (A + B, A - B)
    def addsub(self, A: Val, B: Val):
        if A.tw is not None and B.tw is None:
            # (A + B, A - B) = (B + A, -(B - A)): keep it simple, materialise A
            A = self.materialise_tw(A)
        if B.tw is None:
            s = self.plain_add(A.name, A.rot, B.name, B.rot)
            d = self.plain_add(A.name, A.rot, B.name, B.rot + 2)
            return Val(s, A.rot), Val(d, A.rot)
        A = self.materialise_tw(A)
        re, im, sg = self.tw_parts(B.tw)
        t = self.fma_real(A.name, A.rot, re, B.name, B.rot)
        s = self.fma_real(t, A.rot, im, B.name, B.rot + (1 if sg > 0 else 3))
        d = self.e(f"c2amb({A.name}, {s})")
        return Val(s, A.rot), Val(d, A.rot)

    def rot(self, A: Val, k) -> Val:
        return Val(A.name, A.rot + k, A.tw)

    def cmul_const(self, A: Val, num, den) -> Val:
        """A * exp(s*2*pi*i*num/den), as a pending twiddle unless it is a power of i."""
        assert A.tw is None
        num %= den
        if num == 0:
            return A
        g = gcd(num, den)
        num, den = num // g, den // g
        if den == 2:
            return self.rot(A, 2)
        if den == 4:
            return self.rot(A, (self.s * num) % 4)
        ang = self.s * 2.0 * math.pi * num / den
        # reduce to |angle| <= pi/4 around a multiple of pi/2 so constants stay well conditioned
        return Val(A.name, A.rot, ('c', math.cos(ang), math.sin(ang)))

    def finish(self, A: Val) -> str:
        A = self.materialise_tw(A)
        if A.rot == 0:
            return A.name
        if A.rot == 2:
            return self.e(f"cscale({A.name}, -1.0f)")
        return self.e(f"{'caddi' if A.rot == 1 else 'csubi'}(czero(), {A.name})")

    # ---- DFTs
    def dft2(self, x):
        return list(self.addsub(x[0], x[1]))

    def dft4(self, x):
        s = self.s
        s0, s1 = self.addsub(x[0], x[2])
        s2, s3 = self.addsub(x[1], x[3])
        X0, X2 = self.addsub(s0, s2)
        X1, X3 = self.addsub(s1, self.rot(s3, s % 4))
        return [X0, X1, X2, X3]

    def dft(self, x):
        R = len(x)
        if R == 1:
            return x
        if R == 2:
            return self.dft2(x)
        if R == 4:
            return self.dft4(x)
        P, Q = SPLIT[R]
        A = [[None] * Q for _ in range(P)]
        for j2 in range(Q):
            col = self.dft([x[j1 * Q + j2] for j1 in range(P)])
            for k1 in range(P):
                A[k1][j2] = self.cmul_const(col[k1], j2 * k1, R)
        out = [None] * R
        for k1 in range(P):
            row = self.dft(A[k1])
            for k2 in range(Q):
                out[k1 + P * k2] = row[k2]
        return out


def codelet(R: int, sign: int, tw: bool) -> str:
    g = Gen(sign)
    xin = []
    for j in range(R):
        g.n += 1
        g.lines.append(f"    const cplx t{g.n} = x[{j}];")
        t = None
        if tw and j > 0:
            t = ('r', f"w[{j}]", sign > 0)
        xin.append(Val(f"t{g.n}", 0, t))
    y = g.dft(xin)
    outs = [g.finish(v) for v in y]
    name = f"cdft{R}v_{'f' if sign < 0 else 'b'}" + (("_tw" if sign < 0 else "_twc") if tw else "")
    body = "\n".join(g.lines)
    wb = "\n".join(f"    x[{k}] = {outs[k]};" for k in range(R))
    args = "cplx* __restrict__ x" + (", const cplx* __restrict__ w" if tw else "")
    what = f"radix {R} {'forward' if sign < 0 else 'backward'}" + \
           ((" of x[j]*w[j]" if sign < 0 else " of x[j]*conj(w[j])") if tw else "")
    return f"// {what}: {g.ops} packed instructions\nBLUR_HD void {name}({args})\n{{\n{body}\n{wb}\n}}\n"


def main() -> None:
    out = ["// GENERATED by gen_codelets3.py - do not edit.\n#pragma once\n#include \"cplx2.cuh\"\n", "namespace blurb200 {\n"]
    for R in RADICES:
        out.append(codelet(R, -1, False))
        out.append(codelet(R, +1, False))
        out.append(codelet(R, -1, True))
        out.append(codelet(R, +1, True))
    out.append("template <int R> struct CDftV;")
    for R in RADICES:
        out.append(f"template <> struct CDftV<{R}> {{\n"
                   f"    static BLUR_HD void fwd(cplx* x) {{ cdft{R}v_f(x); }}\n"
                   f"    static BLUR_HD void bwd(cplx* x) {{ cdft{R}v_b(x); }}\n"
                   f"    static BLUR_HD void fwd_tw(cplx* x, const cplx* w) {{ cdft{R}v_f_tw(x, w); }}\n"
                   f"    static BLUR_HD void bwd_twc(cplx* x, const cplx* w) {{ cdft{R}v_b_twc(x, w); }}\n}};")
    out.append("\n}  // namespace blurb200\n")
    sys.stdout.write("\n".join(out))


if __name__ == "__main__":
    main()
