/*
 * oracle/cpu_fft.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE (see cpu_fft.h).
 *
 * Recursive decimation-in-time mixed-radix complex FFT in float32, plus the
 * textbook even-length real<->half-complex split.  Twiddles are evaluated in
 * double and rounded once to float, which is what pffft and pocketfft (float
 * instantiation) do as well.
 */
#include "cpu_fft.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

struct offt_plan {
    int n;
    int nfac;
    int fac[64];   /* pairs (p, m): radix p, remaining length m */
    ocpx *tw;      /* tw[k] = exp(-2*pi*i*k/n), k in [0,n) */
    ocpx *rtw;     /* rtw[k] = exp(-2*pi*i*k/(2n)), k in [0,n/2]: real split */
};

static void factorize(int n, int *fac, int *nfac)
{
    int p = 4, cnt = 0;
    /* radix 4 first, then 2, 3, 5, then whatever odd primes are left */
    while (n > 1) {
        while (n % p) {
            switch (p) {
            case 4: p = 2; break;
            case 2: p = 3; break;
            default: p += 2; break;
            }
            if ((long)p * p > n) p = n;
        }
        n /= p;
        fac[cnt++] = p;
        fac[cnt++] = n;
    }
    *nfac = cnt / 2;
}

offt_plan *offt_plan_create(int n)
{
    offt_plan *p = (offt_plan *)calloc(1, sizeof(*p));
    if (!p || n < 1) { free(p); return NULL; }
    p->n = n;
    factorize(n, p->fac, &p->nfac);
    p->tw = (ocpx *)malloc(sizeof(ocpx) * (size_t)n);
    p->rtw = (ocpx *)malloc(sizeof(ocpx) * (size_t)(n / 2 + 1));
    for (int k = 0; k < n; ++k) {
        double a = -2.0 * M_PI * (double)k / (double)n;
        p->tw[k].re = (float)cos(a);
        p->tw[k].im = (float)sin(a);
    }
    for (int k = 0; k <= n / 2; ++k) {
        double a = -M_PI * (double)k / (double)n;
        p->rtw[k].re = (float)cos(a);
        p->rtw[k].im = (float)sin(a);
    }
    return p;
}

void offt_plan_destroy(offt_plan *p)
{
    if (!p) return;
    free(p->tw);
    free(p->rtw);
    free(p);
}

int offt_plan_length(const offt_plan *p) { return p->n; }

static inline ocpx cmul(ocpx a, ocpx b)
{
    ocpx r = { a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re };
    return r;
}
static inline ocpx cadd(ocpx a, ocpx b) { ocpx r = { a.re + b.re, a.im + b.im }; return r; }
static inline ocpx csub(ocpx a, ocpx b) { ocpx r = { a.re - b.re, a.im - b.im }; return r; }

/* twiddle lookup honouring direction */
static inline ocpx twd(const offt_plan *st, size_t idx, int inverse)
{
    ocpx w = st->tw[idx];
    if (inverse) w.im = -w.im;
    return w;
}

static void bfly2(ocpx *out, size_t fstride, const offt_plan *st, int m, int inv)
{
    for (int k = 0; k < m; ++k) {
        ocpx t = cmul(out[k + m], twd(st, fstride * k, inv));
        out[k + m] = csub(out[k], t);
        out[k] = cadd(out[k], t);
    }
}

static void bfly3(ocpx *out, size_t fstride, const offt_plan *st, int m, int inv)
{
    const float s60 = (float)(inv ? 0.86602540378443864676 : -0.86602540378443864676);
    for (int k = 0; k < m; ++k) {
        ocpx a = cmul(out[k + m], twd(st, fstride * k, inv));
        ocpx b = cmul(out[k + 2 * m], twd(st, 2 * fstride * k, inv));
        ocpx s = cadd(a, b), d = csub(a, b);
        ocpx h = { out[k].re - 0.5f * s.re, out[k].im - 0.5f * s.im };
        d.re *= s60; d.im *= s60;
        out[k] = cadd(out[k], s);
        out[k + m].re = h.re - d.im;      out[k + m].im = h.im + d.re;
        out[k + 2 * m].re = h.re + d.im;  out[k + 2 * m].im = h.im - d.re;
    }
}

static void bfly4(ocpx *out, size_t fstride, const offt_plan *st, int m, int inv)
{
    for (int k = 0; k < m; ++k) {
        ocpx a = out[k];
        ocpx b = cmul(out[k + m], twd(st, fstride * k, inv));
        ocpx c = cmul(out[k + 2 * m], twd(st, 2 * fstride * k, inv));
        ocpx d = cmul(out[k + 3 * m], twd(st, 3 * fstride * k, inv));
        ocpx s0 = cadd(a, c), s1 = csub(a, c), s2 = cadd(b, d), s3 = csub(b, d);
        out[k] = cadd(s0, s2);
        out[k + 2 * m] = csub(s0, s2);
        if (inv) {
            out[k + m].re = s1.re - s3.im;     out[k + m].im = s1.im + s3.re;
            out[k + 3 * m].re = s1.re + s3.im; out[k + 3 * m].im = s1.im - s3.re;
        } else {
            out[k + m].re = s1.re + s3.im;     out[k + m].im = s1.im - s3.re;
            out[k + 3 * m].re = s1.re - s3.im; out[k + 3 * m].im = s1.im + s3.re;
        }
    }
}

static void bfly5(ocpx *out, size_t fstride, const offt_plan *st, int m, int inv)
{
    const ocpx ya = twd(st, fstride * m, inv);      /* exp(-/+2pi i/5)   */
    const ocpx yb = twd(st, 2 * fstride * m, inv);  /* exp(-/+2pi i*2/5) */
    for (int k = 0; k < m; ++k) {
        ocpx x0 = out[k];
        ocpx x1 = cmul(out[k + m], twd(st, fstride * k, inv));
        ocpx x2 = cmul(out[k + 2 * m], twd(st, 2 * fstride * k, inv));
        ocpx x3 = cmul(out[k + 3 * m], twd(st, 3 * fstride * k, inv));
        ocpx x4 = cmul(out[k + 4 * m], twd(st, 4 * fstride * k, inv));
        ocpx s14 = cadd(x1, x4), d14 = csub(x1, x4);
        ocpx s23 = cadd(x2, x3), d23 = csub(x2, x3);
        out[k].re = x0.re + s14.re + s23.re;
        out[k].im = x0.im + s14.im + s23.im;
        ocpx c1 = { x0.re + s14.re * ya.re + s23.re * yb.re, x0.im + s14.im * ya.re + s23.im * yb.re };
        ocpx c2 = { x0.re + s14.re * yb.re + s23.re * ya.re, x0.im + s14.im * yb.re + s23.im * ya.re };
        /* e1 = i*(ya.im*d14 + yb.im*d23), e2 = i*(yb.im*d14 - ya.im*d23) */
        ocpx e1 = { -(d14.im * ya.im + d23.im * yb.im), d14.re * ya.im + d23.re * yb.im };
        ocpx e2 = { -(d14.im * yb.im - d23.im * ya.im), d14.re * yb.im - d23.re * ya.im };
        out[k + m] = cadd(c1, e1);
        out[k + 4 * m] = csub(c1, e1);
        out[k + 2 * m] = cadd(c2, e2);
        out[k + 3 * m] = csub(c2, e2);
    }
}

static void bfly_generic(ocpx *out, size_t fstride, const offt_plan *st, int m, int p, int inv)
{
    int n = st->n;
    ocpx *scr = (ocpx *)malloc(sizeof(ocpx) * (size_t)p);
    for (int u = 0; u < m; ++u) {
        for (int q = 0; q < p; ++q) scr[q] = out[u + q * m];
        int k = u;
        for (int q = 0; q < p; ++q) {
            size_t tidx = 0;
            ocpx acc = scr[0];
            for (int j = 1; j < p; ++j) {
                tidx += fstride * (size_t)k;
                if (tidx >= (size_t)n) tidx %= (size_t)n;
                acc = cadd(acc, cmul(scr[j], twd(st, tidx, inv)));
            }
            out[k] = acc;
            k += m;
        }
    }
    free(scr);
}

static void work(ocpx *out, const ocpx *in, size_t fstride, const int *fac, const offt_plan *st, int inv)
{
    const int p = fac[0], m = fac[1];
    if (m == 1) {
        for (int k = 0; k < p; ++k) out[k] = in[(size_t)k * fstride];
    } else {
        for (int k = 0; k < p; ++k)
            work(out + (size_t)k * m, in + (size_t)k * fstride, fstride * p, fac + 2, st, inv);
    }
    switch (p) {
    case 2: bfly2(out, fstride, st, m, inv); break;
    case 3: bfly3(out, fstride, st, m, inv); break;
    case 4: bfly4(out, fstride, st, m, inv); break;
    case 5: bfly5(out, fstride, st, m, inv); break;
    default: bfly_generic(out, fstride, st, m, p, inv); break;
    }
}

void offt_c2c(const offt_plan *p, const ocpx *in, ocpx *out, int inverse)
{
    if (p->n == 1) { out[0] = in[0]; return; }
    work(out, in, 1, p->fac, p, inverse);
}

void offt_r2c(const offt_plan *ph, const float *in, ocpx *out, ocpx *scratch)
{
    const int h = ph->n;            /* n = 2h */
    offt_c2c(ph, (const ocpx *)in, scratch, 0);
    /* X[k] = E[k] + w^k O[k];  E = (Z[k]+conj Z[h-k])/2, O = (Z[k]-conj Z[h-k])/(2i) */
    out[0].re = scratch[0].re + scratch[0].im; out[0].im = 0.f;
    out[h].re = scratch[0].re - scratch[0].im; out[h].im = 0.f;
    for (int k = 1; k <= h / 2; ++k) {
        ocpx a = scratch[k], b = scratch[h - k];
        ocpx e = { 0.5f * (a.re + b.re), 0.5f * (a.im - b.im) };
        ocpx o = { 0.5f * (a.im + b.im), -0.5f * (a.re - b.re) };
        ocpx w = ph->rtw[k];
        ocpx t = cmul(o, w);
        out[k].re = e.re + t.re;      out[k].im = e.im + t.im;
        out[h - k].re = e.re - t.re;  out[h - k].im = -(e.im - t.im);
    }
}

void offt_c2r(const offt_plan *ph, const ocpx *in, float *out, ocpx *scratch)
{
    const int h = ph->n;
    /* Z[k] = E[k] + i*O[k], E = X[k]+conj X[h-k], O = (X[k]-conj X[h-k]) * conj(w^k)
     * (the factor 2 relative to r2c's split makes the pair unscaled: out = n * x) */
    scratch[0].re = in[0].re + in[h].re;
    scratch[0].im = in[0].re - in[h].re;
    for (int k = 1; k <= h / 2; ++k) {
        ocpx a = in[k], b = in[h - k];
        ocpx e = { a.re + b.re, a.im - b.im };
        ocpx d = { a.re - b.re, a.im + b.im };
        ocpx w = ph->rtw[k]; w.im = -w.im;
        ocpx o = cmul(d, w);
        scratch[k].re = e.re - o.im;      scratch[k].im = e.im + o.re;
        scratch[h - k].re = e.re + o.im;  scratch[h - k].im = -(e.im - o.re);
    }
    offt_c2c(ph, scratch, (ocpx *)out, 1);
}
