/*
 * oracle/cpu_fft.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C float32 mixed-radix (2/3/4/5) FFT used by the CPU oracle and by the
 * API shims that stand in for the reference's absent third-party FFT engines
 * (pffft @ fbc40586, pocketfft @ bb87ca50).  It restates the
 * *published* algorithm those libraries implement (Cooley–Tukey decimation in
 * time, float32 data, twiddles rounded from double), not their source.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may link or call this.
 */
#ifndef ORACLE_CPU_FFT_H
#define ORACLE_CPU_FFT_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float re, im; } ocpx;

typedef struct offt_plan offt_plan;

/* Complex plan of length n (n must factor into 2,3,5; any other prime factor
 * falls back to a slow generic butterfly so every n works). */
offt_plan *offt_plan_create(int n);
void offt_plan_destroy(offt_plan *p);
int offt_plan_length(const offt_plan *p);

/* out[k] = sum_n in[n] * exp(-/+ 2*pi*i*k*n/N); inverse!=0 selects '+'.
 * Unscaled.  in and out must not alias. */
void offt_c2c(const offt_plan *p, const ocpx *in, ocpx *out, int inverse);

/* Real transforms of even length n built on a complex plan of length n/2.
 * r2c: out has n/2+1 bins (unscaled, forward sign '-').
 * c2r: in has n/2+1 bins, out n reals, unscaled (gives n * x).
 * plan_half must have length n/2.  scratch: n/2 ocpx (caller provided). */
void offt_r2c(const offt_plan *plan_half, const float *in, ocpx *out, ocpx *scratch);
void offt_c2r(const offt_plan *plan_half, const ocpx *in, float *out, ocpx *scratch);

#ifdef __cplusplus
}
#endif
#endif
