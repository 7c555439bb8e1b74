/*
 * oracle/blur_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE (see blur_oracle.h).
 *
 * CPU restatement of the reference hot path.  Citations are into /root/reference
 * (michelerenzullo/Blur_algorithms @ 2f873a4e).  The FFT arithmetic of the absent
 * pffft / pocketfft submodules is restated by oracle/cpu_fft.c.
 */
#include "blur_oracle.h"
#include "cpu_fft.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---------------------------------------------------------------- sizing */

/* Source.cpp:60-73.  `radius` is a float initialised from a double expression;
 * `radius * 2 + .5f` is float arithmetic truncated to int; the clamp happens
 * before the odd fix-up, so the result may exceed max_width by one. */
int orc_gaussian_window(double sigma, int max_width)
{
    const float radius = (float)(sigma * sqrt(2 * log(255.0)) - 1);
    int width = (int)(radius * 2 + .5f);
    if (max_width) width = width < max_width ? width : max_width;
    if (width % 2 == 0) ++width;
    return width;
}

/* Source.cpp:75-102.  Loop variable is a float, y*y is float, the division,
 * exp and normalisation are double, storage is float.  Then rotate left by
 * width/2 so the centre tap sits at index 0 (Source.cpp:96-100). */
void orc_get_gaussian(float *kernel, double sigma, int width, int fft_length)
{
    if (!width) width = orc_gaussian_window(sigma, 0);
    const int size = fft_length ? fft_length : width;
    const float mid_w = (width - 1) / 2.f;
    const double s = 2. * sigma * sigma;
    int i = 0;
    for (float y = -mid_w; y <= mid_w; ++y, ++i)
        kernel[i] = (float)((exp(-(double)(y * y) / s)) / (3.14159265358979323846 * s));
    double acc = 0.;
    for (i = 0; i < width; ++i) acc += kernel[i];
    const double sum = 1. / acc;
    for (i = 0; i < width; ++i) kernel[i] = (float)(kernel[i] * sum);
    if (fft_length) {
        const int sh = width / 2;
        float *tmp = (float *)malloc(sizeof(float) * (size_t)(sh > 0 ? sh : 1));
        memcpy(tmp, kernel, sizeof(float) * (size_t)sh);
        memmove(kernel, kernel + sh, sizeof(float) * (size_t)(size - sh));
        memcpy(kernel + (size - sh), tmp, sizeof(float) * (size_t)sh);
        free(tmp);
    }
}

/* Utils.hpp:141-148 */
int orc_is_valid_size(int N)
{
    const int N_min = 32;
    int R = N;
    while (R >= 5 * N_min && (R % 5) == 0) R /= 5;
    while (R >= 3 * N_min && (R % 3) == 0) R /= 3;
    while (R >= 2 * N_min && (R % 2) == 0) R /= 2;
    return (R == N_min) ? 1 : 0;
}

/* Utils.hpp:150-157 */
int orc_nearest_transform_size(int N)
{
    const int N_min = 32;
    if (N < N_min) N = N_min;
    N = N_min * ((N + N_min - 1) / N_min);
    while (!orc_is_valid_size(N)) N += N_min;
    return N;
}

/* The reference's compile-time macro `boxblur` (Source.cpp:12) as a run-time switch: when on, the
 * pipelines below take the `#ifdef boxblur` branches (Source.cpp:152-156,207-211,224-228,287-291,
 * 325-329,338-342,437-441,468-472,491-495): a tent kernel (box convolved with itself) instead of
 * the Gaussian.  Not thread safe; the tests set it around single calls. */
static int g_boxblur = 0;
void orc_set_boxblur(int on) { g_boxblur = on ? 1 : 0; }

/* Source.cpp:128-140: 1-D box kernel convolved by itself, accumulated in float storage exactly like the
 * reference (`kernel[i] += double` rounds to float on every addition); the icol loop runs to kLen + 1
 * like the reference's (those terms clamp to 0).  kernel must be zeroed, fft_length floats. */
void orc_box_kernel_1d(float *kernel, int kLen, int fft_length)
{
    const double scale = 1. / pow((double)kLen, 4);
    for (int irow = -kLen + 1; irow <= kLen - 1; irow++)
        for (int icol = -kLen + 1; icol <= kLen + 1; icol++) {
            const double kval = (double)((kLen - abs(irow)) * (kLen - abs(icol)));
            double v = kval * scale;
            v = v < 0. ? 0. : (v > 1. ? 1. : v);
            float *k = &kernel[(icol + fft_length) % fft_length];
            *k = (float)(*k + v);
        }
}

/* nsmooth / kSize / pad of either build: Source.cpp:283-292 = :432-442 = :148-157 */
static void kernel_geometry(int rows, int cols, double *nsmooth, int *kSize, int *pad)
{
    *kSize = orc_gaussian_window(*nsmooth, rows > cols ? rows : cols);
    int passes = 1;
    if (g_boxblur) {
        int m = (int)*nsmooth * (int)*nsmooth;
        const int lim = (rows < cols ? rows : cols) - 1;
        if (lim < m) m = lim;
        *nsmooth = sqrt((double)m);
        passes = 2;
        *kSize = (int)(*nsmooth * *nsmooth);
    }
    *pad = (*kSize - 1) / 2 * passes;
}

/* taps of either build, centred at index 0 of an fft_length array (zeroed by the caller) */
static void kernel_taps(float *taps, double nsmooth, int kSize, int fft_length)
{
    if (g_boxblur) orc_box_kernel_1d(taps, (int)(nsmooth * nsmooth), fft_length);
    else orc_get_gaussian(taps, nsmooth, kSize, fft_length);
}

/* Source.cpp:434-457 (identical in pocketfft_1D :284-306) */
void orc_gauss_sizes(int rows, int cols, double sigma, int out[6])
{
    int kSize, pad;
    kernel_geometry(rows, cols, &sigma, &kSize, &pad);
    int sizes[2] = { rows + pad * 2, cols + pad * 2 };
    int tz[2] = { 0, 0 };
    for (int i = 0; i < 2; ++i) {
        if (!orc_is_valid_size(sizes[i])) {
            int ns = orc_nearest_transform_size(sizes[i]);
            tz[i] = ns - sizes[i];
            sizes[i] = ns;
        }
    }
    out[0] = kSize; out[1] = pad; out[2] = sizes[0]; out[3] = sizes[1]; out[4] = tz[0]; out[5] = tz[1];
}

/* ------------------------------------------------------- layout utilities */

/* Utils.hpp:159-184 with T=uint8_t, U=float: round = 0, plain conversion. */
void orc_deinterleave_u8_f32(const uint8_t *in, float *b, float *g, float *r, uint32_t total)
{
#pragma omp parallel for schedule(static)
    for (int64_t x = 0; x < (int64_t)total; ++x) {
        b[x] = in[x * 3 + 0];
        g[x] = in[x * 3 + 1];
        r[x] = in[x * 3 + 2];
    }
}

/* float -> uint8 as the reference's author's build behaves (Utils.hpp:204-206:
 * `T = U + 0.5f` is a truncating conversion with no clamp; negative inputs are UB
 * in ISO C++ and saturate to 0 on the author's ARM build per the DFT_image
 * fixtures).  Values >= 256 wrap modulo 256 like x86 cvttss2si+mov. */
static inline uint8_t f32_to_u8(float v)
{
    float t = v + 0.5f;
    if (!(t > 0.f)) return 0;
    return (uint8_t)(int)t;
}

/* Utils.hpp:186-210 with T=uint8_t, U=float */
void orc_interleave_f32_u8(const float *b, const float *g, const float *r, uint8_t *out, uint32_t total)
{
#pragma omp parallel for schedule(static)
    for (int64_t x = 0; x < (int64_t)total; ++x) {
        out[x * 3 + 0] = f32_to_u8(b[x]);
        out[x * 3 + 1] = f32_to_u8(g[x]);
        out[x * 3 + 2] = f32_to_u8(r[x]);
    }
}

/* Utils.hpp:212-243 */
void orc_reflect_101_u8(const uint8_t *in, uint8_t *out, int top, int bottom, int left, int right,
                        int rows, int cols, int C)
{
    if (top > rows - 1) top = rows - 1;         /* :217-220 */
    if (bottom > rows - 1) bottom = rows - 1;
    if (left > cols - 1) left = cols - 1;
    if (right > cols - 1) right = cols - 1;
    const int stride1 = cols * C;
    const int prow = rows + top + bottom;
    const int pcol = stride1 + (left + right) * C;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < prow; ++i) {
        uint8_t *row = out + (size_t)i * pcol;
        int src;
        if (i < prow - bottom) src = abs(i - top);      /* :231-232 */
        else src = 2 * (rows - 1) + top - i;            /* :233-234 */
        memcpy(row + left * C, in + (size_t)stride1 * src, (size_t)stride1);
        for (int j = 0; j < left; ++j)                  /* :236-237 */
            memcpy(row + j * C, row + (2 * left - j) * C, (size_t)C);
        for (int j = 0; j < right; ++j) {               /* :239-240 */
            int dst = left + cols + j;
            int s = 2 * (left + cols - 1) - dst;
            memcpy(row + dst * C, row + s * C, (size_t)C);
        }
    }
}

/* flip_block<float,1>(in, out, w, h): in is h x w row-major, out is w x h. */
void orc_flip_block_f32(const float *in, float *out, int w, int h)
{
    const int B = 32;
#pragma omp parallel for schedule(static)
    for (int y0 = 0; y0 < h; y0 += B)
        for (int x0 = 0; x0 < w; x0 += B) {
            int y1 = y0 + B < h ? y0 + B : h, x1 = x0 + B < w ? x0 + B : w;
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x)
                    out[(size_t)x * h + y] = in[(size_t)y * w + x];
        }
}

/* ------------------------------------------------- 1-D tile Gaussian path */

typedef struct {
    int n;                 /* FFT length (sizes[i]) */
    offt_plan *half;       /* complex plan of length n/2 */
    float *kre;            /* Re K[k], k in [0, n/2] */
} axis_plan;

/* kernel taps + spectrum: Source.cpp:465-503 / :322-346 */
static int axis_plan_init(axis_plan *ap, int n, double sigma, int kSize)
{
    ap->n = n;
    ap->half = offt_plan_create(n / 2);
    ap->kre = (float *)malloc(sizeof(float) * (size_t)(n / 2 + 1));
    float *taps = (float *)calloc((size_t)n, sizeof(float));
    ocpx *spec = (ocpx *)malloc(sizeof(ocpx) * (size_t)(n / 2 + 1));
    ocpx *scr = (ocpx *)malloc(sizeof(ocpx) * (size_t)(n / 2));
    if (!ap->half || !ap->kre || !taps || !spec || !scr) return 1;
    kernel_taps(taps, sigma, kSize, n);
    offt_r2c(ap->half, taps, spec, scr);
    for (int k = 0; k <= n / 2; ++k) ap->kre[k] = spec[k].re;
    free(taps); free(spec); free(scr);
    return 0;
}

static void axis_plan_free(axis_plan *ap)
{
    offt_plan_destroy(ap->half);
    free(ap->kre);
}

/* One 1-D pass over a plane src[R][L] -> dst[R][L] (SURVEY.md Appendix A.3;
 * Source.cpp:520-537 / :353-366).  nyquist_mode 1 restates
 * pffft_sorted_optimized_convolution (Source.cpp:414-427) on pffft's ordered
 * layout [F0, F(N/2), Re1, Im1, ...]: slot pair i=0 scales both DC and Nyquist
 * by kernel_dft[0]*scaler. */
static void pass_1d(const float *src, float *dst, int R, int L, int pad, const axis_plan *ap,
                    int nyquist_mode)
{
    const int N = ap->n, H = N / 2;
    const float scaler = 1.f / N;           /* Source.cpp:506-507 */
#pragma omp parallel
    {
        float *tile = (float *)malloc(sizeof(float) * (size_t)N);
        ocpx *work = (ocpx *)malloc(sizeof(ocpx) * (size_t)(H + 1));
        ocpx *scr = (ocpx *)malloc(sizeof(ocpx) * (size_t)H);
#pragma omp for schedule(static)
        for (int j = 0; j < R; ++j) {
            const float *row = src + (size_t)j * L;
            /* Source.cpp:525-529: left reflect | row | right reflect | trailing zeros */
            for (int k = 0; k < pad; ++k) tile[k] = row[pad - k];
            memcpy(tile + pad, row, sizeof(float) * (size_t)L);
            for (int k = 0; k < pad; ++k) tile[pad + L + k] = row[L - 2 - k];
            for (int k = L + 2 * pad; k < N; ++k) tile[k] = 0.f;

            offt_r2c(ap->half, tile, work, scr);
            if (nyquist_mode) {
                /* Source.cpp:420-426 */
                const float m0 = ap->kre[0] * scaler;
                work[0].re *= m0;
                work[H].re *= m0;       /* ordered slot 1 = F(N/2), scaled by kernel_dft[0] */
                for (int i = 1; i < H; ++i) {
                    const float m = ap->kre[i] * scaler;
                    work[i].re *= m;
                    work[i].im *= m;
                }
                offt_c2r(ap->half, work, tile, scr);
            } else {
                /* Source.cpp:362-363: all N/2+1 bins, 1/N applied by c2r's fct */
                for (int i = 0; i <= H; ++i) {
                    work[i].re *= ap->kre[i];
                    work[i].im *= ap->kre[i];
                }
                offt_c2r(ap->half, work, tile, scr);
                for (int k = 0; k < L; ++k) tile[pad + k] *= scaler;
            }
            memcpy(dst + (size_t)j * L, tile + pad, sizeof(float) * (size_t)L);   /* :536 */
        }
        free(tile); free(work); free(scr);
    }
}

int orc_gauss_blur_1d_f32(const uint8_t *bgr, int rows, int cols, double sigma, int nyquist_mode,
                          float *rowpass_out, float *final_out)
{
    int sz[6];
    orc_gauss_sizes(rows, cols, sigma, sz);
    const int kSize = sz[0], pad = sz[1];
    {   /* boxblur build: nsmooth itself is replaced (Source.cpp:288,438) before the taps are made */
        int ks, pd;
        kernel_geometry(rows, cols, &sigma, &ks, &pd);
    }
    if (pad > rows - 1 || pad > cols - 1 || pad < 0) return 2;   /* README.md:33-38 */
    const size_t nd = (size_t)rows * cols;
    float *planes = (float *)malloc(sizeof(float) * nd * 3);
    float *resf = (float *)malloc(sizeof(float) * nd);
    if (!planes || !resf) { free(planes); free(resf); return 1; }
    orc_deinterleave_u8_f32(bgr, planes, planes + nd, planes + 2 * nd, (uint32_t)nd);  /* :461 */

    axis_plan row_ap, col_ap;
    axis_plan_init(&row_ap, sz[3], sigma, kSize);
    axis_plan_init(&col_ap, sz[2], sigma, kSize);

    for (int c = 0; c < 3; ++c) {                                   /* :510 */
        float *p = planes + (size_t)c * nd;
        pass_1d(p, resf, rows, cols, pad, &row_ap, nyquist_mode);   /* :520-537 */
        if (rowpass_out) memcpy(rowpass_out + (size_t)c * nd, resf, sizeof(float) * nd);
        orc_flip_block_f32(resf, p, cols, rows);                    /* :540 */
        pass_1d(p, resf, cols, rows, pad, &col_ap, nyquist_mode);   /* :546-560 */
        orc_flip_block_f32(resf, p, rows, cols);                    /* :562 */
    }
    if (final_out) memcpy(final_out, planes, sizeof(float) * nd * 3);
    axis_plan_free(&row_ap);
    axis_plan_free(&col_ap);
    free(resf);
    free(planes);
    return 0;
}

int orc_gauss_blur_1d(uint8_t *bgr, int rows, int cols, double sigma, int nyquist_mode,
                      float *rowpass_out)
{
    const size_t nd = (size_t)rows * cols;
    float *fin = (float *)malloc(sizeof(float) * nd * 3);
    if (!fin) return 1;
    int rc = orc_gauss_blur_1d_f32(bgr, rows, cols, sigma, nyquist_mode, rowpass_out, fin);
    if (!rc) orc_interleave_f32_u8(fin, fin + nd, fin + 2 * nd, bgr, (uint32_t)nd);   /* :567 */
    free(fin);
    return rc;
}

/* ----------------------------------------------------------- 2-D path */

typedef struct {
    int border[4];      /* top bottom left right, Source.cpp:160-176 */
    int sizes[2];
    int kSize, pad;
    double nsmooth;        /* after the boxblur adjustment (:153) */
} geom2d;

static void geom2d_init(geom2d *g, int rows, int cols, double sigma)
{
    kernel_geometry(rows, cols, &sigma, &g->kSize, &g->pad);            /* :149-157 */
    g->nsmooth = sigma;
    for (int i = 0; i < 4; ++i) g->border[i] = g->pad;
    g->sizes[0] = rows + 2 * g->pad;
    g->sizes[1] = cols + 2 * g->pad;
    for (int i = 0; i < 2; ++i) {
        if (!orc_is_valid_size(g->sizes[i])) {
            int ns = orc_nearest_transform_size(g->sizes[i]);
            int np = ns - g->sizes[i];
            g->sizes[i] = ns;
            g->border[i * 2 + 0] += np / 2;                      /* :172 */
            g->border[i * 2 + 1] += (int)(np / 2.f + 0.5f);      /* :173 */
        }
    }
}

/* 2-D r2c of a real H x W plane into H x (W/2+1): rows r2c, then columns c2c
 * (what pocketfft::r2c does for axes {0,1}: real transform on the last axis,
 * complex transforms on the others; call site Source.cpp:238). */
static void fft2_r2c(const float *plane, ocpx *spec, int H, int W, const offt_plan *half_w,
                     const offt_plan *full_h)
{
    const int WC = W / 2 + 1;
#pragma omp parallel
    {
        ocpx *scr = (ocpx *)malloc(sizeof(ocpx) * (size_t)(W / 2));
#pragma omp for schedule(static)
        for (int i = 0; i < H; ++i) offt_r2c(half_w, plane + (size_t)i * W, spec + (size_t)i * WC, scr);
        free(scr);
    }
#pragma omp parallel
    {
        ocpx *cin = (ocpx *)malloc(sizeof(ocpx) * (size_t)H);
        ocpx *cout = (ocpx *)malloc(sizeof(ocpx) * (size_t)H);
#pragma omp for schedule(static)
        for (int j = 0; j < WC; ++j) {
            for (int i = 0; i < H; ++i) cin[i] = spec[(size_t)i * WC + j];
            offt_c2c(full_h, cin, cout, 0);
            for (int i = 0; i < H; ++i) spec[(size_t)i * WC + j] = cout[i];
        }
        free(cin); free(cout);
    }
}

static void fft2_c2r(ocpx *spec, float *plane, int H, int W, const offt_plan *half_w,
                     const offt_plan *full_h, float fct)
{
    const int WC = W / 2 + 1;
#pragma omp parallel
    {
        ocpx *cin = (ocpx *)malloc(sizeof(ocpx) * (size_t)H);
        ocpx *cout = (ocpx *)malloc(sizeof(ocpx) * (size_t)H);
#pragma omp for schedule(static)
        for (int j = 0; j < WC; ++j) {
            for (int i = 0; i < H; ++i) cin[i] = spec[(size_t)i * WC + j];
            offt_c2c(full_h, cin, cout, 1);
            for (int i = 0; i < H; ++i) spec[(size_t)i * WC + j] = cout[i];
        }
        free(cin); free(cout);
    }
#pragma omp parallel
    {
        ocpx *scr = (ocpx *)malloc(sizeof(ocpx) * (size_t)(W / 2));
#pragma omp for schedule(static)
        for (int i = 0; i < H; ++i) {
            float *row = plane + (size_t)i * W;
            offt_c2r(half_w, spec + (size_t)i * WC, row, scr);
            for (int j = 0; j < W; ++j) row[j] *= fct;
        }
        free(scr);
    }
}

static int run_2d(const uint8_t *bgr, uint8_t *out_u8, int rows, int cols, double sigma,
                  int dft_image, float *spec_f32, int padded_sizes[2])
{
    geom2d g;
    geom2d_init(&g, rows, cols, sigma);
    if (g.pad < 0) return 2;
    const int H = g.sizes[0], W = g.sizes[1], WC = W / 2 + 1;
    /* Reflect_101 silently clamps each border to dim-1 (Utils.hpp:217-220), which would
     * desynchronise the caller's geometry; the restatement refuses such inputs. */
    if (g.border[0] > rows - 1 || g.border[1] > rows - 1 || g.border[2] > cols - 1 ||
        g.border[3] > cols - 1)
        return 2;
    if (padded_sizes) { padded_sizes[0] = H; padded_sizes[1] = W; }
    const size_t np = (size_t)H * W;
    uint8_t *padded = (uint8_t *)malloc(np * 3);
    float *planes = (float *)malloc(sizeof(float) * np * 3);
    if (!padded || !planes) { free(padded); free(planes); return 1; }
    orc_reflect_101_u8(bgr, padded, g.border[0], g.border[1], g.border[2], g.border[3], rows, cols, 3); /* :181 */
    orc_deinterleave_u8_f32(padded, planes, planes + np, planes + 2 * np, (uint32_t)np);               /* :188 */

    offt_plan *half_w = offt_plan_create(W / 2);
    offt_plan *full_h = offt_plan_create(H);
    offt_plan *half_h = offt_plan_create(H / 2);

    /* kernel spectra, Source.cpp:204-232 (column spectrum mirrored to full length, :218) */
    float *kcol = (float *)malloc(sizeof(float) * (size_t)H);
    float *krow = (float *)malloc(sizeof(float) * (size_t)WC);
    {
        float *taps = (float *)calloc((size_t)(H > W ? H : W), sizeof(float));
        ocpx *spec = (ocpx *)malloc(sizeof(ocpx) * (size_t)((H > W ? H : W) / 2 + 1));
        ocpx *scr = (ocpx *)malloc(sizeof(ocpx) * (size_t)((H > W ? H : W) / 2));
        kernel_taps(taps, g.nsmooth, g.kSize, H);
        offt_r2c(half_h, taps, spec, scr);
        for (int k = 0; k <= H / 2; ++k) kcol[k] = spec[k].re;
        for (int k = H / 2 + 1; k < H; ++k) kcol[k] = kcol[H - k];
        memset(taps, 0, sizeof(float) * (size_t)(H > W ? H : W));
        kernel_taps(taps, g.nsmooth, g.kSize, W);
        offt_r2c(half_w, taps, spec, scr);
        for (int k = 0; k < WC; ++k) krow[k] = spec[k].re;
        free(taps); free(spec); free(scr);
    }

    ocpx *resf = (ocpx *)malloc(sizeof(ocpx) * (size_t)H * WC);
    for (int c = 0; c < 3; ++c) {                         /* hybrid_loop(3,...) :236 */
        float *p = planes + (size_t)c * np;
        fft2_r2c(p, resf, H, W, half_w, full_h);          /* :238 */
        if (dft_image) {
            /* Source.cpp:240-252 */
#pragma omp parallel for schedule(static)
            for (int row = 0; row < H; ++row)
                for (int col = 0; col < W; ++col) {
                    int row_ = (row + (H % 2 == 0 ? H : (H + 1)) / 2) % H;
                    int col_ = (col + (W % 2 == 0 ? W : (W + 1)) / 2) % W;
                    int cval = col_ < (W / 2 + 1) ? col_ : ((W / 2) - col_ % (W / 2));
                    p[(size_t)row * W + col] =
                        20 * log10f(fabsf(resf[(size_t)row_ * WC + cval].re) + 0.00001f);
                }
        } else {
            /* Source.cpp:255-263 */
#pragma omp parallel for schedule(static)
            for (int i = 0; i < H; ++i)
                for (int j = 0; j < WC; ++j) {
                    float m = krow[j] * kcol[i];
                    resf[(size_t)i * WC + j].re *= m;
                    resf[(size_t)i * WC + j].im *= m;
                }
            fft2_c2r(resf, p, H, W, half_w, full_h, 1.f / (float)((int)np));
        }
    }
    if (spec_f32) memcpy(spec_f32, planes, sizeof(float) * np * 3);
    orc_interleave_f32_u8(planes, planes + np, planes + 2 * np, padded, (uint32_t)np);   /* :268 */
    /* crop, Source.cpp:271-276 */
    for (int i = 0; i < H - g.border[1] - g.border[0]; ++i)
        memcpy(out_u8 + (size_t)i * cols * 3,
               padded + ((size_t)(i + g.border[0]) * W + g.border[2]) * 3, (size_t)(W - g.border[3] - g.border[2]) * 3);

    free(resf); free(kcol); free(krow);
    offt_plan_destroy(half_w); offt_plan_destroy(full_h); offt_plan_destroy(half_h);
    free(padded); free(planes);
    return 0;
}

int orc_gauss_blur_2d(uint8_t *bgr, int rows, int cols, double sigma)
{
    uint8_t *tmp = (uint8_t *)malloc((size_t)rows * cols * 3);
    if (!tmp) return 1;
    int rc = run_2d(bgr, tmp, rows, cols, sigma, 0, NULL, NULL);
    if (!rc) memcpy(bgr, tmp, (size_t)rows * cols * 3);
    free(tmp);
    return rc;
}

int orc_dft_image(const uint8_t *bgr, uint8_t *out_u8, int rows, int cols, double nsmooth,
                  float *spec_f32, int padded_sizes[2])
{
    return run_2d(bgr, out_u8, rows, cols, nsmooth, 1, spec_f32, padded_sizes);
}

/* ----------------------------------------------------------- box blur */

static inline int refl101(int i, int L)
{
    if (i < 0) i = -i;
    if (i >= L) i = 2 * (L - 1) - i;
    return i;
}

/* one horizontal sliding-accumulator pass over an h x w x C image (declared spec) */
static void box_h(const uint8_t *in, uint8_t *out, int w, int h, int C, int r)
{
    const uint32_t k = 2u * (uint32_t)r + 1u;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y) {
        const uint8_t *row = in + (size_t)y * w * C;
        uint8_t *orow = out + (size_t)y * w * C;
        for (int c = 0; c < C; ++c) {
            uint32_t acc = 0;
            for (int i = -r; i <= r; ++i) acc += row[refl101(i, w) * C + c];
            for (int x = 0; x < w; ++x) {
                orow[x * C + c] = (uint8_t)((2u * acc + k) / (2u * k));
                acc += row[refl101(x + r + 1, w) * C + c];
                acc -= row[refl101(x - r, w) * C + c];
            }
        }
    }
}

static void transpose_u8(const uint8_t *in, uint8_t *out, int w, int h, int C)
{
    const int B = 32;
#pragma omp parallel for schedule(static)
    for (int y0 = 0; y0 < h; y0 += B)
        for (int x0 = 0; x0 < w; x0 += B) {
            int y1 = y0 + B < h ? y0 + B : h, x1 = x0 + B < w ? x0 + B : w;
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x)
                    memcpy(out + ((size_t)x * h + y) * C, in + ((size_t)y * w + x) * C, (size_t)C);
        }
}

int orc_fastboxblur(uint8_t *img, int w, int h, int C, int ksize, int passes)
{
    if (ksize < 1 || C < 1 || w < 1 || h < 1 || passes < 0) return 2;
    const int r = (ksize - 1) / 2;
    if (r > w - 1 || r > h - 1) return 2;
    const size_t nb = (size_t)w * h * C;
    uint8_t *a = (uint8_t *)malloc(nb), *b = (uint8_t *)malloc(nb);
    if (!a || !b) { free(a); free(b); return 1; }
    for (int p = 0; p < passes; ++p) {
        box_h(img, a, w, h, C, r);          /* horizontal */
        transpose_u8(a, b, w, h, C);        /* flip_block */
        box_h(b, a, h, w, C, r);            /* vertical as horizontal on the transpose */
        transpose_u8(a, img, h, w, C);      /* flip back */
    }
    free(a); free(b);
    return 0;
}
