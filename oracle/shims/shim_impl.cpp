/*
 * oracle/shims/shim_impl.cpp — TEST INFRASTRUCTURE.
 *
 * Backs the pffft / pocketfft API stand-ins (see the shim headers) with oracle/cpu_fft.c so
 * that the reference's Source.cpp can be compiled UNMODIFIED into oracle/_ref.
 */
#include "pffft_pommier/pffft.h"
#include "pocketfft/pocketfft_hdronly.h"
#include "../cpu_fft.h"

#include <map>
#include <mutex>

namespace {
std::mutex g_mu;
std::map<int, offt_plan *> g_plans;   // like pocketfft's internal plan cache

const offt_plan *get_plan(int n)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_plans.find(n);
    if (it != g_plans.end()) return it->second;
    offt_plan *p = offt_plan_create(n);
    g_plans[n] = p;
    return p;
}

struct Scratch {
    std::vector<ocpx> a, b;
    void need(size_t n) { if (a.size() < n) { a.resize(n); b.resize(n); } }
};
thread_local Scratch tls;
}  // namespace

struct PFFFT_Setup {
    int N;
    const offt_plan *half;
};

extern "C" PFFFT_Setup *pffft_new_setup(int N, pffft_transform_t transform)
{
    if (transform != PFFFT_REAL || N % 2) return nullptr;
    return new PFFFT_Setup{N, get_plan(N / 2)};
}

extern "C" void pffft_destroy_setup(PFFFT_Setup *s) { delete s; }

extern "C" void pffft_transform_ordered(PFFFT_Setup *s, const float *input, float *output, float * /*work*/,
                                        pffft_direction_t direction)
{
    const int N = s->N, H = N / 2;
    tls.need((size_t)H + 1);
    ocpx *spec = tls.a.data(), *scr = tls.b.data();
    if (direction == PFFFT_FORWARD) {
        offt_r2c(s->half, input, spec, scr);
        output[0] = spec[0].re;
        output[1] = spec[H].re;
        for (int k = 1; k < H; ++k) { output[2 * k] = spec[k].re; output[2 * k + 1] = spec[k].im; }
    } else {
        spec[0].re = input[0]; spec[0].im = 0.f;
        spec[H].re = input[1]; spec[H].im = 0.f;
        for (int k = 1; k < H; ++k) { spec[k].re = input[2 * k]; spec[k].im = input[2 * k + 1]; }
        offt_c2r(s->half, spec, output, scr);
    }
}

namespace pocketfft {

static void check_contig(const shape_t &shape, const stride_t &sr, const stride_t &sc, bool ok_axes)
{
    bool ok = ok_axes;
    if (shape.size() == 1) ok = ok && sr[0] == (ptrdiff_t)sizeof(float) && sc[0] == (ptrdiff_t)sizeof(std::complex<float>);
    else if (shape.size() == 2)
        ok = ok && sr[1] == (ptrdiff_t)sizeof(float) && sr[0] == (ptrdiff_t)(sizeof(float) * shape[1]) &&
             sc[1] == (ptrdiff_t)sizeof(std::complex<float>) &&
             sc[0] == (ptrdiff_t)(sizeof(std::complex<float>) * (shape[1] / 2 + 1));
    else ok = false;
    if (!ok) { std::fprintf(stderr, "pocketfft shim: unsupported layout\n"); std::abort(); }
}

void shim_r2c_f32(const shape_t &shape, const stride_t &stride_in, const stride_t &stride_out,
                  const shape_t &axes, bool forward, const float *in, std::complex<float> *out_, float fct)
{
    bool ax_ok = forward && ((axes.size() == 1 && axes[0] == 0 && shape.size() == 1) ||
                             (axes.size() == 2 && axes[0] == 0 && axes[1] == 1 && shape.size() == 2));
    check_contig(shape, stride_in, stride_out, ax_ok);
    ocpx *out = reinterpret_cast<ocpx *>(out_);
    if (shape.size() == 1) {
        const int N = (int)shape[0];
        tls.need((size_t)N / 2 + 1);
        offt_r2c(get_plan(N / 2), in, out, tls.b.data());
        if (fct != 1.f) for (int k = 0; k <= N / 2; ++k) { out[k].re *= fct; out[k].im *= fct; }
        return;
    }
    const int H = (int)shape[0], W = (int)shape[1], WC = W / 2 + 1;
    const offt_plan *half_w = get_plan(W / 2), *full_h = get_plan(H);
#pragma omp parallel
    {
        tls.need((size_t)(H > W ? H : W) + 1);
#pragma omp for schedule(static)
        for (int i = 0; i < H; ++i) offt_r2c(half_w, in + (size_t)i * W, out + (size_t)i * WC, tls.b.data());
#pragma omp for schedule(static)
        for (int j = 0; j < WC; ++j) {
            ocpx *cin = tls.a.data(), *cout = tls.b.data();
            for (int i = 0; i < H; ++i) cin[i] = out[(size_t)i * WC + j];
            offt_c2c(full_h, cin, cout, 0);
            for (int i = 0; i < H; ++i) { out[(size_t)i * WC + j].re = cout[i].re * fct; out[(size_t)i * WC + j].im = cout[i].im * fct; }
        }
    }
}

void shim_c2r_f32(const shape_t &shape, const stride_t &stride_in, const stride_t &stride_out,
                  const shape_t &axes, bool forward, const std::complex<float> *in_, float *out, float fct)
{
    bool ax_ok = !forward && ((axes.size() == 1 && axes[0] == 0 && shape.size() == 1) ||
                              (axes.size() == 2 && axes[0] == 0 && axes[1] == 1 && shape.size() == 2));
    check_contig(shape, stride_out, stride_in, ax_ok);
    const ocpx *in = reinterpret_cast<const ocpx *>(in_);
    if (shape.size() == 1) {
        const int N = (int)shape[0];
        tls.need((size_t)N / 2 + 1);
        offt_c2r(get_plan(N / 2), in, out, tls.b.data());
        if (fct != 1.f) for (int k = 0; k < N; ++k) out[k] *= fct;
        return;
    }
    const int H = (int)shape[0], W = (int)shape[1], WC = W / 2 + 1;
    const offt_plan *half_w = get_plan(W / 2), *full_h = get_plan(H);
    std::vector<ocpx> tmp((size_t)H * WC);
#pragma omp parallel
    {
        tls.need((size_t)(H > W ? H : W) + 1);
#pragma omp for schedule(static)
        for (int j = 0; j < WC; ++j) {
            ocpx *cin = tls.a.data(), *cout = tls.b.data();
            for (int i = 0; i < H; ++i) cin[i] = in[(size_t)i * WC + j];
            offt_c2c(full_h, cin, cout, 1);
            for (int i = 0; i < H; ++i) tmp[(size_t)i * WC + j] = cout[i];
        }
#pragma omp for schedule(static)
        for (int i = 0; i < H; ++i) {
            float *row = out + (size_t)i * W;
            offt_c2r(half_w, tmp.data() + (size_t)i * WC, row, tls.b.data());
            for (int j = 0; j < W; ++j) row[j] *= fct;
        }
    }
}
}  // namespace pocketfft
