// fft_plan.hpp — host-side planning: window/tap semantics of the reference, FFT factorisation,
// twiddle and kernel-spectrum tables.  Pure C++ (no CUDA) so the CPU emulator can use it too.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "fft_core.cuh"

namespace blurb200 {

// ---- sizing (mirrors the integer/float semantics of the reference exactly) ----------------

// Reference: Source.cpp:60-73.  radius is float(double expr); width is float math truncated;
// clamp to max_width happens BEFORE the make-odd step.
inline int gaussian_window(double sigma, int max_width = 0)
{
    const float radius = static_cast<float>(sigma * std::sqrt(2.0 * std::log(255.0)) - 1.0);
    int width = static_cast<int>(radius * 2.f + .5f);
    if (max_width) width = std::min(width, max_width);
    if (width % 2 == 0) ++width;
    return width;
}

// Reference: Utils.hpp:141-148 (pffft's rule: N = 32 * 2^a 3^b 5^c).
inline int is_valid_size(int n)
{
    int r = n;
    while (r >= 160 && r % 5 == 0) r /= 5;
    while (r >= 96 && r % 3 == 0) r /= 3;
    while (r >= 64 && r % 2 == 0) r /= 2;
    return r == 32 ? 1 : 0;
}

// Reference: Utils.hpp:150-157.
inline int nearest_transform_size(int n)
{
    if (n < 32) n = 32;
    n = 32 * ((n + 31) / 32);
    while (!is_valid_size(n)) n += 32;
    return n;
}

// ---- kernel kinds ----------------------------------------------------------------------------
// Throughout the planning code a NEGATIVE `sigma` selects the FFT box blur of the reference's
// `#define boxblur` build (Source.cpp:12; box_kernel :128-140; branches :287-291,437-441,468-495):
// sigma = -kLen, taps = a box of width kLen convolved with itself (tent of 2 kLen - 1 taps), pad = kLen - 1.
// It flows through get_gaussian / gauss_geometry / kernel_spectrum like a Gaussian's (sigma, width) pair.
inline bool is_box_sigma(double sigma) { return sigma < 0; }
inline int box_klen(double sigma) { return static_cast<int>(-sigma + 0.5); }

// Source.cpp:287-291 = :437-441: the boxblur build replaces nsmooth by sqrt(min((int)ns^2, min(rows-1, cols-1)))
// and takes kSize = kLen = (int)(nsmooth'^2) (double arithmetic, truncation).
inline int boxfft_klen(int rows, int cols, double nsmooth)
{
    const int m = std::min(static_cast<int>(nsmooth) * static_cast<int>(nsmooth), std::min(rows - 1, cols - 1));
    if (m < 0) return 0;
    const double ns = std::sqrt(static_cast<double>(m));
    return static_cast<int>(ns * ns);
}

// Source.cpp:128-140, float storage accumulated exactly like the reference (`kernel[i] += double`
// rounds on every addition); the icol loop runs to kLen + 1 as there (those terms clamp to 0).
inline void box_kernel_1d(std::vector<float>& kernel, int kLen, int fft_length)
{
    kernel.assign(static_cast<size_t>(fft_length), 0.f);
    const double scale = 1.0 / std::pow(static_cast<double>(kLen), 4);
    for (int irow = -kLen + 1; irow <= kLen - 1; ++irow)
        for (int icol = -kLen + 1; icol <= kLen + 1; ++icol) {
            const double kval = static_cast<double>((kLen - std::abs(irow)) * (kLen - std::abs(icol)));
            const double v = std::min(std::max(kval * scale, 0.0), 1.0);
            float& k = kernel[static_cast<size_t>((icol + fft_length) % fft_length)];
            k = static_cast<float>(k + v);
        }
}

// Reference: Source.cpp:75-102.  float loop variable, float y*y, double exp/normalise, float
// storage, then rotate left by width/2 so the centre tap is element 0.
inline void get_gaussian(std::vector<float>& kernel, double sigma, int width = 0, int fft_length = 0)
{
    if (is_box_sigma(sigma)) {          // tent taps of the boxblur build, centre at element 0 (fft_length > 2 kLen + 1)
        box_kernel_1d(kernel, box_klen(sigma), fft_length ? fft_length : 2 * box_klen(sigma) + 3);
        return;
    }
    if (!width) width = gaussian_window(sigma);
    kernel.assign(static_cast<size_t>(fft_length ? fft_length : width), 0.f);
    const float mid_w = (width - 1) / 2.f;
    const double s = 2.0 * sigma * sigma;
    int i = 0;
    for (float y = -mid_w; y <= mid_w; ++y, ++i)
        kernel[i] = static_cast<float>(std::exp(-static_cast<double>(y * y) / s) / (3.14159265358979323846 * s));
    double acc = 0.0;
    for (i = 0; i < width; ++i) acc += kernel[i];
    const double inv = 1.0 / acc;
    for (i = 0; i < width; ++i) kernel[i] = static_cast<float>(kernel[i] * inv);
    if (fft_length) std::rotate(kernel.begin(), kernel.begin() + width / 2, kernel.end());
}

// Geometry of the 1-D tile pipeline, Source.cpp:434-457.
struct GaussGeometry {
    int rows, cols, ksize, pad;
    int n_row;   // sizes[1]: FFT length of the row pass
    int n_col;   // sizes[0]: FFT length of the column pass
};

inline GaussGeometry gauss_geometry(int rows, int cols, double sigma)
{
    GaussGeometry g{};
    g.rows = rows; g.cols = cols;
    g.ksize = is_box_sigma(sigma) ? 2 * box_klen(sigma) - 1 : gaussian_window(sigma, std::max(rows, cols));
    g.pad = (g.ksize - 1) / 2;          // box: kLen - 1 = (kSize - 1) / 2 * passes of Source.cpp:292 for odd kLen
    g.n_col = rows + 2 * g.pad;
    g.n_row = cols + 2 * g.pad;
    if (!is_valid_size(g.n_col)) g.n_col = nearest_transform_size(g.n_col);
    if (!is_valid_size(g.n_row)) g.n_row = nearest_transform_size(g.n_row);
    return g;
}

// ---- FFT factorisation ---------------------------------------------------------------------

static const int kAllowedRadix[] = {2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16, 18, 20, 24, 25, 27, 30, 32};

inline bool split_radices(int n, int k, int min_r, std::vector<int>& cur, std::vector<int>& best)
{
    // enumerate non-decreasing factorizations of n into exactly k allowed radices, keep the
    // one with the smallest maximum factor
    if (k == 1) {
        for (int r : kAllowedRadix)
            if (r == n && r >= min_r) {
                cur.push_back(r);
                if (best.empty() || cur.back() < best.back()) best = cur;
                cur.pop_back();
                return true;
            }
        return false;
    }
    bool any = false;
    for (int r : kAllowedRadix) {
        if (r < min_r || n % r) continue;
        cur.push_back(r);
        any |= split_radices(n / r, k - 1, r, cur, best);
        cur.pop_back();
    }
    return any;
}

inline uint64_t magic40(int d) { return ((1ull << 40) + static_cast<uint64_t>(d) - 1) / static_cast<uint64_t>(d); }

// Returns false if N is not a valid size or cannot be planned within kMaxPass passes.
// `override_radices` (optional, e.g. from BLUR_B200_PLAN) lists the strided radices.
inline bool make_geom(int N, FftGeom& g, const std::vector<int>* override_radices = nullptr)
{
    std::memset(&g, 0, sizeof(g));
    if (N < 32 || !is_valid_size(N)) return false;
    const int rest = N / kLastRadix;
    std::vector<int> rad;
    if (override_radices && !override_radices->empty()) {
        int prod = 1;
        for (int r : *override_radices) prod *= r;
        if (prod != rest) return false;
        rad = *override_radices;
    } else if (rest == 1) {
        rad = {1};
    } else {
        for (int k = 1; k < kMaxPass && rad.empty(); ++k) {
            std::vector<int> cur, best;
            if (split_radices(rest, k, 2, cur, best)) rad = best;
        }
        if (rad.empty()) return false;
    }
    if (static_cast<int>(rad.size()) + 1 > kMaxPass) return false;
    g.N = N;
    g.npass = static_cast<int>(rad.size()) + 1;
    int prod = 1;
    for (int i = 0; i < g.npass; ++i) {
        g.radix[i] = (i + 1 == g.npass) ? kLastRadix : rad[i];
        prod *= g.radix[i];
        g.M[i] = N / prod;
        g.W[i] = N / g.radix[i];
        g.magicM[i] = magic40(g.M[i]);
        g.magicW[i] = magic40(g.W[i]);
    }
    g.slots = N + N / 32;
    return true;
}

// Launch shape: B transforms per CTA, T threads.
struct LaunchShape {
    int batch;
    int threads;
    size_t smem_bytes;
    int ctas_per_sm;
};

inline LaunchShape choose_launch(const FftGeom& g, int max_threads = 576, size_t smem_per_sm = 227 * 1024)
{
    LaunchShape s{};
    const size_t per = static_cast<size_t>(g.slots) * 8;
    const int wlast = g.N / kLastRadix;
    int B = (256 + wlast - 1) / wlast;
    B = std::max(1, std::min(B, 8));
    while (B > 1 && per * B > 80 * 1024) --B;
    s.batch = B;
    s.smem_bytes = per * B;
    s.ctas_per_sm = static_cast<int>(std::max<size_t>(1, std::min<size_t>(8, smem_per_sm / (s.smem_bytes + 1024))));
    const int tmin = s.ctas_per_sm == 1 ? 384 : 192;
    const int tmax = s.ctas_per_sm == 1 ? max_threads : 384;
    double best = 1e300;
    for (int T = tmin; T <= tmax; T += 32) {
        double cost = 0;
        for (int i = 0; i < g.npass; ++i) {
            const double per_item = (i + 1 == g.npass) ? 2.0 * 32 * 5 : g.radix[i] * std::log2(std::max(2, g.radix[i]));
            const long items = static_cast<long>(B) * g.W[i];
            cost += static_cast<double>((items + T - 1) / T) * T * per_item * (i + 1 == g.npass ? 1 : 2);
        }
        // mild preference for more resident warps at equal padded work
        cost *= 1.0 + 0.02 * (tmax - T) / 32.0;
        if (cost < best) { best = cost; s.threads = T; }
    }
    if (const char* e = std::getenv("BLUR_B200_THREADS")) {
        int T = std::atoi(e);
        if (T >= 32 && T <= max_threads && T % 32 == 0) s.threads = T;
    }
    return s;
}

// ---- tables --------------------------------------------------------------------------------

inline void make_twiddles(int N, std::vector<float2>& tw)
{
    tw.resize(static_cast<size_t>(N));
    for (int k = 0; k < N; ++k) {
        const double a = -2.0 * 3.14159265358979323846 * static_cast<double>(k) / static_cast<double>(N);
        tw[k] = make_float2(static_cast<float>(std::cos(a)), static_cast<float>(std::sin(a)));
    }
}

// Re K[k] for k in [0, N): DFT (double accumulation) of the rotated float taps of
// get_gaussian(.., width, N).  Source.cpp:465-503 computes the same spectrum with a float32
// FFT; only the real part is ever used (:362,:423).
inline void kernel_spectrum(double sigma, int width, int N, std::vector<double>& K)
{
    std::vector<float> taps;
    get_gaussian(taps, sigma, width, N);
    std::vector<double> cs(static_cast<size_t>(N));
    for (int m = 0; m < N; ++m) cs[m] = std::cos(2.0 * 3.14159265358979323846 * m / N);
    std::vector<std::pair<int, double>> nz;
    for (int n = 0; n < N; ++n)
        if (taps[n] != 0.f) nz.emplace_back(n, static_cast<double>(taps[n]));
    K.assign(static_cast<size_t>(N), 0.0);
    for (int k = 0; k <= N / 2; ++k) {
        double acc = 0.0;
        for (const auto& t : nz) acc += t.second * cs[static_cast<size_t>((static_cast<int64_t>(k) * t.first) % N)];
        K[k] = acc;
        if (k) K[N - k] = acc;
    }
}

// Scrambled + transposed multiplier table for mid_pass: kt[q * (N/32) + u] multiplies logical
// position p = 32u + q.  1/N is folded in as a float product like Source.cpp:423 / :506-507.
// nyquist_mode 1 reproduces pffft_sorted_optimized_convolution (Source.cpp:414-427): in pffft's
// ordered layout slot pair 0 holds (F0, F(N/2)), so bin N/2 is scaled by K[0], not K[N/2].
inline void make_kt(const FftGeom& g, const std::vector<double>& K, int nyquist_mode, std::vector<float>& kt)
{
    const int N = g.N, W = N / kLastRadix;
    const float scaler = 1.f / static_cast<float>(N);
    kt.resize(static_cast<size_t>(N));
    for (int p = 0; p < N; ++p) {
        int k = scrambled_freq(g, p);
        if (nyquist_mode == 1 && k == N / 2) k = 0;
        const int u = p / kLastRadix, q = p % kLastRadix;
        kt[static_cast<size_t>(q) * W + u] = static_cast<float>(K[k]) * scaler;
    }
}

}  // namespace blurb200
