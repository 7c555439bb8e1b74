// ols_plan.hpp — host-side planning of the overlap-save passes: block length choice and the
// twiddle / kernel-spectrum tables.  Pure C++ (the CPU emulator uses it too).
#pragma once
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "fft_plan.hpp"
#include "ols_core.cuh"

namespace blurb200 {

static const int kOlsRadices[] = {4, 6, 8, 12, 16, 24, 32};

// Picks the block length with the least total FFT work for lines of `len` samples and a
// reflect pad of `pad`.  `align` = granularity of a block's first sample; max_r0 caps the block
// length (the column pass keeps four blocks in shared memory at once).  Returns false if no block length leaves at least a quarter
// of each block as valid output (very wide kernels: the caller uses the full-length path).
inline bool choose_ols_axis(int len, int pad, int align, int max_r0, OlsAxis& ax, int force_r0 = 0)
{
    ax = OlsAxis{};
    double best = 1e300;
    for (int r0 : kOlsRadices) {
        if ((force_r0 && r0 != force_r0) || r0 > max_r0) continue;
        const int L = 256 * r0;
        const int halo = ((pad + align - 1) / align) * align;
        const int V = L - 2 * halo;
        if (V < L / 4 && !force_r0) continue;
        if (V <= 0) continue;
        const int nb = (len + V - 1) / V;
        // work ~ samples transformed x (log2 L + constant for the passes' fixed costs)
        const double cost = static_cast<double>(nb) * L * (std::log2(static_cast<double>(L)) + 6.0);
        if (cost < best) {
            best = cost;
            ax.R0 = r0; ax.L = L; ax.halo = halo; ax.V = V; ax.nblocks = nb; ax.len = len; ax.pad = pad;
        }
    }
    return ax.R0 != 0;
}

// Two-factor kernels (ols2_kernels.cuh): L = 32 * R2 with R2 = 32, halo = pad exactly.  Chosen when at
// least `min_frac` of a block is valid output (narrow and medium kernels: sigma <= ~65); wider
// kernels keep the three-factor kernels and their longer blocks.
// `even_halo`: round the halo up to an even number so that V is a multiple of 4 and every block's
// first output is 16-byte aligned in Q (row axis: the epilogue leaves through bulk shared->global copies).
inline bool choose_ols2_axis(int len, int pad, OlsAxis& ax, bool even_halo = false, double min_frac = 0.55)
{
    const int R2 = 32, L = 32 * R2;
    const int halo = even_halo ? pad + (pad & 1) : pad;
    const int V = L - 2 * halo;
    if (V < static_cast<int>(min_frac * L)) return false;
    ax = OlsAxis{};
    ax.R0 = L / 256; ax.L = L; ax.halo = halo; ax.V = V; ax.nblocks = (len + V - 1) / V; ax.len = len; ax.pad = pad; ax.R2 = R2;
    return true;
}

// How one image is cut over `nranks` GPUs (nranks = 1: the whole image) and which algorithm
// runs: the single source of truth for blur_b200.cu, the CPU emulator and dist.py.
struct ShardGeom {
    int algo;            // 0 = one full-length FFT per line (fft_kernels.cuh), 1 = overlap-save
    OlsAxis ols_row, ols_col;
    int stripe_rows;     // image rows per rank in the row pass (capacity; the last rank may own fewer)
    int cpp;             // column pairs per rank in the column pass
    int qpitch, cpart;   // overlap-save Q: floats per Q row / image columns per part (= 2*cpp)
    int row0, valid_rows, valid_ncp;
    size_t block_floats; // floats per (stripe, part) block of Q = 3 * stripe_rows * 2*cpp
};

inline ShardGeom shard_geometry(const GaussGeometry& gg, int nyquist_mode, int rank, int nranks)
{
    ShardGeom g{};
    const int rows = gg.rows, cols = gg.cols;
    // Overlap-save unless the kernel is too wide for the block menu.  The Nyquist quirk of pffft
    // (defined on the reference's full-length tile) is a rank-one correction per line that the
    // overlap-save epilogues add; in a sharded plan the caller sums the two small correction
    // buffers over the ranks (blur_b200_gauss_quirk_local / _finish).  BLUR_B200_QUIRK_SHARDED=full
    // keeps round 1's full-length kernel for sharded pffft-mode plans.
    int want_ols = 1;
    if (nyquist_mode != 0 && nranks > 1) {
        const char* q = std::getenv("BLUR_B200_QUIRK_SHARDED");
        if (q && !std::strcmp(q, "full")) want_ols = 0;
    }
    if (const char* e = std::getenv("BLUR_B200_ALGO")) {
        if (!std::strcmp(e, "full")) want_ols = 0;
    }
    int force_r0 = 0;
    if (const char* e = std::getenv("BLUR_B200_OLS_R0")) force_r0 = std::atoi(e);
    if (want_ols && choose_ols_axis(cols, gg.pad, 1, 32, g.ols_row, force_r0) &&
        choose_ols_axis(rows, gg.pad, 1, 24, g.ols_col, force_r0))
        g.algo = 1;
    // two-factor kernels where the kernel is narrow enough for their 1024-sample blocks
    // (BLUR_B200_OLS2=0: keep the three-factor kernels; =row / =col: one axis only — A/B measurements)
    if (g.algo == 1 && !force_r0) {
        const char* e = std::getenv("BLUR_B200_OLS2");
        const bool want_row = !e || !std::strcmp(e, "1") || !std::strcmp(e, "row");
        const bool want_col = !e || !std::strcmp(e, "1") || !std::strcmp(e, "col");
        // ... except on an axis where the 1024-sample blocks fit the line badly: the last of several two-factor
        // blocks nearly empty while fewer, longer three-factor blocks cover the line with at least 20 % fewer samples
        // (1080 rows at sigma = 5: one block of 1536 instead of two of 1024 -> column pass 0.043 -> 0.037 ms, measured;
        // every other measured shape — 4K sigma 25, 4096^2 sigma 10, 8192^2 and 16384^2 sigma 50 — keeps 1024).
        // Only the line length and the pad enter, never the rank count: sharded and single-GPU plans agree bit for bit.
        auto prefer3 = [](const OlsAxis& a3, const OlsAxis& a2, int len) {
            if (a2.nblocks < 2) return false;
            const int last_valid = len - (a2.nblocks - 1) * a2.V;
            return last_valid * 4 < a2.V && static_cast<long long>(a3.nblocks) * a3.L * 5 < static_cast<long long>(a2.nblocks) * a2.L * 4;
        };
        const bool forced = e != nullptr;
        OlsAxis a2;
        if (want_row && choose_ols2_axis(cols, gg.pad, a2, true) && (forced || !prefer3(g.ols_row, a2, cols))) g.ols_row = a2;
        if (want_col && choose_ols2_axis(rows, gg.pad, a2) && (forced || !prefer3(g.ols_col, a2, rows))) g.ols_col = a2;
    }
    const int cp_total = (cols + 1) / 2;
    int sr = (rows + nranks - 1) / nranks;
    if (g.algo == 1) {
        g.cpp = 4 * ((((cols + 7) / 8) + nranks - 1) / nranks);     // whole column octets per rank
        g.qpitch = g.cpart = 2 * g.cpp;
    } else {
        g.cpp = (cp_total + nranks - 1) / nranks;
    }
    sr += sr & 1;        // stripes start on even rows: the same row pairs share an FFT for every nranks
    g.stripe_rows = sr;
    g.row0 = rank * sr;
    g.valid_rows = std::max(0, std::min(sr, rows - g.row0));
    g.valid_ncp = std::max(0, std::min(g.cpp, cp_total - rank * g.cpp));
    g.block_floats = static_cast<size_t>(3) * g.cpp * sr * 2;
    return g;
}

// twA[(q-1)*256 + t] = exp(-2 pi i q t / L), q = 1..R0-1, t < 256
inline void make_ols_twiddles_a(int R0, std::vector<float2>& tw)
{
    const int L = 256 * R0;
    tw.resize(static_cast<size_t>(std::max(1, R0 - 1)) * 256);
    for (int q = 1; q < R0; ++q)
        for (int t = 0; t < 256; ++t) {
            const double a = -2.0 * 3.14159265358979323846 * static_cast<double>(q) * t / L;
            tw[static_cast<size_t>(q - 1) * 256 + t] = make_float2(static_cast<float>(std::cos(a)), static_cast<float>(std::sin(a)));
        }
}

// twB[(q-1)*16 + t] = exp(-2 pi i q t / 256), q = 1..15, t < 16
inline void make_ols_twiddles_b(std::vector<float2>& tw)
{
    tw.resize(15 * 16);
    for (int q = 1; q < 16; ++q)
        for (int t = 0; t < 16; ++t) {
            const double a = -2.0 * 3.14159265358979323846 * static_cast<double>(q) * t / 256.0;
            tw[static_cast<size_t>(q - 1) * 16 + t] = make_float2(static_cast<float>(std::cos(a)), static_cast<float>(std::sin(a)));
        }
}

// kt[q * (16 R0) + u] = Re K_L[k0 + R0*(k1 + 16 q)] * (1/L),  u = 16*k0 + k1.
// K_L = DFT of the reference's rotated float taps (Source.cpp:75-102) zero-extended to L; only
// its real part is used, like Source.cpp:362,423.  1/L folded in as a float product (:506-507).
inline void make_ols_kt(int R0, double sigma, int ksize, std::vector<float>& kt)
{
    const int L = 256 * R0, W = 16 * R0;
    std::vector<double> K;
    kernel_spectrum(sigma, ksize, L, K);
    const float scaler = 1.f / static_cast<float>(L);
    kt.resize(static_cast<size_t>(L));
    for (int u = 0; u < W; ++u)
        for (int q = 0; q < 16; ++q) {
            const int k0 = u >> 4, k1 = u & 15;
            const int k = k0 + R0 * (k1 + 16 * q);
            kt[static_cast<size_t>(q) * W + u] = static_cast<float>(K[k]) * scaler;
        }
}

// Tables of the two-factor kernels (ols2_core.cuh), L = 32 * R2:
//   tw2[j * 32 + t] = exp(-2 pi i j t / L),  j < R2, t < 32     (thread constants W_L^{t j})
//   kt[k2 * 32 + k1] = Re K_L[k1 + 32 k2] * (1/L)               (same provenance as make_ols_kt)
inline void make_ols2_tables(int R2, double sigma, int ksize, std::vector<float2>& tw2, std::vector<float>& kt)
{
    const int L = 32 * R2;
    tw2.resize(static_cast<size_t>(R2) * 32);
    for (int j = 0; j < R2; ++j)
        for (int t = 0; t < 32; ++t) {
            const double a = -2.0 * 3.14159265358979323846 * static_cast<double>(j) * t / L;
            tw2[static_cast<size_t>(j) * 32 + t] = make_float2(static_cast<float>(std::cos(a)), static_cast<float>(std::sin(a)));
        }
    std::vector<double> K;
    kernel_spectrum(sigma, ksize, L, K);
    const float scaler = 1.f / static_cast<float>(L);
    kt.resize(static_cast<size_t>(L));
    for (int k2 = 0; k2 < R2; ++k2)
        for (int k1 = 0; k1 < 32; ++k1) kt[static_cast<size_t>(k2) * 32 + k1] = static_cast<float>(K[k1 + 32 * k2]) * scaler;
}

}  // namespace blurb200
