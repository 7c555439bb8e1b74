// tests/emul/emulate.cpp — TEST INFRASTRUCTURE (never linked into the product library).
//
// Executes the device code of blur_algorithms_b200/csrc (fft_core.cuh, blur_io.cuh,
// fft_plan.hpp — the very same host/device functions the CUDA kernels call) on the CPU, one
// "thread" after another, with the kernel's barrier structure turned into sequential phases.
// Lets the `-m "not gpu"` suite check the FFT decomposition, scrambled multiply, tile placement
// and HBM layouts against the oracle without a GPU.
#include <algorithm>
#include <cstdio>
#include <vector>
#include "../../blur_algorithms_b200/csrc/fft_plan.hpp"
#include "../../blur_algorithms_b200/csrc/blur_io.cuh"
#include "../../blur_algorithms_b200/csrc/fft_dispatch.cuh"
#include "../../blur_algorithms_b200/csrc/ols_plan.hpp"
#include "../../blur_algorithms_b200/csrc/ols2_core.cuh"

using namespace blurb200;

template <class IO>
static void run_units(const FftGeom& g, const IO& io, const std::vector<float2>& tw, const std::vector<float>& kt)
{
#pragma omp parallel
    {
        std::vector<float2> smem(static_cast<size_t>(g.slots));
        float2* s = smem.data();
#pragma omp for schedule(dynamic, 4)
        for (int unit = 0; unit < io.nunits; ++unit) {
#define F0(R) for (int u = 0; u < g.W[0]; ++u) fwd_pass<R, true>(s, g, 0, u, tw.data(), io, unit)
            BLUR_RADIX_SWITCH(g.radix[0], F0)
#undef F0
            for (int p = 1; p < g.npass - 1; ++p) {
#define F1(R) for (int u = 0; u < g.W[p]; ++u) fwd_pass<R, false>(s, g, p, u, tw.data(), io, unit)
                BLUR_RADIX_SWITCH(g.radix[p], F1)
#undef F1
            }
            for (int u = 0; u < g.N / kLastRadix; ++u) mid_pass(s, g, u, kt.data());
            for (int p = g.npass - 2; p >= 1; --p) {
#define F2(R) for (int u = 0; u < g.W[p]; ++u) inv_pass<R, false>(s, g, p, u, tw.data(), io, unit)
                BLUR_RADIX_SWITCH(g.radix[p], F2)
#undef F2
            }
#define F3(R) for (int u = 0; u < g.W[0]; ++u) inv_pass<R, true>(s, g, 0, u, tw.data(), io, unit)
            BLUR_RADIX_SWITCH(g.radix[0], F3)
#undef F3
        }
    }
}

// ---- overlap-save path (ols_core.cuh) -----------------------------------------------------------
#define OLS_R0_SWITCH(r, F)                                                                         \
    switch (r) {                                                                                    \
        case 4: F(4); break;   case 6: F(6); break;   case 8: F(8); break;   case 12: F(12); break; \
        case 16: F(16); break; case 24: F(24); break; case 32: F(32); break;                        \
        default: break;                                                                             \
    }

struct OlsTables {
    std::vector<float2> twA, twB;
    std::vector<float> kt;
    OlsTables(int R0, double sigma, int ksize)
    {
        make_ols_twiddles_a(R0, twA);
        make_ols_twiddles_b(twB);
        make_ols_kt(R0, sigma, ksize, kt);
    }
};

template <int R0>
static void ols_row_units(const OlsRowIO& io, const OlsTables& T)
{
#pragma omp parallel
    {
        std::vector<float2> smem(static_cast<size_t>(ols_tile_slots(R0)));
        std::vector<uint8_t> raw(2 * ols_raw_pitch<R0>());
        float2* s = smem.data();
        const uint8_t* lo = io.in;
        const uint8_t* hi = io.in + static_cast<size_t>(io.rows) * io.cols * 3;
#pragma omp for schedule(dynamic, 4)
        for (int unit = 0; unit < io.nunits; ++unit) {
            const OlsRowIO::Ctx ctx = io.begin(unit, raw.data(), ols_raw_pitch<R0>());
            // what ols_row_stage's 16-byte copies deliver
            std::fill(raw.begin(), raw.end(), 0xCD);
            for (int k = 0; k < 2 && ctx.ra + k < io.rows; ++k) {
                const OlsRowIO::Stage st = io.stage_of(ctx.ra + k, ctx.s, ctx.w0, ctx.w1);
                for (int e = 0; e < 16 * st.nvec; ++e)
                    if (st.src + e >= lo && st.src + e < hi) raw[k * ols_raw_pitch<R0>() + st.dst + e] = st.src[e];
            }
            if (!ctx.interior)
                for (int e = 0; e < 2 * 256 * R0; ++e) io.fixup(ctx, e);
            for (int ch = 0; ch < 3; ++ch) {
                for (int t = 0; t < 256; ++t) ols_pass_a_fwd<R0, 1>(s, ch, t, T.twA.data(), io, ctx);
                for (int it = 0; it < 16 * R0; ++it) ols_pass_b<false, 1>(s, it, T.twB.data());
                for (int it = 0; it < 16 * R0; ++it) ols_mid<R0, 1>(s, it, T.kt.data());
                for (int it = 0; it < 16 * R0; ++it) ols_pass_b<true, 1>(s, it, T.twB.data());
                { cplx w[R0]; ols_tw_load<R0>(w, T.twA.data(), 0);
                  for (int t = 0; t < 256; ++t) ols_pass_a_inv<R0, 1>(s, ch, t, t + 1, T.twA.data(), w, io, ctx); }
            }
        }
    }
}

// q = Q as received: [3][rows_total][qpitch]
template <int R0>
static void ols_col_units(const OlsColIO& io, const OlsTables& T)
{
    constexpr int F = 4, L = 256 * R0;
#pragma omp parallel
    {
        std::vector<float2> smem(static_cast<size_t>(F) * ols_tile_slots(R0));
#pragma omp for schedule(dynamic, 4)
        for (int unit = 0; unit < io.nunits; ++unit) {
            const OlsColIO::Ctx ctx = io.begin(unit);
            // what ols_col_fetch's copies deliver: rows outside Q are zero-filled
            std::fill(smem.begin(), smem.end(), make_float2(-777.f, 777.f));
            for (int i = 0; i < L; ++i)
                for (int X = 0; X < F; ++X) {
                    const float* p = io.src(ctx, X, i);
                    smem[static_cast<size_t>(F) * ols_pad(i) + X] = p ? cpack(p[0], p[1]) : czero();
                }
            if (!ctx.interior)
                for (int e = 0; e < F * L; ++e) ols_col_patch(smem.data(), e, ctx.s, io.ax);
            for (int X = 0; X < F; ++X) {
                float2* sx = smem.data() + X;
                for (int t = 0; t < 256; ++t) ols_pass_a_fwd<R0, F>(sx, X, t, T.twA.data(), io, ctx);
                for (int it = 0; it < 16 * R0; ++it) ols_pass_b<false, F>(sx, it, T.twB.data());
                for (int it = 0; it < 16 * R0; ++it) ols_mid<R0, F>(sx, it, T.kt.data());
                for (int it = 0; it < 16 * R0; ++it) ols_pass_b<true, F>(sx, it, T.twB.data());
                { cplx w[R0]; ols_tw_load<R0>(w, T.twA.data(), 0);
                  for (int t = 0; t < 256; ++t) ols_pass_a_inv<R0, F>(sx, X, t, t + 1, T.twA.data(), w, io, ctx); }
            }
        }
    }
}

// ---- two-factor overlap-save path (ols2_core.cuh): the kernels' phases, one item after another ----
struct Ols2Tables {
    std::vector<float2> tw2;
    std::vector<float> kt;
    Ols2Tables(int R2, double sigma, int ksize) { make_ols2_tables(R2, sigma, ksize, tw2, kt); }
    // thread constants of item t: w[j] = W_L^{t j}
    template <int R2> void item_twiddles(int t, cplx (&w)[R2]) const
    {
        w[0] = cpack(1.f, 0.f);
        for (int j = 1; j < R2; ++j) w[j] = tw2[static_cast<size_t>(j) * 32 + t];
    }
};

template <int R2>
static void ols2_row_units(const OlsRowIO& io, const Ols2Tables& T)
{
#pragma omp parallel
    {
        std::vector<float2> tile(static_cast<size_t>(ols2_tile_slots<R2>()));
        std::vector<uint8_t> raw(2 * ols2_raw_pitch<R2>());
        const uint8_t* lo = io.in;
        const uint8_t* hi = io.in + static_cast<size_t>(io.rows) * io.cols * 3;
#pragma omp for schedule(dynamic, 4)
        for (int unit = 0; unit < io.nunits; ++unit) {
            const OlsRowIO::Ctx ctx = io.begin(unit, raw.data(), ols2_raw_pitch<R2>());
            std::fill(raw.begin(), raw.end(), 0xCD);
            for (int k = 0; k < 2 && ctx.ra + k < io.rows; ++k) {
                const OlsRowIO::Stage st = io.stage_of(ctx.ra + k, ctx.s, ctx.w0, ctx.w1);
                for (int e = 0; e < 16 * st.nvec; ++e)
                    if (st.src + e >= lo && st.src + e < hi) raw[k * ols2_raw_pitch<R2>() + st.dst + e] = st.src[e];
            }
            if (!ctx.interior)
                for (int e = 0; e < 2 * ols2_len<R2>(); ++e) io.fixup(ctx, e);
            for (int ch = 0; ch < 3; ++ch) {
                std::fill(tile.begin(), tile.end(), make_float2(-777.f, 777.f));
                for (int n2 = 0; n2 < R2; ++n2) ols2_row_pass1<R2>(tile.data(), ctx, ch, n2);
                for (int k1 = 0; k1 < 32; ++k1) { cplx w[R2]; T.item_twiddles<R2>(k1, w); ols2_pass2<R2, 1>(tile.data(), k1, w, T.kt.data()); }
                for (int n2 = 0; n2 < R2; ++n2) { cplx w[32]; T.item_twiddles<32>(n2, w); if (io.nparts > 1) ols2_row_pass3<R2, true>(tile.data(), io, ctx, ch, n2, w); else ols2_row_pass3<R2, false>(tile.data(), io, ctx, ch, n2, w); }
            }
        }
    }
}

template <int R2>
static void ols2_col_units(const OlsColIO& io, const Ols2Tables& T)
{
    constexpr int F = 4, L = ols2_len<R2>();
#pragma omp parallel
    {
        std::vector<float2> smem(static_cast<size_t>(F) * ols2_tile_slots<R2>());
#pragma omp for schedule(dynamic, 4)
        for (int unit = 0; unit < 2 * io.nunits; ++unit) {
            // sub-unit = member m of the (strip, block) pair q, as ols2_col_kernel walks them
            const int q = unit >> 1, m = unit & 1;
            const Ols2ColWalk wk = Ols2ColWalk::at(q, io.nbl);
            const OlsColIO::Ctx ctx = ols2_col_ctx(io, wk.strip, wk.b, m);
            if (!ctx.valid) continue;
            std::fill(smem.begin(), smem.end(), make_float2(-777.f, 777.f));
            for (int i = 0; i < L; ++i)
                for (int X = 0; X < F; ++X) {
                    const float* p = io.src(ctx, X, i);
                    smem[static_cast<size_t>(F) * ols2_slot<R2>(i) + X] = p ? cpack(p[0], p[1]) : czero();
                }
            if (!ctx.interior)
                for (int e = 0; e < F * L; ++e) ols2_col_patch<R2>(smem.data(), e, ctx.s, io.ax);
            for (int X = 0; X < F; ++X) {
                float2* sx = smem.data() + X;
                for (int n2 = 0; n2 < R2; ++n2) ols2_col_pass1<R2>(sx, n2);
                for (int k1 = 0; k1 < 32; ++k1) { cplx w[R2]; T.item_twiddles<R2>(k1, w); ols2_pass2<R2, F>(sx, k1, w, T.kt.data()); }
                for (int n2 = 0; n2 < R2; ++n2) {
                    cplx w[32], x[32];
                    T.item_twiddles<32>(n2, w);
                    ols2_col_pass3_load<R2>(sx, n2, x);
                    ols2_col_pass3_store<R2>(io, ctx, X, n2, x, w);
                }
            }
        }
    }
}

static int ols_row_pass(const uint8_t* in_stripe, float* q, const GaussGeometry& gg, const ShardGeom& sg, double sigma,
                        int nranks, float* const* qpeer = nullptr, int rank = 0)
{
    if (sg.valid_rows <= 0) return 0;
    OlsRowIO io{};
    io.in = in_stripe; io.q = q; io.rows = sg.valid_rows; io.cols = gg.cols; io.npairs = (sg.valid_rows + 1) / 2;
    io.srows = sg.stripe_rows; io.qpitch = sg.qpitch; io.cpart = sg.cpart; io.nparts = nranks;
    io.ax = sg.ols_row; io.nunits = io.ax.nblocks * io.npairs;
    if (qpeer) { io.peer = 1; io.rank = rank; for (int p = 0; p < nranks; ++p) io.qpeer[p] = qpeer[p]; }
    if (io.ax.R2 == 32) { ols2_row_units<32>(io, Ols2Tables(32, sigma, gg.ksize)); return 0; }
    const OlsTables T(io.ax.R0, sigma, gg.ksize);
#define F_(R) ols_row_units<R>(io, T)
    OLS_R0_SWITCH(io.ax.R0, F_)
#undef F_
    return 0;
}

// out_cols: uint8 [rows][2*cpp][3] (this rank's column stripe)
static int ols_col_pass(const float* q_recv, uint8_t* out_cols, const GaussGeometry& gg, const ShardGeom& sg, double sigma,
                        int rank, int nranks)
{
    const int rows = gg.rows;
    OlsColIO io{};
    io.noct = std::min(sg.qpitch / 8, (std::max(0, gg.cols - rank * sg.cpart) + 7) / 8);
    if (io.noct <= 0) return 0;
    std::vector<uint8_t> U(static_cast<size_t>(3) * io.noct * rows * 8);
    io.u = U.data(); io.rows = rows; io.ax = sg.ols_col; io.nunits = io.ax.nblocks * 3 * io.noct;
    io.q = q_recv; io.qpitch = sg.qpitch; io.rows_total = nranks * sg.stripe_rows;
    if (io.ax.R2 == 32) {
        io.nbl = io.ax.nblocks; io.nunits = ols2_col_npairs(io.noct, io.nbl);
        ols2_col_units<32>(io, Ols2Tables(32, sigma, gg.ksize));
    } else {
        const OlsTables T(io.ax.R0, sigma, gg.ksize);
#define F_(R) ols_col_units<R>(io, T)
        OLS_R0_SWITCH(io.ax.R0, F_)
#undef F_
    }
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < io.noct * 8 && c < 2 * sg.cpp; ++c)
            for (int ch = 0; ch < 3; ++ch)
                out_cols[(static_cast<size_t>(r) * 2 * sg.cpp + c) * 3 + ch] =
                    U[((static_cast<size_t>(ch) * io.noct + (c >> 3)) * rows + r) * 8 + (c & 7)];
    return 0;
}

extern "C" {

// Overlap-save Gaussian path on the CPU through the device code, cut over `nparts` emulated
// ranks (row stripes -> exchange by copies -> column stripes).  Returns 0, or 4 if the plan would
// not pick the overlap-save algorithm.  info: [0]=L_row [1]=L_col [2]=nblocks_row [3]=nblocks_col
int emul_ols_gauss(const uint8_t* in, uint8_t* out, int rows, int cols, double sigma, int nparts, float* rowpass_planar,
                   int* info)
{
    const GaussGeometry gg = gauss_geometry(rows, cols, sigma);
    if (gg.pad < 0 || gg.pad > rows - 1 || gg.pad > cols - 1) return 2;
    const bool peer = nparts < 0;          // negative: fused exchange (row pass stores into the owners' buffers)
    const int P = peer ? -nparts : nparts;
    std::vector<ShardGeom> sg;
    for (int r = 0; r < P; ++r) sg.push_back(shard_geometry(gg, 0, r, P));
    if (sg[0].algo != 1) return 4;
    if (info) { info[0] = sg[0].ols_row.L; info[1] = sg[0].ols_col.L; info[2] = sg[0].ols_row.nblocks; info[3] = sg[0].ols_col.nblocks; }
    const size_t bf = sg[0].block_floats, chf = bf / 3;      // floats per (stripe, part) block / per channel of it
    if (peer) {
        std::vector<std::vector<float>> recv(P, std::vector<float>(bf * P, 0.f));
        std::vector<float*> ptrs;
        for (auto& v : recv) ptrs.push_back(v.data());
        for (int r = 0; r < P; ++r)
            ols_row_pass(in + static_cast<size_t>(sg[r].row0) * cols * 3, nullptr, gg, sg[r], sigma, P, ptrs.data(), r);
        for (int p = 0; p < P; ++p) {
            std::vector<uint8_t> oc(static_cast<size_t>(rows) * 2 * sg[p].cpp * 3, 0);
            ols_col_pass(recv[p].data(), oc.data(), gg, sg[p], sigma, p, P);
            for (int r = 0; r < rows; ++r)
                for (int c = 0; c < 2 * sg[p].cpp; ++c) {
                    const int x = p * 2 * sg[p].cpp + c;
                    if (x >= cols) continue;
                    for (int ch = 0; ch < 3; ++ch)
                        out[(static_cast<size_t>(r) * cols + x) * 3 + ch] = oc[(static_cast<size_t>(r) * 2 * sg[p].cpp + c) * 3 + ch];
                }
        }
        return 0;
    }
    std::vector<std::vector<float>> send(P, std::vector<float>(bf * P, 0.f));
    for (int r = 0; r < P; ++r)
        ols_row_pass(in + static_cast<size_t>(sg[r].row0) * cols * 3, send[r].data(), gg, sg[r], sigma, P);
    if (rowpass_planar)
        for (int s = 0; s < P; ++s)
            for (int ch = 0; ch < 3; ++ch)
                for (int r = 0; r < sg[s].valid_rows; ++r)
                    for (int x = 0; x < cols; ++x) {
                        const int part = x / sg[s].cpart, xl = x - part * sg[s].cpart;
                        rowpass_planar[(static_cast<size_t>(ch) * rows + sg[s].row0 + r) * cols + x] =
                            send[s][((static_cast<size_t>(ch) * P + part) * sg[s].stripe_rows + r) * sg[s].qpitch + xl];
                    }
    for (int p = 0; p < P; ++p) {
        // "all-to-all per channel": recv[ch][src] = send[src][ch][p]
        std::vector<float> recv(bf * P);
        for (int ch = 0; ch < 3; ++ch)
            for (int s = 0; s < P; ++s)
                std::copy(send[s].begin() + (static_cast<size_t>(ch) * P + p) * chf,
                          send[s].begin() + (static_cast<size_t>(ch) * P + p + 1) * chf,
                          recv.begin() + (static_cast<size_t>(ch) * P + s) * chf);
        std::vector<uint8_t> oc(static_cast<size_t>(rows) * 2 * sg[p].cpp * 3, 0);
        ols_col_pass(recv.data(), oc.data(), gg, sg[p], sigma, p, P);
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < 2 * sg[p].cpp; ++c) {
                const int x = p * 2 * sg[p].cpp + c;
                if (x >= cols) continue;
                for (int ch = 0; ch < 3; ++ch)
                    out[(static_cast<size_t>(r) * cols + x) * 3 + ch] = oc[(static_cast<size_t>(r) * 2 * sg[p].cpp + c) * 3 + ch];
            }
    }
    return 0;
}

// Gaussian 1-D tile path on the CPU through the device code.  nparts > 1 exercises the
// multi-GPU Q layout (row stripes x column-pair parts) inside one process.
// rowpass_planar (optional): float [3][rows][cols].  Returns 0 on success.
int emul_gauss(const uint8_t* in, uint8_t* out, int rows, int cols, double sigma, int nyquist_mode,
               int nparts, float* rowpass_planar, int* radices_row, int* radices_col)
{
    const GaussGeometry gg = gauss_geometry(rows, cols, sigma);
    if (gg.pad < 0 || gg.pad > rows - 1 || gg.pad > cols - 1) return 2;
    FftGeom grow, gcol;
    if (!make_geom(gg.n_row, grow) || !make_geom(gg.n_col, gcol)) return 3;
    if (radices_row) for (int i = 0; i < kMaxPass; ++i) radices_row[i] = i < grow.npass ? grow.radix[i] : 0;
    if (radices_col) for (int i = 0; i < kMaxPass; ++i) radices_col[i] = i < gcol.npass ? gcol.radix[i] : 0;
    std::vector<float2> tw_row, tw_col;
    make_twiddles(grow.N, tw_row);
    make_twiddles(gcol.N, tw_col);
    std::vector<double> K;
    std::vector<float> kt_row, kt_col;
    kernel_spectrum(sigma, gg.ksize, grow.N, K);
    make_kt(grow, K, nyquist_mode, kt_row);
    kernel_spectrum(sigma, gg.ksize, gcol.N, K);
    make_kt(gcol, K, nyquist_mode, kt_col);

    // "GPUs": stripe s owns rows [s*srows, ...), part p owns column pairs [p*cpp, ...)
    const int P = nparts;
    const int cp_total = (cols + 1) / 2;
    const int cpp = (cp_total + P - 1) / P;
    int srows = (rows + P - 1) / P;
    srows += srows & 1;                                   // stripes start on even rows
    const size_t part_elems = static_cast<size_t>(3) * cpp * srows * 2;   // floats per (stripe, part)
    std::vector<std::vector<float>> Q(P, std::vector<float>(part_elems * P, 0.f));   // Q[stripe][part][...]
    for (int s = 0; s < P; ++s) {
        const int r0 = s * srows, nr = std::max(0, std::min(srows, rows - r0));
        if (nr <= 0) continue;
        RowIO io{};
        io.in = in + static_cast<size_t>(r0) * cols * 3;
        io.q = Q[s].data();
        io.rows = nr; io.cols = cols; io.pad = gg.pad; io.N = grow.N;
        io.qrows = srows; io.cpp = cpp;
        io.nunits = ((nr + 3) / 4) * 6;
        run_units(grow, io, tw_row, kt_row);
    }
    if (rowpass_planar) {
        for (int s = 0; s < P; ++s)
            for (int ch = 0; ch < 3; ++ch)
                for (int r = s * srows; r < std::min(rows, (s + 1) * srows); ++r)
                    for (int x = 0; x < cols; ++x) {
                        const int cp = x >> 1, part = cp / cpp, cpl = cp - part * cpp;
                        rowpass_planar[(static_cast<size_t>(ch) * rows + r) * cols + x] =
                            Q[s][((static_cast<size_t>(part * 3 + ch) * cpp + cpl) * srows + (r - s * srows)) * 2 + (x & 1)];
                    }
    }
    // "all-to-all": part p gathers block p of every stripe -> chunks [stripe][3][cpp][srows][2]
    for (int p = 0; p < P; ++p) {
        std::vector<float> recv(part_elems * P);
        for (int s = 0; s < P; ++s)
            std::copy(Q[s].begin() + part_elems * p, Q[s].begin() + part_elems * (p + 1), recv.begin() + part_elems * s);
        const int cp0 = p * cpp, ncp = std::max(0, std::min(cpp, cp_total - cp0));
        if (ncp <= 0) continue;
        std::vector<uint8_t> U(static_cast<size_t>(3) * 2 * cpp * rows);
        ColIO io{};
        io.q = reinterpret_cast<const float2*>(recv.data());
        io.chunk_stride = part_elems / 2;
        io.chunk_rows = srows;
        io.u = U.data();
        io.rows = rows; io.ncp = cpp; io.valid_ncp = ncp; io.ucols = 2 * cpp; io.pad = gg.pad; io.N = gcol.N;
        io.nunits = 3 * cpp;
        run_units(gcol, io, tw_col, kt_col);
        for (int ch = 0; ch < 3; ++ch)
            for (int c = 0; c < 2 * ncp; ++c) {
                const int x = 2 * cp0 + c;
                if (x >= cols) continue;
                for (int r = 0; r < rows; ++r)
                    out[(static_cast<size_t>(r) * cols + x) * 3 + ch] = U[(static_cast<size_t>(ch) * 2 * cpp + c) * rows + r];
            }
    }
    return 0;
}

// Per-rank pieces with the exact buffer contract of blur_b200_gauss_row_pass / _col_pass, so the
// CPU (gloo) tests can run blur_algorithms_b200/dist.py without a GPU.
int emul_row_pass(const uint8_t* in_stripe, float* q, int rows, int cols, double sigma, int nyquist_mode,
                  int rank, int nranks)
{
    const GaussGeometry gg = gauss_geometry(rows, cols, sigma);
    {
        const ShardGeom sg = shard_geometry(gg, nyquist_mode, rank, nranks);
        if (sg.algo == 1) return ols_row_pass(in_stripe, q, gg, sg, sigma, nranks);
    }
    FftGeom grow;
    if (!make_geom(gg.n_row, grow)) return 3;
    std::vector<float2> tw; make_twiddles(grow.N, tw);
    std::vector<double> K; std::vector<float> kt;
    kernel_spectrum(sigma, gg.ksize, grow.N, K); make_kt(grow, K, nyquist_mode, kt);
    const int cp_total = (cols + 1) / 2, cpp = (cp_total + nranks - 1) / nranks;
    int srows = (rows + nranks - 1) / nranks; srows += srows & 1;
    const int r0 = rank * srows, nr = std::max(0, std::min(srows, rows - r0));
    if (nr <= 0) return 0;
    RowIO io{};
    io.in = in_stripe; io.q = q; io.rows = nr; io.cols = cols; io.pad = gg.pad; io.N = grow.N;
    io.qrows = srows; io.cpp = cpp; io.nunits = ((nr + 3) / 4) * 6;
    run_units(grow, io, tw, kt);
    return 0;
}

int emul_col_pass(const float* q_chunks, uint8_t* out_cols, int rows, int cols, double sigma, int nyquist_mode,
                  int rank, int nranks)
{
    const GaussGeometry gg = gauss_geometry(rows, cols, sigma);
    {
        const ShardGeom sg = shard_geometry(gg, nyquist_mode, rank, nranks);
        if (sg.algo == 1) return ols_col_pass(q_chunks, out_cols, gg, sg, sigma, rank, nranks);
    }
    FftGeom gcol;
    if (!make_geom(gg.n_col, gcol)) return 3;
    std::vector<float2> tw; make_twiddles(gcol.N, tw);
    std::vector<double> K; std::vector<float> kt;
    kernel_spectrum(sigma, gg.ksize, gcol.N, K); make_kt(gcol, K, nyquist_mode, kt);
    const int cp_total = (cols + 1) / 2, cpp = (cp_total + nranks - 1) / nranks;
    int srows = (rows + nranks - 1) / nranks; srows += srows & 1;
    const size_t part_elems = static_cast<size_t>(3) * cpp * srows * 2;
    std::vector<uint8_t> U(static_cast<size_t>(3) * 2 * cpp * rows);
    ColIO io{};
    io.q = reinterpret_cast<const float2*>(q_chunks);
    io.chunk_stride = part_elems / 2; io.chunk_rows = srows; io.u = U.data();
    io.rows = rows; io.ncp = cpp; io.valid_ncp = std::max(0, std::min(cpp, cp_total - rank * cpp));
    io.ucols = 2 * cpp; io.pad = gg.pad; io.N = gcol.N; io.nunits = 3 * cpp;
    run_units(gcol, io, tw, kt);
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < 2 * cpp; ++c)
            for (int ch = 0; ch < 3; ++ch)
                out_cols[(static_cast<size_t>(r) * 2 * cpp + c) * 3 + ch] = U[(static_cast<size_t>(ch) * 2 * cpp + c) * rows + r];
    return 0;
}

// raw complex round trip through the passes for one length (no kernel): out = ifft(fft(in) * 1/N)
int emul_fft_roundtrip(int N, const float* in_re, const float* in_im, float* out_re, float* out_im,
                       float* spec_re, float* spec_im)
{
    FftGeom g;
    if (!make_geom(N, g)) return 3;
    std::vector<float2> tw;
    make_twiddles(N, tw);
    struct IO {
        const float *ir, *ii; float *orr, *oi; int nunits;
        struct Ctx {};
        Ctx begin(int) const { return Ctx(); }
        void load(const Ctx&, int n, float& re, float& im) const { re = ir[n]; im = ii[n]; }
        void store(const Ctx&, int n, float re, float im) const { orr[n] = re; oi[n] = im; }
    } io{in_re, in_im, out_re, out_im, 1};
    std::vector<float> kt(static_cast<size_t>(N), 1.f / N);
    run_units(g, io, tw, kt);
    if (spec_re) {
        // forward only, unscrambled, for checking against numpy.fft
        std::vector<float2> s(static_cast<size_t>(g.slots));
#define F0(R) for (int u = 0; u < g.W[0]; ++u) fwd_pass<R, true>(s.data(), g, 0, u, tw.data(), io, 0)
        BLUR_RADIX_SWITCH(g.radix[0], F0)
#undef F0
        for (int p = 1; p < g.npass - 1; ++p) {
#define F1(R) for (int u = 0; u < g.W[p]; ++u) fwd_pass<R, false>(s.data(), g, p, u, tw.data(), io, 0)
            BLUR_RADIX_SWITCH(g.radix[p], F1)
#undef F1
        }
        for (int u = 0; u < N / 32; ++u) {
            float xr[32], xi[32];
            for (int j = 0; j < 32; ++j) { xr[j] = s[u * 33 + j].x; xi[j] = s[u * 33 + j].y; }
            Dft<32, false>::run(xr, xi);
            for (int q = 0; q < 32; ++q) { int k = scrambled_freq(g, 32 * u + q); spec_re[k] = xr[q]; spec_im[k] = xi[q]; }
        }
    }
    return 0;
}
}
