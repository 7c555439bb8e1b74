// tests/cpp/dropin_test.cpp — C++ parity test of the drop-in host layer (include/blur_algorithms.hpp).
//
// Reads like the reference's Test(image, flag, nsmooth) dispatcher (Source.cpp:574-608): every
// flag calls the function of the same name on image.data, and the result is compared with the CPU
// oracle (oracle/blur_oracle.h — test infrastructure) on the same seeded input.
//   g++ -std=c++17 -I include -I oracle tests/cpp/dropin_test.cpp -L blur_algorithms_b200 -lblur_b200 \
//       -L oracle -loracle -o tests/cpp/dropin_test
// `dropin_test link` only checks that every entry point links and the GPU-free helpers agree with
// the oracle (runs in the CPU container); without arguments it needs a CUDA device.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>

#include "blur_algorithms.hpp"
#include "blur_oracle.h"

namespace {
int g_fail = 0;
void expect(bool ok, const char* what)
{
    std::printf("%-58s %s\n", what, ok ? "ok" : "FAIL");
    if (!ok) ++g_fail;
}

std::vector<uint8_t> synth(int rows, int cols, uint32_t seed)
{
    std::vector<uint8_t> v(static_cast<size_t>(rows) * cols * 3);
    uint32_t s = seed * 2654435761u + 12345u;
    for (auto& b : v) { s = s * 1664525u + 1013904223u; b = static_cast<uint8_t>(s >> 24); }
    return v;
}

int max_abs_diff(const std::vector<uint8_t>& a, const std::vector<uint8_t>& b)
{
    int m = 0;
    for (size_t i = 0; i < a.size(); ++i) { const int d = std::abs(int(a[i]) - int(b[i])); if (d > m) m = d; }
    return m;
}

void host_helpers()
{
    bool ok = true;
    for (int n = 0; n < 40000; n += 7)
        ok = ok && isValidSize(n) == orc_is_valid_size(n) && nearestTransformSize(n) == orc_nearest_transform_size(n);
    expect(ok, "isValidSize / nearestTransformSize == oracle");
    ok = true;
    for (double s : {0.3, 1.0, 2.0, 5.0, 25.0, 50.0, 106.770783})
        for (int mw : {0, 100, 331, 332})
            ok = ok && gaussian_window(s, mw) == orc_gaussian_window(s, mw);
    expect(ok, "gaussian_window == oracle");
    std::vector<float> k, ko(64);
    getGaussian(k, 5.0, 31, 64);
    orc_get_gaussian(ko.data(), 5.0, 31, 64);
    expect(k.size() == 64 && !std::memcmp(k.data(), ko.data(), 64 * sizeof(float)), "getGaussian(5, 31, 64) bit-identical to oracle");
}
}  // namespace

int main(int argc, char** argv)
{
    host_helpers();
    if (argc > 1 && !std::strcmp(argv[1], "link")) {
        // take the address of every device-backed entry so a missing symbol fails at link time
        void (*fns[])(uint8_t*, int, int, double) = {pffft_, pocketfft_1D, pocketfft_2D, pocketfft_2D_DFT_image};
        expect(fns[0] && fns[3] && &fastboxblur && &Reflect_101<uint8_t, 3> && &flip_block<float, 1>, "device-backed entry points link");
        return g_fail ? 1 : 0;
    }
    const int rows = 270, cols = 480;
    const double nsmooth = 5.0;
    const std::vector<uint8_t> img = synth(rows, cols, 1);

    {   // flag 5: pocketfft_1D (Source.cpp:598-ish dispatcher -> :280)
        std::vector<uint8_t> a = img, b = img;
        pocketfft_1D(a.data(), rows, cols, nsmooth);
        orc_gauss_blur_1d(b.data(), rows, cols, nsmooth, 0, nullptr);
        expect(max_abs_diff(a, b) <= 1, "pocketfft_1D within 1 LSB of oracle");
        std::vector<uint8_t> c = img;
        pocketfft_2D(c.data(), rows, cols, nsmooth);
        std::vector<uint8_t> d = img;
        orc_gauss_blur_2d(d.data(), rows, cols, nsmooth);
        expect(max_abs_diff(c, d) <= 1, "pocketfft_2D within 1 LSB of oracle");
    }
    {   // flag 3: pffft_ (Nyquist quirk, Source.cpp:414-427)
        std::vector<uint8_t> a = img, b = img;
        pffft_(a.data(), rows, cols, nsmooth);
        orc_gauss_blur_1d(b.data(), rows, cols, nsmooth, 1, nullptr);
        expect(max_abs_diff(a, b) <= 1, "pffft_ within 1 LSB of oracle");
    }
    {   // plan reuse: same result as the one-shot call, twice
        blur_b200::GaussPlan plan(rows, cols, nsmooth);
        std::vector<uint8_t> a = img, b(img.size()), c = img;
        plan.run(a.data(), a.data());
        plan.run(img.data(), b.data());
        pocketfft_1D(c.data(), rows, cols, nsmooth);
        expect(a == b && a == c, "GaussPlan reuse == one-shot call, in place == out of place");
    }
    {   // streamed batches: GaussPlan::run_batch and fastboxblur_batch == the per-image calls
        blur_b200::GaussPlan plan(rows, cols, nsmooth);
        const std::vector<uint8_t> img2 = synth(rows, cols, 2);
        std::vector<uint8_t> a = img, b = img2, c = img, wa(img.size()), wb(img.size()), wc(img.size());
        plan.run(img.data(), wa.data());
        plan.run(img2.data(), wb.data());
        const uint8_t* ins[3] = {a.data(), b.data(), c.data()};
        uint8_t* outs[3] = {a.data(), b.data(), c.data()};
        plan.run_batch(ins, outs, 3);
        expect(a == wa && b == wb && c == wa, "GaussPlan::run_batch == run per image");
        std::vector<uint8_t> two(2 * img.size()), x = img, y = img2;
        std::copy(img.begin(), img.end(), two.begin());
        std::copy(img2.begin(), img2.end(), two.begin() + img.size());
        fastboxblur_batch(two.data(), cols, rows, 3, 25, 2, 2);
        fastboxblur(x.data(), cols, rows, 3, 25, 2);
        fastboxblur(y.data(), cols, rows, 3, 25, 2);
        expect(std::equal(x.begin(), x.end(), two.begin()) && std::equal(y.begin(), y.end(), two.begin() + img.size()),
               "fastboxblur_batch == fastboxblur per image");
    }
    {   // flag 4: fastboxblur(image.data, w, h, channels, nsmooth*nsmooth, 2) — bit exact
        std::vector<uint8_t> a = img, b = img;
        fastboxblur(a.data(), cols, rows, 3, 25, 2);
        orc_fastboxblur(b.data(), cols, rows, 3, 25, 2);
        expect(a == b, "fastboxblur bit-exact vs oracle");
    }
    {   // DFT_image
        std::vector<uint8_t> a = img, b(img.size());
        int padded[2];
        pocketfft_2D_DFT_image(a.data(), rows, cols, 0.1);
        orc_dft_image(img.data(), b.data(), rows, cols, 0.1, nullptr, padded);
        expect(max_abs_diff(a, b) <= 1, "DFT_image within 1 LSB of oracle");
    }
    {   // Reflect_101<uint8_t,3> and <float,1>
        const int sz[2] = {rows, cols};
        const int t = 7, bt = 9, l = 5, r = 11;
        std::vector<uint8_t> a(static_cast<size_t>(rows + t + bt) * (cols + l + r) * 3), b(a.size());
        Reflect_101<uint8_t, 3>(img.data(), a.data(), t, bt, l, r, sz);
        orc_reflect_101_u8(img.data(), b.data(), t, bt, l, r, rows, cols, 3);
        expect(a == b, "Reflect_101<uint8_t,3> bit-exact vs oracle");
        std::vector<float> f(static_cast<size_t>(rows) * cols), fo(static_cast<size_t>(rows + t + bt) * (cols + l + r)), fr(fo.size());
        for (size_t i = 0; i < f.size(); ++i) f[i] = img[i] * 0.25f;
        Reflect_101<float, 1>(f.data(), fo.data(), t, bt, l, r, sz);
        orc_reflect_101_u8(reinterpret_cast<const uint8_t*>(f.data()), reinterpret_cast<uint8_t*>(fr.data()), t, bt, l, r, rows, cols, 4);
        expect(!std::memcmp(fo.data(), fr.data(), fo.size() * sizeof(float)), "Reflect_101<float,1> bit-exact vs oracle");
    }
    {   // deinterleave_BGR / interleave_BGR / flip_block
        const uint32_t total = static_cast<uint32_t>(rows) * cols;
        std::vector<float> p(3 * size_t(total)), q(3 * size_t(total));
        float* pp[3] = {p.data(), p.data() + total, p.data() + 2 * size_t(total)};
        deinterleave_BGR(img.data(), pp, total);
        orc_deinterleave_u8_f32(img.data(), q.data(), q.data() + total, q.data() + 2 * size_t(total), total);
        expect(p == q, "deinterleave_BGR bit-exact vs oracle");
        for (auto& v : p) v += 0.4f;
        std::vector<uint8_t> a(img.size()), b(img.size());
        const float* cp[3] = {pp[0], pp[1], pp[2]};
        interleave_BGR(cp, a.data(), total);
        orc_interleave_f32_u8(pp[0], pp[1], pp[2], b.data(), total);
        expect(a == b, "interleave_BGR bit-exact vs oracle");
        std::vector<float> fa(total), fb(total);
        flip_block<float, 1>(pp[0], fa.data(), cols, rows);
        orc_flip_block_f32(pp[0], fb.data(), cols, rows);
        expect(fa == fb, "flip_block<float,1> bit-exact vs oracle");
        // other instantiations of the templates (Utils.hpp:159-210): <uint16_t,float>, <float,float>, <uint8_t,uint16_t>
        {
            std::vector<uint16_t> w16(img.size()), back16(img.size());
            for (size_t i = 0; i < img.size(); ++i) w16[i] = static_cast<uint16_t>(img[i] * 257u);
            deinterleave_BGR<uint16_t, float>(w16.data(), pp, total);
            bool ok = true;
            for (uint32_t x = 0; x < total && ok; ++x)
                for (int c = 0; c < 3; ++c) ok = ok && pp[c][x] == static_cast<float>(w16[3 * size_t(x) + c]);       // round = 0
            expect(ok, "deinterleave_BGR<uint16_t,float> exact");
            for (auto& v : p) v += 0.4f;
            interleave_BGR<uint16_t, float>(cp, back16.data(), total);
            ok = true;
            for (uint32_t x = 0; x < total && ok; ++x)
                for (int c = 0; c < 3; ++c) ok = ok && back16[3 * size_t(x) + c] == static_cast<uint16_t>(pp[c][x] + 0.5f);   // Utils.hpp:189
            expect(ok, "interleave_BGR<uint16_t,float> = (uint16_t)(v + 0.5f)");
            std::vector<float> fi(img.size()), fo2(img.size());
            for (size_t i = 0; i < img.size(); ++i) fi[i] = img[i] * 0.37f - 3.f;
            std::vector<float> pl(3 * size_t(total));
            float* plp[3] = {pl.data(), pl.data() + total, pl.data() + 2 * size_t(total)};
            deinterleave_BGR<float, float>(fi.data(), plp, total);
            const float* cpl[3] = {plp[0], plp[1], plp[2]};
            interleave_BGR<float, float>(cpl, fo2.data(), total);
            expect(fi == fo2, "deinterleave/interleave_BGR<float,float> round trip is the identity (round = 0)");
            std::vector<uint16_t> p16(3 * size_t(total));
            uint16_t* p16p[3] = {p16.data(), p16.data() + total, p16.data() + 2 * size_t(total)};
            deinterleave_BGR<uint8_t, uint16_t>(img.data(), p16p, total);
            ok = true;
            for (uint32_t x = 0; x < total && ok; ++x)
                for (int c = 0; c < 3; ++c) ok = ok && p16p[c][x] == img[3 * size_t(x) + c];
            expect(ok, "deinterleave_BGR<uint8_t,uint16_t> exact (integral -> integral adds nothing)");
            std::vector<uint8_t> t3(img.size()), t3b(img.size());
            flip_block<uint8_t, 3>(img.data(), t3.data(), cols, rows);
            ok = true;
            for (int y = 0; y < rows && ok; ++y)
                for (int x = 0; x < cols && ok; ++x) ok = !std::memcmp(&t3[(size_t(x) * rows + y) * 3], &img[(size_t(y) * cols + x) * 3], 3);
            expect(ok, "flip_block<uint8_t,3> transposes whole pixels");
        }
    }
    {   // error behaviour: pad > min(rows, cols) - 1 is refused, not read out of bounds
        std::vector<uint8_t> tiny = synth(8, 400, 3);
        bool threw = false;
        try { pocketfft_1D(tiny.data(), 8, 400, 50.0); } catch (const blur_b200::Error& e) { threw = e.code == BLUR_B200_EPAD; }
        expect(threw, "pad > min(rows,cols)-1 -> BLUR_B200_EPAD");
    }
    std::printf("%s (%llu kernel launches)\n", g_fail ? "FAILED" : "all ok", (unsigned long long)blur_b200_kernel_launches());
    return g_fail ? 1 : 0;
}
