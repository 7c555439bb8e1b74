"""CPU: pins the oracle (oracle/liboracle.so) against the reference's own golden material.

  * known-answer values printed by the reference's verbatim helpers (tests/golden/known_answers.json)
  * outputs of the reference's pipeline code on seeded inputs (tests/golden/reference_outputs.npz)
  * two of the reference's DFT_image PNG pairs (tests/golden/spectrum_*.png)
  * live comparison with oracle/_ref where it has been built (the build container)
  * an independent float64 direct convolution and a numpy.fft (= pocketfft) restatement
"""
import json
import os

import numpy as np
import pytest

from util import GOLDEN, GOLDEN_BOX_CASES, GOLDEN_CASES, box_numpy, direct_gauss_f64, golden_outputs, synth


@pytest.fixture(scope="module")
def ka():
    with open(os.path.join(GOLDEN, "known_answers.json")) as f:
        return json.load(f)


def test_sizing_known_answers(oracle, ka):
    for s, w in ka["gaussian_window"].items():
        assert oracle.gaussian_window(float(s)) == w
    for m, w in ka["gaussian_window_max50"].items():
        assert oracle.gaussian_window(50.0, int(m)) == w
    for n, v in ka["nearest_transform_size"].items():
        assert oracle.nearest_transform_size(int(n)) == v
    assert [n for n in range(0, 4097) if oracle.is_valid_size(n)] == ka["valid_sizes_upto_4096"]
    # SURVEY.md Appendix B
    assert oracle.gaussian_window(50) == 331 and oracle.nearest_transform_size(16714) == 17280
    assert not oracle.is_valid_size(2160) and oracle.is_valid_size(1920)


def test_taps_known_answers(oracle, ka):
    assert np.array_equal(oracle.get_gaussian(1.0, 5, 0), np.array(ka["gaussian_s1_w5"], np.float32))
    assert np.array_equal(oracle.get_gaussian(2.0, 11, 32), np.array(ka["gaussian_s2_w11_n32"], np.float32))
    assert np.array_equal(oracle.get_gaussian(5.0, 31, 64), np.array(ka["gaussian_s5_w31_n64"], np.float32))
    assert np.array_equal(oracle.get_gaussian(50.0, 331, 17280)[:170], np.array(ka["gaussian_s50_w331_n17280_head"], np.float32))
    k = oracle.get_gaussian(5.0, 31, 64)
    assert k[1] == k[63] and k[16] == 0 and k[48] == 0 and abs(k.astype(np.float64).sum() - 1) < 1e-7


def test_reflect_readme_example(oracle):
    # README.md:47-53: "ABCDEFG" pad 6 -> GFEDCBABCDEFGFEDCBA
    row = np.frombuffer(b"ABCDEFG", np.uint8).reshape(1, 7)
    out = oracle.reflect_101(row, 0, 0, 6, 6)
    assert bytes(out[0]) == b"GFEDCBABCDEFGFEDCBA"
    m = np.arange(1, 13, dtype=np.uint8).reshape(3, 4)
    exp = np.array([[7, 6, 5, 6, 7, 8, 7], [3, 2, 1, 2, 3, 4, 3], [7, 6, 5, 6, 7, 8, 7], [11, 10, 9, 10, 11, 12, 11],
                    [7, 6, 5, 6, 7, 8, 7], [3, 2, 1, 2, 3, 4, 3]], np.uint8)
    assert np.array_equal(oracle.reflect_101(m, 1, 2, 2, 1), exp)
    img = synth("noise", 13, 17)
    for pads in [(1, 2, 2, 1), (5, 7, 3, 9), (12, 12, 16, 16)]:
        t, b, l, r = pads
        assert np.array_equal(oracle.reflect_101(img, *pads), np.pad(img, ((t, b), (l, r), (0, 0)), mode="reflect"))


def test_interleave_rounding(oracle):
    pl = np.array([[0.4, 0.5, 254.6, 255.4]] * 3, np.float32)
    assert list(oracle.interleave(pl)) == [0, 0, 0, 1, 1, 1, 255, 255, 255, 255, 255, 255]
    img = synth("noise", 4, 5)
    assert np.array_equal(oracle.interleave(oracle.deinterleave(img)).reshape(img.shape), img)


def test_gauss_vs_reference_outputs(oracle):
    gold = golden_outputs()
    for kind, rows, cols, sigma in GOLDEN_CASES:
        img = synth(kind, rows, cols)
        key = f"{kind}_{rows}x{cols}_s{sigma}"
        assert np.array_equal(oracle.gauss_blur_1d(img, sigma, 0), gold[key + "_pocketfft1d"]), key
        assert np.array_equal(oracle.gauss_blur_1d(img, sigma, 1), gold[key + "_pffft"]), key
        assert np.array_equal(oracle.gauss_blur_2d(img, sigma), gold[key + "_pocketfft2d"]), key


def test_box_vs_reference_outputs_and_numpy(oracle):
    gold = golden_outputs()
    for kind, rows, cols, ns in GOLDEN_BOX_CASES:
        img = synth(kind, rows, cols)
        out = oracle.fastboxblur(img, int(ns * ns), 2)
        assert np.array_equal(out, gold[f"{kind}_{rows}x{cols}_n{ns}_box"])
        assert np.array_equal(out, box_numpy(img, int(ns * ns), 2))
    # even ksize behaves as the next smaller odd width; 1 channel and 4 channels
    g = synth("noise", 40, 50)[..., 0]
    assert np.array_equal(oracle.fastboxblur(g, 8, 1), box_numpy(g, 7, 1))
    rgba = np.concatenate([synth("noise", 33, 47), synth("noise", 33, 47, 1)[..., :1]], -1)
    assert np.array_equal(oracle.fastboxblur(rgba, 11, 3), box_numpy(rgba, 11, 3))
    with pytest.raises(ValueError):
        oracle.fastboxblur(g, 2 * 40 + 1, 1)


def test_box_float_and_integer_rounding_agree():
    # the upstream lineage rounds acc*(1.f/k)+0.5f in float; the declared integer rule gives the
    # same byte for every reachable accumulator at the widths used
    for k in (3, 9, 31, 225, 961):
        acc = np.arange(0, 255 * k + 1, dtype=np.int64)
        f = (acc.astype(np.float32) * np.float32(1.0 / k) + np.float32(0.5)).astype(np.int64)
        assert np.array_equal(f, (2 * acc + k) // (2 * k)), k


def test_gauss_vs_float64_direct(oracle):
    for kind, rows, cols, sigma in [("noise", 240, 320, 5.0), ("stripes", 96, 128, 2.5), ("noise", 135, 241, 6.0)]:
        img = synth(kind, rows, cols)
        out, rp, fin = oracle.gauss_blur_1d(img, sigma, 0, want_f32=True)
        y, z = direct_gauss_f64(oracle, img, sigma)
        assert np.abs(rp - y.transpose(2, 0, 1)).max() < 255 * 1e-5
        assert np.abs(fin - z.transpose(2, 0, 1)).max() < 255 * 1e-5
        u = (z + 0.5).astype(np.uint8)
        assert np.abs(u.astype(int) - out.astype(int)).max() <= 1


def test_gauss_numpy_pocketfft_twin(oracle):
    # numpy.fft is mreineck's pocketfft: restate pocketfft_1D (Source.cpp:353-384) with it
    img = synth("noise", 100, 75)
    sigma = 3.0
    s = oracle.gauss_sizes(100, 75, sigma)
    pad = s["pad"]

    def pass1d(p, N):
        L = p.shape[1]
        kre = np.fft.rfft(oracle.get_gaussian(sigma, s["kSize"], N).astype(np.float64)).real
        tile = np.zeros((p.shape[0], N))
        tile[:, :pad] = p[:, pad:0:-1]
        tile[:, pad:pad + L] = p
        tile[:, pad + L:pad + L + pad] = p[:, L - 2:L - 2 - pad:-1]
        return np.fft.irfft(np.fft.rfft(tile) * kre, N)[:, pad:pad + L]
    ref = np.stack([pass1d(pass1d(img[..., c].astype(np.float64), s["n_row"]).T, s["n_col"]).T for c in range(3)])
    out, rp, fin = oracle.gauss_blur_1d(img, sigma, 0, want_f32=True)
    assert np.abs(fin - ref).max() < 255 * 1e-5


def test_nyquist_quirk_formula(oracle):
    # pffft mode = exact + (K0 - K_{N/2})/N * X[N/2] * (-1)^(pad+x) on the padded tile
    img = synth("noise", 64, 96, 3)
    sigma = 2.0
    s = oracle.gauss_sizes(64, 96, sigma)
    _, rp0, _ = oracle.gauss_blur_1d(img, sigma, 0, want_f32=True)
    _, rp1, _ = oracle.gauss_blur_1d(img, sigma, 1, want_f32=True)
    N, pad, L = s["n_row"], s["pad"], 96
    taps = oracle.get_gaussian(sigma, s["kSize"], N).astype(np.float64)
    K = np.fft.rfft(taps).real
    row = img[5, :, 1].astype(np.float64)
    tile = np.zeros(N)
    tile[:pad] = row[pad:0:-1]; tile[pad:pad + L] = row; tile[pad + L:pad + L + pad] = row[L - 2:L - 2 - pad:-1]
    xn = (tile * (-1.0) ** np.arange(N)).sum()
    corr = (K[0] - K[N // 2]) / N * xn * (-1.0) ** (pad + np.arange(L))
    assert np.abs((rp1[1, 5] - rp0[1, 5]) - corr).max() < 2e-4


def test_dft_image_golden_pairs(oracle):
    cv2 = pytest.importorskip("cv2")
    im = cv2.imread(os.path.join(GOLDEN, "spectrum_in_blur200.png"), cv2.IMREAD_COLOR)
    gold = cv2.imread(os.path.join(GOLDEN, "spectrum_out_blur200.png"), cv2.IMREAD_COLOR)
    d = np.abs(oracle.dft_image(im, 0.1).astype(int) - gold.astype(int))
    assert d.max() <= 1 and d.mean() < 0.01
    im = cv2.imread(os.path.join(GOLDEN, "spectrum_in_noise200.png"), cv2.IMREAD_COLOR)
    gold = cv2.imread(os.path.join(GOLDEN, "spectrum_out_noise200_rows384_640.png"), cv2.IMREAD_COLOR)
    d = np.abs(oracle.dft_image(im, 0.1)[384:640].astype(int) - gold.astype(int))
    assert d.max() <= 1 and d.mean() < 0.01


def test_dft_image_all_reference_pairs(oracle):
    """All 22 + 1 pairs, only where the reference tree is mounted (build container)."""
    cv2 = pytest.importorskip("cv2")
    base = "/root/reference/test_images/spectrum_analysis"
    if not os.path.isdir(base):
        pytest.skip("reference tree not mounted")
    for r in [10, 20, 30, 40, 50, 60, 70, 80, 90, 100, 200]:
        for a, b in [(f"{base}/blur/spatial domain/blur_radius{r}.png", f"{base}/blur/time domain/radius_{r}.png"),
                     (f"{base}/noisy/spatial domain/noise_gauss{r}.png", f"{base}/noisy/time domain/{r}.png")]:
            d = np.abs(oracle.dft_image(cv2.imread(a, cv2.IMREAD_COLOR), 0.1).astype(int) - cv2.imread(b, cv2.IMREAD_COLOR).astype(int))
            assert d.max() <= 1 and d.mean() < 0.01, a
    c0 = cv2.imread(f"{base}/blur/collage/0.png", cv2.IMREAD_COLOR)[:1024]
    d = np.abs(oracle.dft_image(np.ascontiguousarray(c0), 0.1).astype(int) - cv2.imread(f"{base}/blur/time domain/radius_0.png", cv2.IMREAD_COLOR).astype(int))
    assert d.max() <= 1


def test_oracle_vs_reference_live(oracle, reference):
    """Where oracle/_ref exists: the restatement equals the reference's own code bit for bit."""
    for s in [0.3, 0.5, 1, 2, 3, 5, 10, 25, 38.729833, 50, 100]:
        assert oracle.gaussian_window(s) == reference.gaussian_window(s)
    for n in range(0, 36000, 13):
        assert oracle.is_valid_size(n) == reference.is_valid_size(n)
    for n in [1, 31, 33, 1110, 1950, 2324, 4004, 16714, 33098]:
        assert oracle.nearest_transform_size(n) == reference.nearest_transform_size(n)
    for s, w, N in [(1, 5, 0), (2, 11, 32), (5, 31, 64), (50, 331, 17280), (25, 165, 2400)]:
        assert np.array_equal(oracle.get_gaussian(s, w, N), reference.get_gaussian(s, w, N))
    img = synth("noise", 37, 53, 7)
    for pads in [(1, 2, 2, 1), (5, 7, 3, 9), (36, 36, 52, 52), (50, 50, 80, 80)]:
        assert np.array_equal(oracle.reflect_101(img, *pads), reference.reflect_101(img, *pads))
    assert np.array_equal(oracle.deinterleave(img), reference.deinterleave(img))
    for rows, cols, sigma in [(64, 96, 2.0), (135, 241, 6.0), (300, 200, 9.0)]:
        im = synth("noise", rows, cols, 11)
        assert np.array_equal(oracle.gauss_blur_1d(im, sigma, 0), reference.test(im, 5, sigma))
        assert np.array_equal(oracle.gauss_blur_1d(im, sigma, 1), reference.test(im, 3, sigma))
        assert np.array_equal(oracle.gauss_blur_2d(im, sigma), reference.test(im, 2, sigma))
        assert np.array_equal(oracle.fastboxblur(im, 9, 2), reference.test(im, 4, 3.0))


def test_oracle_equals_reference_at_1080p_and_4k(oracle):
    """BASELINE configs 1 and 2: the C restatement reproduces the reference's own output (oracle/_ref,
    generated by tests/golden/make_golden_large.py) bit for bit — SHA-256 of the whole image."""
    import hashlib
    from util import LARGE_GOLDEN_CASES, large_golden, lattice
    gold = large_golden()
    for rows, cols, sigma in LARGE_GOLDEN_CASES:
        img = synth("noise", rows, cols, 0)
        for mode, name in ((0, "pocketfft1d"), (1, "pffft")):
            out = oracle.gauss_blur_1d(img, sigma, mode)
            key = f"noise_{rows}x{cols}_s{sigma}_{name}"
            assert hashlib.sha256(out.tobytes()).digest() == gold[key + "_sha256"].tobytes(), key
            assert np.array_equal(lattice(out), gold[key + "_lattice"]), key


def test_boxblur_build_tent_kernel_and_reference(oracle):
    """`#define boxblur` (Source.cpp:12): box_kernel's 1-D taps are the tent (kLen - |i|) / kLen^2 (:128-140), the
    pipeline is a reflect-101 tent convolution per axis, and the restatement equals the reference's own boxblur
    build (oracle/_ref/libblurref_box.so) bit for bit."""
    k = 9
    taps = oracle.box_kernel_1d(k, 64)
    want = np.zeros(64)
    for i in range(-k + 1, k):
        want[i % 64] = (k - abs(i)) / k ** 2
    assert np.abs(taps - want).max() < 1e-7 and abs(float(taps.sum()) - 1.0) < 1e-6
    img = synth("noise", 120, 160, 3)
    t = np.array([(k - abs(i)) / k ** 2 for i in range(-k + 1, k)])

    def conv(a, axis):
        ap = np.pad(a.astype(np.float64), [(k - 1, k - 1) if ax == axis else (0, 0) for ax in range(3)], mode="reflect")
        out = np.zeros(a.shape)
        for j, w in enumerate(t):
            sl = [slice(None)] * 3
            sl[axis] = slice(j, j + a.shape[axis])
            out += w * ap[tuple(sl)]
        return out
    exact = np.floor(conv(conv(img, 1), 0) + 0.5).astype(int)
    got = oracle.boxfft_blur(img, 3.0, 0)
    assert np.abs(got.astype(int) - exact).max() <= 1 and (got.astype(int) != exact).mean() < 1e-3
    from oracle_lib import Reference
    if not Reference.available(boxblur=True):
        pytest.skip("oracle/_ref/libblurref_box.so not built (needs /root/reference)")
    ref = Reference(boxblur=True)
    for rows, cols, ns in [(120, 160, 3.0), (300, 200, 5.0), (50, 400, 9.0)]:
        im = synth("noise", rows, cols, 3)
        for flag, mode in ((5, 0), (3, 1)):
            assert np.array_equal(ref.test(im, flag, ns), oracle.boxfft_blur(im, ns, mode)), (rows, cols, ns, flag)
        # pocketfft_2D's boxblur branches (Source.cpp:152-156,207-211,224-228); the restatement refuses windows that
        # nearly fill the shorter side, where the reference's Reflect_101 clamps its pads (Utils.hpp:217-220)
        if 2 * ns * ns > min(rows, cols):
            continue
        oracle.lib.orc_set_boxblur(1)
        try:
            two_d = oracle.gauss_blur_2d(im, ns)
        finally:
            oracle.lib.orc_set_boxblur(0)
        assert np.array_equal(ref.test(im, 2, ns), two_d), (rows, cols, ns)


def test_pffft_quirk_is_rank_one_correction(oracle):
    """The identity DESIGN.md 3.1a builds on: the oracle's pffft-mode row pass (Source.cpp:414-427)
    equals the exact row pass + c (-1)^x A per line, A = weighted alternating sum of the line
    (weight 2 on the reflect-mirrored samples), c = (K[0] - K[N/2]) / N."""
    import blur_algorithms_b200 as bb
    rows, cols, sigma = 24, 97, 4.0
    img = np.random.default_rng(5).integers(0, 256, (rows, cols, 3), dtype=np.uint8)
    _, rp_exact, _ = oracle.gauss_blur_1d(img, sigma, 0, want_f32=True)
    _, rp_quirk, _ = oracle.gauss_blur_1d(img, sigma, 1, want_f32=True)
    g = bb.gauss_geometry(rows, cols, sigma)
    n, pad = g["n_row"], g["pad"]
    taps = bb.getGaussian(sigma, g["kSize"], n).astype(np.float64)
    sign = np.where(np.arange(n) % 2 == 0, 1.0, -1.0)
    c = (taps.sum() - (taps * sign).sum()) / n
    x = np.arange(cols)
    w = 1.0 + ((x >= 1) & (x <= pad)) + ((x >= cols - 1 - pad) & (x <= cols - 2))
    sx = np.where(x % 2 == 0, 1.0, -1.0)
    planes = img.astype(np.float64).transpose(2, 0, 1)               # [3, rows, cols]
    A = (planes * (w * sx)).sum(axis=2)                              # [3, rows]
    want = rp_exact.astype(np.float64) + c * A[:, :, None] * sx[None, None, :]
    assert np.abs(want - rp_quirk).max() < 2e-3
    assert np.abs(rp_quirk - rp_exact).max() > 10 * 2e-3             # the correction is not negligible here
