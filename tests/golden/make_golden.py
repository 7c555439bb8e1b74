"""Regenerates tests/golden/ fixtures.  Run in the build container (needs /root/reference and
oracle/_ref built with `make -C oracle ref`); the fixtures travel to the GPU box, the reference
does not.

  1. spectrum pairs: two of the reference's own DFT_image golden pairs
     (test_images/spectrum_analysis) — inputs copied byte-for-byte, one
     spectrum in full and one as a 256-row centre crop to keep the repo small.
  2. reference_outputs.npz: outputs of the reference's own pipeline code (Source.cpp compiled
     unmodified against the API shims, oracle/_ref) on seeded inputs: Test() flags 5
     (pocketfft_1D), 3 (pffft_), 2 (pocketfft_2D), 4 (fastboxblur).
  3. known_answers.json: values printed by the reference's verbatim helpers.
"""
import json
import os
import shutil
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import Reference  # noqa: E402

BASE = "/root/reference/test_images/spectrum_analysis"


def synth(kind: str, rows: int, cols: int, seed: int = 0) -> np.ndarray:
    """Deterministic inputs shared with tests/util.py (keep in sync)."""
    rng = np.random.default_rng(seed)
    if kind == "noise":
        return rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8)
    if kind == "ramp":
        y, x = np.mgrid[0:rows, 0:cols]
        return np.stack([(x * 255 // max(cols - 1, 1)), (y * 255 // max(rows - 1, 1)), ((x + y) % 256)], -1).astype(np.uint8)
    if kind == "stripes":
        img = np.zeros((rows, cols, 3), np.uint8)
        img[:, ::2, 0] = 255
        img[::2, :, 1] = 255
        img[::3, ::3, 2] = 255
        return img
    if kind == "impulse":
        img = np.zeros((rows, cols, 3), np.uint8)
        img[rows // 2, cols // 2] = 255
        img[0, 0] = 255
        img[rows - 1, cols - 1] = 255
        return img
    raise ValueError(kind)


CASES = [  # (kind, rows, cols, sigma)
    ("noise", 64, 96, 2.0),
    ("noise", 100, 75, 3.0),
    ("ramp", 120, 160, 4.0),
    ("stripes", 96, 128, 2.5),
    ("impulse", 80, 80, 3.0),
    ("noise", 240, 320, 5.0),
    ("noise", 135, 241, 6.0),   # odd dimensions
]
BOX_CASES = [  # (kind, rows, cols, nsmooth) -> ksize = nsmooth^2, passes 2 (Source.cpp:587)
    ("noise", 64, 96, 3.0),
    ("noise", 100, 75, 2.0),
    ("stripes", 96, 128, 4.0),
    ("noise", 135, 241, 5.0),
]


def main() -> None:
    ref = Reference()
    out = {}
    for kind, rows, cols, sigma in CASES:
        img = synth(kind, rows, cols)
        key = f"{kind}_{rows}x{cols}_s{sigma}"
        out[key + "_pocketfft1d"] = ref.test(img, 5, sigma)
        out[key + "_pffft"] = ref.test(img, 3, sigma)
        out[key + "_pocketfft2d"] = ref.test(img, 2, sigma)
    for kind, rows, cols, ns in BOX_CASES:
        img = synth(kind, rows, cols)
        out[f"{kind}_{rows}x{cols}_n{ns}_box"] = ref.test(img, 4, ns)
    np.savez_compressed(os.path.join(HERE, "reference_outputs.npz"), **out)

    # spectrum pairs
    shutil.copy(f"{BASE}/blur/spatial domain/blur_radius200.png", os.path.join(HERE, "spectrum_in_blur200.png"))
    shutil.copy(f"{BASE}/blur/time domain/radius_200.png", os.path.join(HERE, "spectrum_out_blur200.png"))
    shutil.copy(f"{BASE}/noisy/spatial domain/noise_gauss200.png", os.path.join(HERE, "spectrum_in_noise200.png"))
    g = cv2.imread(f"{BASE}/noisy/time domain/200.png", cv2.IMREAD_COLOR)
    cv2.imwrite(os.path.join(HERE, "spectrum_out_noise200_rows384_640.png"), g[384:640])

    ka = {
        "gaussian_window": {str(s): ref.gaussian_window(s) for s in [0.3, 0.5, 1, 2, 3, 5, 10, 25, 38.729833, 50, 100, 106.770783]},
        "gaussian_window_max50": {str(m): ref.gaussian_window(50, m) for m in [100, 200, 330, 331, 332]},
        "nearest_transform_size": {str(n): ref.nearest_transform_size(n) for n in
                                   [1, 31, 32, 33, 100, 1000, 1110, 1256, 1756, 1950, 2324, 4004, 4126, 4426, 8308, 12108, 16714, 20000, 33098]},
        "valid_sizes_upto_4096": [n for n in range(0, 4097) if ref.is_valid_size(n)],
        "gaussian_s1_w5": [float(v) for v in ref.get_gaussian(1.0, 5, 0)],
        "gaussian_s2_w11_n32": [float(v) for v in ref.get_gaussian(2.0, 11, 32)],
        "gaussian_s5_w31_n64": [float(v) for v in ref.get_gaussian(5.0, 31, 64)],
        "gaussian_s50_w331_n17280_head": [float(v) for v in ref.get_gaussian(50.0, 331, 17280)[:170]],
    }
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump(ka, f, indent=1)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
