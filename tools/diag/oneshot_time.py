"""One-shot drop-in call (pffft_-style, host image in place) repeated on the same geometry: first call builds the plan."""
import os, sys, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import blur_algorithms_b200 as bb
for rows, cols, sigma in ((1080, 1920, 5.0), (2160, 3840, 25.0)):
    img = np.random.default_rng(0).integers(0, 256, (rows, cols, 3), dtype=np.uint8)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter(); bb.pocketfft_1D(img, sigma); ts.append((time.perf_counter() - t0) * 1e3)
    print(rows, cols, sigma, "ms per call:", [round(t, 2) for t in ts])
