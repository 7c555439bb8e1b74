#!/usr/bin/env python
"""Device-resident timings of every BASELINE.json config that is not bench.py's headline line
(C1, C2, C4, C5, plus the pffft-quirk full-length kernel), each with its HBM roofline fraction
(algorithmic bytes per pixel from SURVEY.md §8d / DESIGN.md §3).  One JSON line per config.

    python tools/bench_configs.py [--steps K] [--warmup W]

Inputs are synthetic (default_rng(0)); small configs fit in the 126 MB L2, so between timed
iterations a 256 MB buffer is overwritten to flush it (the flush is outside the event pair).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb                      # noqa: E402
from blur_algorithms_b200._lib import check, lib       # noqa: E402


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def timed(fn, steps, warmup, flush):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(steps):
        if flush is not None:
            flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    pk, pk_src = peak()
    rng = np.random.default_rng(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def emit(name, px, ms, bpp, extra=None):
        gbs = bpp * px / (ms * 1e-3) / 1e9
        line = {"config": name, "ms": ms, "megapixels_per_s": px / 1e6 / (ms * 1e-3), "alg_bytes_per_px": bpp,
                "achieved_GBps": gbs, "hbm_peak_GBps": pk, "peak_source": pk_src, "frac": gbs / pk,
                "timing": f"median of {a.steps} CUDA-event pairs after {a.warmup} warm-ups, L2 flushed between iterations"}
        line.update(extra or {})
        print(json.dumps(line), flush=True)

    for name, rows, cols, sigma, mode in [
            ("C1 gauss 1080x1920 sigma=5 exact (pocketfft_1D semantics)", 1080, 1920, 5.0, bb.NYQUIST_EXACT),
            ("C1 gauss 1080x1920 sigma=5 pffft quirk (pffft_ semantics)", 1080, 1920, 5.0, bb.NYQUIST_PFFFT),
            ("C2 gauss 2160x3840 sigma=25 exact", 2160, 3840, 25.0, bb.NYQUIST_EXACT),
            ("C2 gauss 2160x3840 sigma=25 pffft quirk (pffft_ semantics)", 2160, 3840, 25.0, bb.NYQUIST_PFFFT),
            ("C3 gauss 16384x16384 sigma=50 pffft quirk (pffft_ semantics)", 16384, 16384, 50.0, bb.NYQUIST_PFFFT)]:
        d_in = torch.from_numpy(rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8)).to(dev)
        d_out = torch.empty_like(d_in)
        plan = bb.GaussPlan(rows, cols, sigma, mode, device=0)
        ms = timed(lambda: plan.run(d_in, d_out), a.steps, a.warmup, flush)
        emit(name, rows * cols, ms, 30.0, {"plan": {k: plan.info[k] for k in ("algo", "n_row", "n_col", "ols_L_row", "ols_L_col")}})
        plan.close()
        del d_in, d_out

    # C4: DFT_image, 4096x4096 (three planes at once through the 3-channel entry; per-plane pixels reported)
    rows = cols = 4096
    d_in = torch.from_numpy(rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8)).to(dev)
    d_out = torch.empty_like(d_in)
    padded = (C.c_int * 2)()
    for nsmooth, tag in ((0.1, "pad=0 (4096x4096 transforms)"), (5.0, "nsmooth=5 (4320x4320 transforms)")):
        fn = lambda: check(lib.blur_b200_dft_image_device(d_in.data_ptr(), d_out.data_ptr(), rows, cols, nsmooth, None, padded, stream))  # noqa: E731
        ms = timed(fn, a.steps, a.warmup, flush)
        emit(f"C4 DFT_image 4096x4096 x3 planes, {tag}", 3 * padded[0] * padded[1], ms, 10.0,
             {"note": "pixels = 3 planes x padded size; 10 B/px = 1 (u8 in) + 8 (complex64 half spectrum written+read) + 1 (u8 out)"})
    del d_in, d_out

    # C5: FastBoxBlur batch of 256 1080p RGB images, radius 15 (ksize 31), 2 passes
    w, h, batch = 1920, 1080, 256
    d_img = torch.from_numpy(rng.integers(0, 256, (batch, h, w, 3), dtype=np.uint8)).to(dev)
    d_scr = torch.empty_like(d_img)
    ms = timed(lambda: bb.fastboxblur_device(d_img, w, h, 3, 31, 2, batch, d_scr), max(3, a.steps // 4), 2, None)
    emit("C5 fastboxblur 256 x 1080x1920 RGB ksize=31 passes=2", batch * w * h, ms, 6.0,
         {"note": "in place, so every timed iteration blurs the previous result (data-independent cost); 3.2 GB batch > L2"})
    # the same batch end to end: pinned host images, uploads / kernels / downloads overlapped
    import time
    h_imgs = d_img.cpu().pin_memory()
    del d_img, d_scr
    bb.fastboxblur_batch(h_imgs, 31, 2)
    t0 = time.perf_counter()
    bb.fastboxblur_batch(h_imgs, 31, 2)
    dt = time.perf_counter() - t0
    print(json.dumps({"config": "C5 fastboxblur 256 x 1080x1920 RGB, end to end from pinned host memory (blur_b200_fastboxblur_batch_host)",
                      "ms": dt * 1e3, "megapixels_per_s": batch * w * h / 1e6 / dt, "h2d_bytes": int(h_imgs.numel()),
                      "d2h_bytes": int(h_imgs.numel()), "timing": "wall clock around one synchronous call"}), flush=True)


if __name__ == "__main__":
    main()
