"""Command-line front end with the reference's argument order (Source.cpp:611-641):

    python -m blur_algorithms_b200.cli FLAG NSMOOTH FILE [-o OUT] [--sweep] [--compare-cv2]

FLAG (Source.cpp:613-619, dispatcher `Test` :574-608):
    5 pocketfft_1D   4 fastboxblur (ksize = nsmooth*nsmooth, 2 passes, :586-587)   3 pffft_
    2 pocketfft_2D   1 OpenCV (the reference's own CPU baseline: only available when cv2 is
                       importable, and never routed through this package's kernels)
The blurred image is written to OUT (default: FILE with a `_blur` suffix), like the reference's
final `imwrite` (:636), and the elapsed milliseconds are printed like its `printf`s.

--sweep reproduces the commented benchmark loop (:627-634): 45 sizes starting at 1500x1000,
growing by 225x150, sigma = sqrt(rows); one JSON line per size (device-resident and host-call ms).
FILE may be `synthetic` (uniform noise, default_rng(0)).

Image files: PNG/JPEG/... through cv2 when it is importable, binary PPM (P6) otherwise.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time

import numpy as np

NAMES = {5: "pocketfft_1D", 4: "fastboxblur", 3: "pffft_", 2: "pocketfft_2D", 1: "OpenCV blur"}


def _cv2():
    try:
        import cv2
        return cv2
    except Exception:
        return None


def read_image(path: str) -> np.ndarray:
    """uint8 [rows, cols, 3], BGR like cv::imread (PPM fallback: RGB order swapped to BGR)."""
    if path == "synthetic":
        return np.random.default_rng(0).integers(0, 256, (1500, 1000, 3), dtype=np.uint8)
    cv2 = _cv2()
    if cv2 is not None and not path.lower().endswith(".ppm"):
        img = cv2.imread(path, cv2.IMREAD_COLOR)
        if img is None:
            raise SystemExit(f"cannot read {path}")
        return np.ascontiguousarray(img)
    with open(path, "rb") as f:
        data = f.read()
    tok, pos = [], 0
    while len(tok) < 4:                      # magic, width, height, maxval (comments allowed)
        while data[pos:pos + 1].isspace():
            pos += 1
        if data[pos:pos + 1] == b"#":
            pos = data.index(b"\n", pos) + 1
            continue
        end = pos
        while not data[end:end + 1].isspace():
            end += 1
        tok.append(data[pos:end])
        pos = end
    if tok[0] != b"P6" or int(tok[3]) != 255:
        raise SystemExit("only binary PPM (P6, maxval 255) is supported without cv2")
    w, h = int(tok[1]), int(tok[2])
    rgb = np.frombuffer(data, np.uint8, w * h * 3, pos + 1).reshape(h, w, 3)
    return np.ascontiguousarray(rgb[:, :, ::-1])


def write_image(path: str, bgr: np.ndarray) -> None:
    cv2 = _cv2()
    if cv2 is not None and not path.lower().endswith(".ppm"):
        cv2.imwrite(path, bgr)
        return
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (bgr.shape[1], bgr.shape[0]))
        f.write(np.ascontiguousarray(bgr[:, :, ::-1]).tobytes())


def resize(img: np.ndarray, rows: int, cols: int) -> np.ndarray:
    """cv::resize stand-in for the sweep (nearest neighbour without cv2: timing does not depend on content)."""
    cv2 = _cv2()
    if cv2 is not None:
        return np.ascontiguousarray(cv2.resize(img, (cols, rows)))
    ri = (np.arange(rows) * img.shape[0] // rows)[:, None]
    ci = (np.arange(cols) * img.shape[1] // cols)[None, :]
    return np.ascontiguousarray(img[ri, ci])


def run_test(image: np.ndarray, flag: int, nsmooth: float, quiet: bool = False, boxblur: bool = False) -> float:
    """`Test(image, flag, nsmooth)` (Source.cpp:574-608): blurs `image` in place, returns elapsed ms.
    boxblur: the reference built with `#define boxblur` (Source.cpp:12) — flags 5 / 3 / 2 convolve with box_kernel's tent
    through the FFT kernels, flag 1 is cv::blur twice (:597-600)."""
    import blur_algorithms_b200 as bb
    t0 = time.perf_counter()
    if boxblur and flag in (5, 3, 2):
        bb.boxfft_blur(image, nsmooth, bb.NYQUIST_PFFFT if flag == 3 else bb.NYQUIST_EXACT)
    elif boxblur and flag not in (4,):
        cv2 = _cv2()
        if cv2 is None:
            raise SystemExit("flag 1 is the reference's OpenCV baseline; cv2 is not importable here")
        k = int(nsmooth) * int(nsmooth)
        for _ in range(2):
            cv2.blur(image, (k, k), dst=image, borderType=cv2.BORDER_REFLECT_101)
    elif flag == 5:
        bb.pocketfft_1D(image, nsmooth)
    elif flag == 4:
        rows, cols = image.shape[:2]
        bb.fastboxblur(image, cols, rows, image.shape[2], int(nsmooth * nsmooth), 2)
    elif flag == 3:
        bb.pffft_(image, nsmooth)
    elif flag == 2:
        bb.pocketfft_2D(image, nsmooth)
    else:
        cv2 = _cv2()
        if cv2 is None:
            raise SystemExit("flag 1 is the reference's OpenCV baseline; cv2 is not importable here")
        cv2.GaussianBlur(image, (0, 0), nsmooth, dst=image)
    ms = (time.perf_counter() - t0) * 1e3
    if not quiet:
        print(f"{NAMES.get(flag, 'OpenCV blur')}: {ms:f}")
    return ms


def sweep(image: np.ndarray, flag: int) -> None:
    """The reference's 45-size benchmark loop (Source.cpp:627-634) with sigma = sqrt(rows)."""
    import torch
    import blur_algorithms_b200 as bb
    x, y = 1500, 1000
    for i in range(45):
        img = resize(image, x, y)
        sigma = math.sqrt(x)
        line = {"i": i, "rows": x, "cols": y, "sigma": sigma, "flag": flag, "name": NAMES.get(flag)}
        if flag in (2, 3, 5):
            plan = bb.GaussPlan(x, y, sigma, bb.NYQUIST_PFFFT if flag == 3 else bb.NYQUIST_EXACT)
            d = torch.from_numpy(img).cuda()
            o = torch.empty_like(d)
            for _ in range(2):
                plan.run(d, o)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                plan.run(d, o)
            e1.record()
            torch.cuda.synchronize()
            line["device_ms"] = e0.elapsed_time(e1) / 5
            h = img.copy()
            t0 = time.perf_counter()
            plan.run_host(h)
            line["host_call_ms"] = (time.perf_counter() - t0) * 1e3
            plan.close()
        else:
            t0 = time.perf_counter()
            # Source.cpp:627-634 calls Test(image, flag, sqrt(x)) for every flag: FastBoxBlur gets ksize = int(nsmooth^2) ~ x
            try:
                run_test(img, flag, sigma, quiet=True)
            except Exception as e:                      # noqa: BLE001  (a window wider than the kernels take: record it, go on)
                line["error"] = str(e)
                print(json.dumps(line), flush=True)
                x += 225
                y += 150
                continue
            line["host_call_ms"] = (time.perf_counter() - t0) * 1e3
            if flag == 4:
                k = 2 * ((int(sigma * sigma) - 1) // 2) + 1
                line["ksize"] = k
                # k <= 127: packed tile kernel; wider windows: tile kernel while the staged rows fit shared memory, else the two-kernel path
                line["box_path"] = "packed tile kernel" if k <= 127 else "wide-window paths (boxblur.cu dispatcher)"
        line["megapixels_per_s"] = x * y / 1e3 / line.get("device_ms", line["host_call_ms"])
        print(json.dumps(line), flush=True)
        x += 225
        y += 150


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("flag", type=int)
    ap.add_argument("nsmooth", type=float)
    ap.add_argument("file")
    ap.add_argument("-o", "--out")
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--boxblur", action="store_true", help="the reference's `#define boxblur` build: tent kernel through the FFT path")
    ap.add_argument("--compare-cv2", action="store_true",
                    help="also run cv2.GaussianBlur / cv2.blur on the input and report the difference")
    a = ap.parse_args(argv)
    image = read_image(a.file)
    if a.sweep:
        sweep(image, a.flag)
        return 0
    src = image.copy() if a.compare_cv2 else None
    run_test(image, a.flag, a.nsmooth, boxblur=a.boxblur)
    out = a.out or (os.path.splitext(a.file)[0] + "_blur" + (os.path.splitext(a.file)[1] or ".ppm"))
    if a.file == "synthetic" and not a.out:
        out = "synthetic_blur.ppm"
    write_image(out, image)
    if a.compare_cv2:
        cv2 = _cv2()
        if cv2 is None:
            print("cv2 not importable: no comparison", file=sys.stderr)
        else:
            if a.flag == 4:
                k = int(a.nsmooth * a.nsmooth)
                ref = cv2.blur(cv2.blur(src, (k, k), borderType=cv2.BORDER_REFLECT_101), (k, k), borderType=cv2.BORDER_REFLECT_101)
            else:
                ref = cv2.GaussianBlur(src, (0, 0), a.nsmooth, borderType=cv2.BORDER_REFLECT_101)
            d = np.abs(ref.astype(np.int16) - image.astype(np.int16))
            # expected: OpenCV truncates the Gaussian at a different width and filters uint8 images in
            # fixed point; cv2.blur rounds once per 2-D pass, FastBoxBlur after each 1-D pass
            print(json.dumps({"vs_cv2_max_abs": int(d.max()), "vs_cv2_mean_abs": float(d.mean()),
                              "vs_cv2_frac_gt1": float((d > 1).mean())}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
