"""Shared deterministic inputs (keep `synth` in sync with tests/golden/make_golden.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


def synth(kind: str, rows: int, cols: int, seed: int = 0) -> np.ndarray:
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
    if kind == "const":
        return np.full((rows, cols, 3), 200, np.uint8)
    raise ValueError(kind)


GOLDEN_CASES = [
    ("noise", 64, 96, 2.0), ("noise", 100, 75, 3.0), ("ramp", 120, 160, 4.0), ("stripes", 96, 128, 2.5),
    ("impulse", 80, 80, 3.0), ("noise", 240, 320, 5.0), ("noise", 135, 241, 6.0),
]
GOLDEN_BOX_CASES = [("noise", 64, 96, 3.0), ("noise", 100, 75, 2.0), ("stripes", 96, 128, 4.0), ("noise", 135, 241, 5.0)]


# BASELINE.json configs 1 and 2: outputs of the reference itself (tests/golden/make_golden_large.py)
LARGE_GOLDEN_CASES = [(1080, 1920, 5.0), (2160, 3840, 25.0)]


def lattice(img: np.ndarray) -> np.ndarray:
    """The sampled pixels stored per large golden case: a 37 x 53 lattice over the image plus the
    first / last two rows and columns on a finer lattice (borders are where reflect-101 lives)."""
    rows, cols = img.shape[:2]
    parts = [img[::37, ::53].reshape(-1, 3), img[:2, ::7].reshape(-1, 3), img[rows - 2:, ::7].reshape(-1, 3),
             img[::5, :2].reshape(-1, 3), img[::5, cols - 2:].reshape(-1, 3)]
    return np.concatenate(parts)


def large_golden():
    return np.load(os.path.join(GOLDEN, "reference_large.npz"))


def golden_outputs():
    return np.load(os.path.join(GOLDEN, "reference_outputs.npz"))


def direct_gauss_f64(oracle, img: np.ndarray, sigma: float):
    """Independent float64 direct convolution with np.pad(mode='reflect')."""
    rows, cols = img.shape[:2]
    k = oracle.gaussian_window(sigma, max(rows, cols))
    pad = (k - 1) // 2
    g = oracle.get_gaussian(sigma, k, 0).astype(np.float64)
    x = img.astype(np.float64)
    xp = np.pad(x, ((0, 0), (pad, pad), (0, 0)), mode="reflect")
    y = sum(g[i] * xp[:, i:i + cols] for i in range(k))
    yp = np.pad(y, ((pad, pad), (0, 0), (0, 0)), mode="reflect")
    z = sum(g[i] * yp[i:i + rows] for i in range(k))
    return y, z


def box_numpy(img: np.ndarray, ksize: int, passes: int) -> np.ndarray:
    """Slow, obviously-correct restatement of the declared box spec."""
    r = (ksize - 1) // 2
    k = 2 * r + 1
    out = img.astype(np.int64)
    for _ in range(passes):
        for axis in (1, 0):
            padw = [(0, 0)] * out.ndim
            padw[axis] = (r, r)
            p = np.pad(out, padw, mode="reflect")
            acc = np.zeros_like(out)
            n = out.shape[axis]
            for i in range(k):
                sl = [slice(None)] * out.ndim
                sl[axis] = slice(i, i + n)
                acc += p[tuple(sl)]
            out = (2 * acc + k) // (2 * k)
    return out.astype(np.uint8)


class Emulator:
    """tests/emul/libemul.so: the device code of the product run on the CPU (test infra)."""

    def __init__(self):
        so = os.path.join(HERE, "emul", "libemul.so")
        src = os.path.join(HERE, "emul", "emulate.cpp")
        csrc = os.path.join(os.path.dirname(HERE), "blur_algorithms_b200", "csrc")
        deps = [src] + [os.path.join(csrc, f) for f in ("fft_core.cuh", "fft_plan.hpp", "blur_io.cuh", "dft_codelets.cuh", "ols_core.cuh", "ols_plan.hpp",
                                                           "dft_codelets2.cuh", "cplx2.cuh", "ols2_core.cuh", "dft_codelets3.cuh")]
        if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps if os.path.exists(d)):
            subprocess.run(["g++", "-O2", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-o", so, src], check=True)
        L = self.lib = C.CDLL(so)
        f32p = np.ctypeslib.ndpointer(np.float32, flags="C")
        u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
        i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
        L.emul_fft_roundtrip.argtypes = [C.c_int, f32p, f32p, f32p, f32p, f32p, f32p]
        L.emul_gauss.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_void_p, i32p, i32p]
        L.emul_ols_gauss.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_void_p, i32p]
        L.emul_row_pass.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]
        L.emul_col_pass.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]

    def gauss(self, img, sigma, mode=0, nparts=1):
        rows, cols = img.shape[:2]
        out = np.zeros_like(img)
        rp = np.zeros((3, rows, cols), np.float32)
        rr, rc = np.zeros(4, np.int32), np.zeros(4, np.int32)
        code = self.lib.emul_gauss(np.ascontiguousarray(img).reshape(-1), out.reshape(-1), rows, cols, sigma, mode,
                                   nparts, rp.ctypes.data, rr, rc)
        assert code == 0, code
        return out, rp, rr, rc

    def ols_gauss(self, img, sigma, nparts=1):
        """overlap-save path (the plan's default for nyquist_mode 0); returns (out, rowpass, info)"""
        rows, cols = img.shape[:2]
        out = np.zeros_like(img)
        rp = np.zeros((3, rows, cols), np.float32)
        info = np.zeros(4, np.int32)
        code = self.lib.emul_ols_gauss(np.ascontiguousarray(img).reshape(-1), out.reshape(-1), rows, cols, sigma,
                                       nparts, rp.ctypes.data, info)
        assert code == 0, code
        return out, rp, info

    def fft_roundtrip(self, a, b):
        N = a.size
        outr, outi, sr, si = (np.zeros(N, np.float32) for _ in range(4))
        code = self.lib.emul_fft_roundtrip(N, a, b, outr, outi, sr, si)
        return code, outr, outi, sr + 1j * si
