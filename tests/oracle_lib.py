"""ctypes bindings to the CPU oracle (oracle/liboracle.so) and, when it has been built,
to the reference compiled verbatim against API shims (oracle/_ref/libblurref*.so).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build_oracle(with_ref: bool | None = None) -> None:
    """Compile liboracle.so (always) and oracle/_ref (only where /root/reference exists)."""
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True, env=env)
    if with_ref is None:
        with_ref = os.path.exists("/root/reference/Source.cpp")
    if with_ref:
        subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True, env=env)


class Oracle:
    """oracle/liboracle.so — our C restatement."""

    def __init__(self) -> None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle(with_ref=False)
        L = self.lib = C.CDLL(path)
        L.orc_gaussian_window.argtypes = [C.c_double, C.c_int]
        L.orc_gaussian_window.restype = C.c_int
        L.orc_get_gaussian.argtypes = [_f32p, C.c_double, C.c_int, C.c_int]
        L.orc_get_gaussian.restype = None
        L.orc_is_valid_size.argtypes = [C.c_int]
        L.orc_is_valid_size.restype = C.c_int
        L.orc_nearest_transform_size.argtypes = [C.c_int]
        L.orc_nearest_transform_size.restype = C.c_int
        L.orc_deinterleave_u8_f32.argtypes = [_u8p, _f32p, _f32p, _f32p, C.c_uint32]
        L.orc_interleave_f32_u8.argtypes = [_f32p, _f32p, _f32p, _u8p, C.c_uint32]
        L.orc_reflect_101_u8.argtypes = [_u8p, _u8p] + [C.c_int] * 7
        L.orc_flip_block_f32.argtypes = [_f32p, _f32p, C.c_int, C.c_int]
        L.orc_gauss_sizes.argtypes = [C.c_int, C.c_int, C.c_double, _i32p]
        L.orc_gauss_blur_1d_f32.argtypes = [_u8p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_gauss_blur_1d_f32.restype = C.c_int
        L.orc_gauss_blur_1d.argtypes = [_u8p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_void_p]
        L.orc_gauss_blur_1d.restype = C.c_int
        L.orc_gauss_blur_2d.argtypes = [_u8p, C.c_int, C.c_int, C.c_double]
        L.orc_gauss_blur_2d.restype = C.c_int
        L.orc_dft_image.argtypes = [_u8p, _u8p, C.c_int, C.c_int, C.c_double, C.c_void_p, _i32p]
        L.orc_dft_image.restype = C.c_int
        L.orc_fastboxblur.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_fastboxblur.restype = C.c_int
        L.orc_num_threads.restype = C.c_int
        L.orc_set_num_threads.argtypes = [C.c_int]
        L.orc_set_boxblur.argtypes = [C.c_int]
        L.orc_box_kernel_1d.argtypes = [_f32p, C.c_int, C.c_int]

    # --- sizing / taps
    def gaussian_window(self, sigma: float, max_width: int = 0) -> int:
        return self.lib.orc_gaussian_window(sigma, max_width)

    def get_gaussian(self, sigma: float, width: int = 0, fft_length: int = 0) -> np.ndarray:
        if not width:
            width = self.gaussian_window(sigma)
        out = np.zeros(max(width, fft_length), np.float32)
        self.lib.orc_get_gaussian(out, sigma, width, fft_length)
        return out

    def is_valid_size(self, n: int) -> int:
        return self.lib.orc_is_valid_size(n)

    def nearest_transform_size(self, n: int) -> int:
        return self.lib.orc_nearest_transform_size(n)

    def gauss_sizes(self, rows: int, cols: int, sigma: float) -> dict:
        o = np.zeros(6, np.int32)
        self.lib.orc_gauss_sizes(rows, cols, sigma, o)
        return dict(kSize=int(o[0]), pad=int(o[1]), n_col=int(o[2]), n_row=int(o[3]), tz_col=int(o[4]), tz_row=int(o[5]))

    # --- layout
    def deinterleave(self, img: np.ndarray) -> np.ndarray:
        total = img.size // 3
        out = np.empty((3, total), np.float32)
        self.lib.orc_deinterleave_u8_f32(np.ascontiguousarray(img).reshape(-1), out[0], out[1], out[2], total)
        return out

    def interleave(self, planes: np.ndarray) -> np.ndarray:
        planes = np.ascontiguousarray(planes, np.float32)
        total = planes.shape[1]
        out = np.empty(total * 3, np.uint8)
        self.lib.orc_interleave_f32_u8(planes[0], planes[1], planes[2], out, total)
        return out

    def reflect_101(self, img: np.ndarray, top: int, bottom: int, left: int, right: int) -> np.ndarray:
        img = np.ascontiguousarray(img, np.uint8)
        rows, cols = img.shape[:2]
        ch = 1 if img.ndim == 2 else img.shape[2]
        t, b = min(top, rows - 1), min(bottom, rows - 1)
        l, r = min(left, cols - 1), min(right, cols - 1)
        out = np.empty((rows + t + b, cols + l + r, ch), np.uint8)
        self.lib.orc_reflect_101_u8(img.reshape(-1), out.reshape(-1), top, bottom, left, right, rows, cols, ch)
        return out if img.ndim == 3 else out[..., 0]

    def flip_block(self, plane: np.ndarray) -> np.ndarray:
        h, w = plane.shape
        out = np.empty((w, h), np.float32)
        self.lib.orc_flip_block_f32(np.ascontiguousarray(plane, np.float32), out, w, h)
        return out

    # --- pipelines
    def gauss_blur_1d(self, img: np.ndarray, sigma: float, nyquist_mode: int = 0, want_f32: bool = False):
        """Returns uint8 result; with want_f32 also (rowpass[3,rows,cols], final[3,rows,cols]) float32."""
        img = np.ascontiguousarray(img, np.uint8)
        rows, cols = img.shape[:2]
        if want_f32:
            rp = np.empty((3, rows, cols), np.float32)
            fin = np.empty((3, rows, cols), np.float32)
            rc = self.lib.orc_gauss_blur_1d_f32(img.reshape(-1), rows, cols, sigma, nyquist_mode,
                                                rp.ctypes.data, fin.ctypes.data)
            if rc:
                raise ValueError(f"oracle gauss_blur_1d rc={rc}")
            out = self.interleave(fin.reshape(3, -1)).reshape(rows, cols, 3)
            return out, rp, fin
        out = img.copy()
        rc = self.lib.orc_gauss_blur_1d(out.reshape(-1), rows, cols, sigma, nyquist_mode, None)
        if rc:
            raise ValueError(f"oracle gauss_blur_1d rc={rc}")
        return out

    def boxfft_blur(self, img: np.ndarray, nsmooth: float, nyquist_mode: int = 0) -> np.ndarray:
        """The reference's `#define boxblur` build of pocketfft_1D / pffft_ (tent kernel, Source.cpp:287-291,
        437-441, box_kernel :128-140)."""
        self.lib.orc_set_boxblur(1)
        try:
            return self.gauss_blur_1d(img, nsmooth, nyquist_mode)
        finally:
            self.lib.orc_set_boxblur(0)

    def box_kernel_1d(self, klen: int, fft_length: int) -> np.ndarray:
        out = np.zeros(fft_length, np.float32)
        self.lib.orc_box_kernel_1d(out, klen, fft_length)
        return out

    def gauss_blur_2d(self, img: np.ndarray, sigma: float) -> np.ndarray:
        out = np.ascontiguousarray(img, np.uint8).copy()
        rows, cols = out.shape[:2]
        rc = self.lib.orc_gauss_blur_2d(out.reshape(-1), rows, cols, sigma)
        if rc:
            raise ValueError(f"oracle gauss_blur_2d rc={rc}")
        return out

    def dft_image(self, img: np.ndarray, nsmooth: float = 0.1, want_f32: bool = False):
        img = np.ascontiguousarray(img, np.uint8)
        rows, cols = img.shape[:2]
        out = np.empty_like(img)
        ps = np.zeros(2, np.int32)
        if want_f32:
            # size query first
            from math import ceil  # noqa: F401
            k = self.gaussian_window(nsmooth, max(rows, cols))
            pad = (k - 1) // 2
            H, W = rows + 2 * pad, cols + 2 * pad
            H = H if self.is_valid_size(H) else self.nearest_transform_size(H)
            W = W if self.is_valid_size(W) else self.nearest_transform_size(W)
            spec = np.empty((3, H, W), np.float32)
            rc = self.lib.orc_dft_image(img.reshape(-1), out.reshape(-1), rows, cols, nsmooth, spec.ctypes.data, ps)
            if rc:
                raise ValueError(f"oracle dft_image rc={rc}")
            return out, spec
        rc = self.lib.orc_dft_image(img.reshape(-1), out.reshape(-1), rows, cols, nsmooth, None, ps)
        if rc:
            raise ValueError(f"oracle dft_image rc={rc}")
        return out

    def fastboxblur(self, img: np.ndarray, ksize: int, passes: int) -> np.ndarray:
        out = np.ascontiguousarray(img, np.uint8).copy()
        h, w = out.shape[:2]
        ch = 1 if out.ndim == 2 else out.shape[2]
        rc = self.lib.orc_fastboxblur(out.reshape(-1), w, h, ch, ksize, passes)
        if rc:
            raise ValueError(f"oracle fastboxblur rc={rc}")
        return out

    def num_threads(self) -> int:
        return self.lib.orc_num_threads()

    def set_num_threads(self, n: int) -> None:
        self.lib.orc_set_num_threads(int(n))


class Reference:
    """oracle/_ref/libblurref(.so|_dft.so): the reference's Source.cpp + Utils.hpp compiled
    unmodified against API shims.  Raises FileNotFoundError if it has not been built."""

    def __init__(self, dft_image: bool = False, boxblur: bool = False) -> None:
        name = "libblurref_box.so" if boxblur else ("libblurref_dft.so" if dft_image else "libblurref.so")
        path = os.path.join(ORACLE_DIR, "_ref", name)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.lib = C.CDLL(path)
        L.ref_gaussian_window.argtypes = [C.c_double, C.c_int]
        L.ref_gaussian_window.restype = C.c_int
        L.ref_get_gaussian.argtypes = [_f32p, C.c_double, C.c_int, C.c_int]
        L.ref_is_valid_size.argtypes = [C.c_int]
        L.ref_is_valid_size.restype = C.c_int
        L.ref_nearest_transform_size.argtypes = [C.c_int]
        L.ref_nearest_transform_size.restype = C.c_int
        L.ref_deinterleave_u8_f32.argtypes = [_u8p, _f32p, _f32p, _f32p, C.c_uint32]
        L.ref_interleave_f32_u8.argtypes = [_f32p, _f32p, _f32p, _u8p, C.c_uint32]
        L.ref_reflect_101_u8.argtypes = [_u8p, _u8p] + [C.c_int] * 7
        L.ref_test.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_double]
        L.ref_has_dft_image.restype = C.c_int
        L.ref_num_threads.restype = C.c_int
        L.ref_set_num_threads.argtypes = [C.c_int]

    @staticmethod
    def available(dft_image: bool = False, boxblur: bool = False) -> bool:
        name = "libblurref_box.so" if boxblur else ("libblurref_dft.so" if dft_image else "libblurref.so")
        return os.path.exists(os.path.join(ORACLE_DIR, "_ref", name))

    def gaussian_window(self, sigma: float, max_width: int = 0) -> int:
        return self.lib.ref_gaussian_window(sigma, max_width)

    def get_gaussian(self, sigma: float, width: int, fft_length: int = 0) -> np.ndarray:
        out = np.zeros(max(width, fft_length), np.float32)
        self.lib.ref_get_gaussian(out, sigma, width, fft_length)
        return out

    def is_valid_size(self, n: int) -> int:
        return self.lib.ref_is_valid_size(n)

    def nearest_transform_size(self, n: int) -> int:
        return self.lib.ref_nearest_transform_size(n)

    def deinterleave(self, img: np.ndarray) -> np.ndarray:
        total = img.size // 3
        out = np.empty((3, total), np.float32)
        self.lib.ref_deinterleave_u8_f32(np.ascontiguousarray(img).reshape(-1), out[0], out[1], out[2], total)
        return out

    def interleave(self, planes: np.ndarray) -> np.ndarray:
        planes = np.ascontiguousarray(planes, np.float32)
        total = planes.shape[1]
        out = np.empty(total * 3, np.uint8)
        self.lib.ref_interleave_f32_u8(planes[0], planes[1], planes[2], out, total)
        return out

    def reflect_101(self, img: np.ndarray, top: int, bottom: int, left: int, right: int) -> np.ndarray:
        img = np.ascontiguousarray(img, np.uint8)
        rows, cols = img.shape[:2]
        ch = 1 if img.ndim == 2 else img.shape[2]
        t, b = min(top, rows - 1), min(bottom, rows - 1)
        l, r = min(left, cols - 1), min(right, cols - 1)
        out = np.empty((rows + t + b, cols + l + r, ch), np.uint8)
        self.lib.ref_reflect_101_u8(img.reshape(-1), out.reshape(-1), top, bottom, left, right, rows, cols, ch)
        return out if img.ndim == 3 else out[..., 0]

    def test(self, img: np.ndarray, flag: int, nsmooth: float) -> np.ndarray:
        """Test(image, flag, nsmooth) of Source.cpp:574: 5 pocketfft_1D, 4 fastboxblur, 3 pffft_,
        2 pocketfft_2D (spectrum image in the DFT_image build)."""
        out = np.ascontiguousarray(img, np.uint8).copy()
        rows, cols = out.shape[:2]
        self.lib.ref_test(out.reshape(-1), rows, cols, flag, nsmooth)
        return out

    def num_threads(self) -> int:
        return self.lib.ref_num_threads()

    def set_num_threads(self, n: int) -> None:
        """OpenMP threads of the reference build (torchrun exports OMP_NUM_THREADS=1)."""
        self.lib.ref_set_num_threads(int(n))
