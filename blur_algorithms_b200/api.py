"""Host-side mirror of the reference's entry points for its blur hot path.

Same names, argument meaning and in-place behaviour as michelerenzullo/Blur_algorithms
(Source.cpp / Utils.hpp); the work is done by hand-written sm_100a kernels behind the C ABI in
include/blur_b200.h.  Images are numpy uint8 arrays [rows, cols, 3] (what cv::imread returns) and
are modified IN PLACE like `image.data` in the reference; torch CUDA tensors are accepted by the
`GaussPlan` / `*_device` entry points for device-resident use.

There is no CPU fallback: every compute call needs a CUDA device and raises BlurB200Error
otherwise.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import NYQUIST_EXACT, NYQUIST_PFFFT, BlurB200Error, check, lib

__all__ = [
    "isValidSize", "nearestTransformSize", "gaussian_window", "getGaussian", "gauss_geometry",
    "pffft_", "pocketfft_1D", "pocketfft_2D", "boxfft_blur", "boxfft_klen", "fastboxblur", "Reflect_101", "deinterleave_BGR",
    "interleave_BGR", "flip_block", "dft_image", "Test", "GaussPlan", "fastboxblur_device", "fastboxblur_device_noalloc", "fastboxblur_batch",
    "kernel_launches", "BlurB200Error", "NYQUIST_EXACT", "NYQUIST_PFFFT",
]


# ---- sizing helpers (host) --------------------------------------------------------------------

def isValidSize(N: int) -> int:
    """Utils.hpp:141-148."""
    return lib.blur_b200_is_valid_size(int(N))


def nearestTransformSize(N: int) -> int:
    """Utils.hpp:150-157."""
    return lib.blur_b200_nearest_transform_size(int(N))


def gaussian_window(sigma: float, max_width: int = 0) -> int:
    """Source.cpp:60-73."""
    return lib.blur_b200_gaussian_window(float(sigma), int(max_width))


def getGaussian(sigma: float, width: int = 0, FFT_length: int = 0) -> np.ndarray:
    """Source.cpp:75-102.  Returns the float32 taps (rotated so the centre is element 0 when
    FFT_length is given) instead of filling a caller-provided container."""
    if not width:
        width = gaussian_window(sigma)
    out = np.zeros(max(width, FFT_length), np.float32)
    check(lib.blur_b200_get_gaussian(out.ctypes.data, float(sigma), int(width), int(FFT_length)))
    return out


def gauss_geometry(rows: int, cols: int, sigma: float) -> dict:
    """kSize / pad / FFT lengths of the 1-D tile path, Source.cpp:434-457."""
    o = (C.c_int * 4)()
    check(lib.blur_b200_gauss_geometry(rows, cols, float(sigma), o))
    return dict(kSize=o[0], pad=o[1], n_col=o[2], n_row=o[3])


# ---- helpers ------------------------------------------------------------------------------------

def _as_image(image: np.ndarray) -> np.ndarray:
    if not isinstance(image, np.ndarray) or image.dtype != np.uint8 or image.ndim != 3 or image.shape[2] != 3:
        raise BlurB200Error(_lib.EINVAL, "image must be a numpy uint8 array of shape [rows, cols, 3]")
    if not image.flags["C_CONTIGUOUS"] or not image.flags["WRITEABLE"]:
        raise BlurB200Error(_lib.EINVAL, "image must be C-contiguous and writeable (it is modified in place)")
    return image


def _ptr(t) -> int:
    """data pointer of a torch tensor / numpy array / int"""
    if isinstance(t, int):
        return t
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


def _stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


# ---- Gaussian blur ------------------------------------------------------------------------------

def pffft_(image: np.ndarray, nsmooth: float) -> None:
    """pffft_(cv::Mat& image, double nsmooth), Source.cpp:429-570.  In place.  Reproduces the
    Nyquist-bin behaviour of pffft_sorted_optimized_convolution (Source.cpp:414-427)."""
    img = _as_image(image)
    check(lib.blur_b200_gauss_u8c3(img.ctypes.data, img.shape[0], img.shape[1], float(nsmooth), NYQUIST_PFFFT))


def pocketfft_1D(image: np.ndarray, nsmooth: float) -> None:
    """pocketfft_1D(cv::Mat& image, double nsmooth), Source.cpp:280-392.  In place."""
    img = _as_image(image)
    check(lib.blur_b200_gauss_u8c3(img.ctypes.data, img.shape[0], img.shape[1], float(nsmooth), NYQUIST_EXACT))


def boxfft_klen(rows: int, cols: int, nsmooth: float) -> int:
    """kLen of the reference's `#define boxblur` build: (int)(ns'^2), ns' = sqrt(min((int)nsmooth^2, min(rows-1, cols-1)))
    (Source.cpp:287-291)."""
    return lib.blur_b200_boxfft_klen(int(rows), int(cols), float(nsmooth))


def boxfft_blur(image: np.ndarray, nsmooth: float, nyquist_mode: int = NYQUIST_EXACT) -> None:
    """pocketfft_1D (nyquist_mode EXACT) / pffft_ (PFFFT) of the reference built with `#define boxblur`
    (Source.cpp:12): FFT convolution with a box of width kLen convolved with itself (box_kernel, :128-140).  In place."""
    img = _as_image(image)
    check(lib.blur_b200_boxfft_u8c3(img.ctypes.data, img.shape[0], img.shape[1], float(nsmooth), nyquist_mode))


def pocketfft_2D(image: np.ndarray, nsmooth: float) -> None:
    """pocketfft_2D(cv::Mat& image, double nsmooth), Source.cpp:143-277 (blur branch).  In place.
    The 2-D formulation computes the same separable reflect-101 convolution as pocketfft_1D and
    differs from it only by float rounding, so it runs on the same kernels."""
    pocketfft_1D(image, nsmooth)


class GaussPlan:
    """Cached plan (twiddles, separable kernel spectra, workspaces) for device-resident blurs of
    one (rows, cols, sigma) problem.  Replaces the per-call setup of Source.cpp:465-503."""

    def __init__(self, rows: int, cols: int, sigma: float, nyquist_mode: int = NYQUIST_EXACT,
                 device: int = 0, rank: int = 0, nranks: int = 1, boxblur: bool = False, direct: bool | None = None):
        """boxblur=True: the reference's `#define boxblur` build — `sigma` is then its nsmooth and the kernel a box
        of width kLen = boxfft_klen(rows, cols, nsmooth) convolved with itself (Source.cpp:128-140, 287-291).
        direct=False: a single-GPU plan of a narrow kernel (half width <= 20) keeps the FFT kernels for run() /
        run_host*() instead of the fused direct kernel, so that its output equals the sharded, halo and pass-wise
        results bit for bit (the two kernel families agree within 1 LSB only)."""
        self._h = C.c_void_p()
        self.rows, self.cols, self.sigma = rows, cols, float(sigma)
        self.rank, self.nranks = rank, nranks
        if boxblur:
            if nranks != 1:
                raise BlurB200Error(_lib.EINVAL, "the FFT box blur is a single-GPU plan")
            check(lib.blur_b200_boxfft_plan_create(C.byref(self._h), rows, cols, float(sigma), nyquist_mode, device))
        elif nranks == 1:
            check(lib.blur_b200_gauss_plan_create(C.byref(self._h), rows, cols, float(sigma), nyquist_mode, device))
        else:
            check(lib.blur_b200_gauss_plan_create_sharded(C.byref(self._h), rows, cols, float(sigma), nyquist_mode,
                                                          device, rank, nranks))
        info = (C.c_int64 * 24)()
        check(lib.blur_b200_gauss_plan_info(self._h, info))
        keys = ["kSize", "pad", "n_col", "n_row", "stripe_rows", "cpp", "row0", "valid_rows", "block_floats",
                "launches_per_run", "threads_row", "threads_col", "batch_row", "batch_col", "smem_row", "smem_col",
                "algo", "ols_L_row", "ols_L_col", "ols_V_row", "ols_V_col", "ols_nblocks_row", "ols_nblocks_col", "rqs"]
        self.info = {k: int(v) for k, v in zip(keys, info)}
        # whole-image entries on the fused direct-convolution kernel (narrow kernels, single GPU)
        if direct is not None and (not direct or lib.blur_b200_gauss_plan_uses_direct(self._h) == 0):
            if direct and nranks != 1:
                raise BlurB200Error(_lib.EINVAL, "the direct kernel is a single-GPU path")
            check(lib.blur_b200_gauss_plan_set_direct(self._h, int(bool(direct))))
        self.info["direct"] = int(lib.blur_b200_gauss_plan_uses_direct(self._h))

    def close(self) -> None:
        if self._h:
            lib.blur_b200_gauss_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def run(self, d_in, d_out=None, stream: int | None = None):
        """d_in/d_out: torch uint8 CUDA tensors [rows, cols, 3]; in place if d_out is None."""
        d_out = d_in if d_out is None else d_out
        check(lib.blur_b200_gauss_run_device(self._h, _ptr(d_in), _ptr(d_out), _stream() if stream is None else stream))
        return d_out

    def run_host(self, h_in: np.ndarray, h_out: np.ndarray | None = None):
        """numpy / pinned host buffers; H2D + kernels + D2H inside the call."""
        h_out = h_in if h_out is None else h_out
        check(lib.blur_b200_gauss_run_host(self._h, _ptr(h_in), _ptr(h_out)))
        return h_out

    def run_host_batch(self, h_in: list, h_out: list | None = None):
        """Stream a list of host images (numpy arrays or pinned CPU tensors) through the plan:
        uploads, kernels and downloads of consecutive images overlap.  In place if h_out is None."""
        h_out = h_in if h_out is None else h_out
        if len(h_in) != len(h_out):
            raise ValueError("h_in and h_out must have the same length")
        n = len(h_in)
        pin = (C.c_void_p * max(n, 1))(*[_ptr(a) for a in h_in])
        pout = (C.c_void_p * max(n, 1))(*[_ptr(a) for a in h_out])
        check(lib.blur_b200_gauss_run_host_batch(self._h, pin, pout, n))
        return h_out

    def rowpass_f32(self, d_in, d_planar, stream: int | None = None):
        """float32 row-pass result as planar [3, rows, cols] (parity hook for `resf`, Source.cpp:536)."""
        check(lib.blur_b200_gauss_rowpass_f32(self._h, _ptr(d_in), _ptr(d_planar), _stream() if stream is None else stream))
        return d_planar

    def row_pass(self, d_in_stripe, d_q, stream: int | None = None):
        check(lib.blur_b200_gauss_row_pass(self._h, _ptr(d_in_stripe), _ptr(d_q), _stream() if stream is None else stream))

    def row_pass_peer(self, d_in_stripe, q_peer_ptrs, stream: int | None = None):
        """Row pass whose epilogue stores into the peers' receive buffers (fused exchange)."""
        arr = (C.c_void_p * len(q_peer_ptrs))(*[int(p) for p in q_peer_ptrs])
        check(lib.blur_b200_gauss_row_pass_peer(self._h, _ptr(d_in_stripe), arr, len(q_peer_ptrs),
                                                _stream() if stream is None else stream))

    def row_pass_push(self, d_in_stripe, q_peer_ptrs, nchunks: int = 4, stream: int | None = None):
        """Row pass in `nchunks` groups of row pairs, each pushed to the owners' receive buffers by a
        copy kernel over NVLink while the next group is transformed (pipelined exchange)."""
        arr = (C.c_void_p * len(q_peer_ptrs))(*[int(p) for p in q_peer_ptrs])
        check(lib.blur_b200_gauss_row_pass_push(self._h, _ptr(d_in_stripe), arr, len(q_peer_ptrs), int(nchunks),
                                                _stream() if stream is None else stream))

    def col_pass(self, d_q_chunks, d_out_cols, stream: int | None = None):
        check(lib.blur_b200_gauss_col_pass(self._h, _ptr(d_q_chunks), _ptr(d_out_cols), _stream() if stream is None else stream))

    # row-stripe ("halo") sharding on a whole-image plan (blur_b200_gauss_halo_* in include/blur_b200.h)
    def halo_info(self, rank: int, nranks: int):
        """dict(row0, valid_rows, cap, halo, qpitch, q_floats, b0, nbl) of stripe `rank`, or None if the plan cannot
        run row-stripe sharded (pffft mode, three-factor kernels)."""
        o = (C.c_int64 * 8)()
        if lib.blur_b200_gauss_halo_info(self._h, int(rank), int(nranks), o) != _lib.OK:
            return None
        return dict(zip(("row0", "valid_rows", "cap", "halo", "qpitch", "q_floats", "b0", "nbl"), (int(v) for v in o)))

    def halo_row_pass(self, rank: int, nranks: int, d_in_stripe, d_q_local, stream: int | None = None):
        check(lib.blur_b200_gauss_halo_row_pass(self._h, rank, nranks, _ptr(d_in_stripe) if d_in_stripe is not None else None,
                                                _ptr(d_q_local), _stream() if stream is None else stream))

    def halo_push(self, rank: int, nranks: int, d_q_local, q_up, q_down, stream: int | None = None):
        """q_up / q_down: the neighbours' Q_local (tensor, int address or None at the image edges)"""
        check(lib.blur_b200_gauss_halo_push(self._h, rank, nranks, _ptr(d_q_local), _ptr(q_up) if q_up is not None else None,
                                            _ptr(q_down) if q_down is not None else None, _stream() if stream is None else stream))

    def halo_col_pass(self, rank: int, nranks: int, d_q_local, d_out_rows, stream: int | None = None):
        check(lib.blur_b200_gauss_halo_col_pass(self._h, rank, nranks, _ptr(d_q_local), _ptr(d_out_rows),
                                                _stream() if stream is None else stream))

    # pffft-mode sharded plans (blur_b200_gauss_quirk_* in include/blur_b200.h)
    def quirk_buffers(self):
        """(corr_row, insum): zero-copy torch views of the plan's two correction buffers (float32 [3*rows],
        int32 [nseg*3*cols]) that a sharded caller sums over the ranks between quirk_local and quirk_finish;
        None for plans without the overlap-save Nyquist corrections."""
        import torch
        pc, pi = C.c_void_p(), C.c_void_p()
        nc, ni = C.c_int64(), C.c_int64()
        if lib.blur_b200_gauss_quirk_buffers(self._h, C.byref(pc), C.byref(nc), C.byref(pi), C.byref(ni)) != _lib.OK:
            return None

        class _Dev:                                    # __cuda_array_interface__ wrapper around a raw device pointer
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}
        return (torch.as_tensor(_Dev(pc.value, nc.value, "<f4"), device="cuda"),
                torch.as_tensor(_Dev(pi.value, ni.value, "<i4"), device="cuda"))

    def quirk_local(self, d_in_stripe, stream: int | None = None):
        check(lib.blur_b200_gauss_quirk_local(self._h, _ptr(d_in_stripe) if d_in_stripe is not None else None,
                                              _stream() if stream is None else stream))

    def quirk_finish(self, stream: int | None = None):
        check(lib.blur_b200_gauss_quirk_finish(self._h, _stream() if stream is None else stream))

    # staged exchange (blur_b200_gauss_row_pass_staged & co. in include/blur_b200.h)
    def staged_supported(self, nstages: int) -> bool:
        return lib.blur_b200_gauss_staged_supported(self._h, int(nstages)) == _lib.OK

    def row_pass_staged(self, d_in_stripe, q_peer_ptrs, nstages: int, stream: int | None = None, store: bool = False,
                        row_ctas_per_sm: int = 0):
        """Row pass as `nstages` launches over column sub-ranges.  store=False: the copy engines push each finished
        sub-range to its owners under the next launch; store=True: every launch writes its sub-ranges straight into
        the owners' buffers (bulk copies over NVLink), optionally held to `row_ctas_per_sm` CTAs per SM."""
        arr = (C.c_void_p * len(q_peer_ptrs))(*[int(p) for p in q_peer_ptrs])
        check(lib.blur_b200_gauss_row_pass_staged(self._h, _ptr(d_in_stripe), arr, len(q_peer_ptrs), int(nstages), int(bool(store)),
                                                  int(row_ctas_per_sm), _stream() if stream is None else stream))

    def stage_wait(self, stage: int, stream: int | None = None):
        """`stream` waits for this rank's outgoing copies of `stage`."""
        check(lib.blur_b200_gauss_stage_wait(self._h, int(stage), _stream() if stream is None else stream))

    def col_pass_stage(self, d_q_recv, d_out_cols, stage: int, nstages: int, stream: int | None = None, col_ctas_per_sm: int = 0):
        """Column pass + interleave of sub-range `stage` of this rank's column stripe."""
        check(lib.blur_b200_gauss_col_pass_stage(self._h, _ptr(d_q_recv), _ptr(d_out_cols), int(stage), int(nstages),
                                                 int(col_ctas_per_sm), _stream() if stream is None else stream))


# ---- FastBoxBlur --------------------------------------------------------------------------------

def fastboxblur(image: np.ndarray, w: int, h: int, channels: int, ksize: int, passes: int) -> None:
    """fastboxblur(in, w, h, channels, ksize, passes), call site Source.cpp:587.  In place on a
    uint8 buffer of h*w*channels bytes."""
    if not isinstance(image, np.ndarray) or image.dtype != np.uint8 or image.size != w * h * channels:
        raise BlurB200Error(_lib.EINVAL, "image must be a numpy uint8 buffer of w*h*channels bytes")
    if not image.flags["C_CONTIGUOUS"] or not image.flags["WRITEABLE"]:
        raise BlurB200Error(_lib.EINVAL, "image must be C-contiguous and writeable")
    check(lib.blur_b200_fastboxblur(image.ctypes.data, w, h, channels, int(ksize), int(passes)))


def fastboxblur_batch(images, ksize: int, passes: int) -> None:
    """`batch` images [batch, h, w, channels] (numpy uint8 or a pinned CPU torch tensor), each blurred
    like fastboxblur(...) (Source.cpp:587), in place; uploads, kernels and downloads overlap."""
    batch, h, w, channels = images.shape
    check(lib.blur_b200_fastboxblur_batch_host(_ptr(images), w, h, channels, int(ksize), int(passes), batch))


def fastboxblur_device(d_img, w: int, h: int, channels: int, ksize: int, passes: int, batch: int, d_scratch,
                       stream: int | None = None) -> None:
    """Device-resident, batched variant: d_img is `batch` images back to back."""
    check(lib.blur_b200_fastboxblur_device(_ptr(d_img), w, h, channels, int(ksize), int(passes), batch,
                                           _ptr(d_scratch), _stream() if stream is None else stream))


def fastboxblur_device_noalloc(d_img, w: int, h: int, channels: int, ksize: int, passes: int, batch: int,
                               stream: int | None = None) -> None:
    """fastboxblur_device without the scratch argument, for callers that may not allocate GPU memory (a node
    inside another application's memory pool, a real-time loop): the second buffer is a fixed 32 MiB array
    inside the library, shared by all calls on a GPU.  Raises BlurB200Error(ESCRATCH) when
    w*h*channels*batch exceeds it.  Do not let two such calls overlap on one GPU; make the first call on each
    GPU at set-up time.  Results equal fastboxblur_device byte for byte."""
    check(lib.blur_b200_fastboxblur_device_noalloc(_ptr(d_img), w, h, channels, int(ksize), int(passes), batch,
                                                   _stream() if stream is None else stream))


# ---- layout utilities ---------------------------------------------------------------------------

def Reflect_101(image: np.ndarray, pad_top: int, pad_bottom: int, pad_left: int, pad_right: int) -> np.ndarray:
    """Reflect_101<uint8_t,C>(input, output, top, bottom, left, right, original_size),
    Utils.hpp:212-243.  Returns the padded array (pads are clamped to dim-1 like the reference)."""
    if image.dtype != np.uint8 or image.ndim not in (2, 3):
        raise BlurB200Error(_lib.EINVAL, "image must be uint8 [rows, cols] or [rows, cols, C]")
    img = np.ascontiguousarray(image)
    rows, cols = img.shape[:2]
    ch = 1 if img.ndim == 2 else img.shape[2]
    if min(pad_top, pad_bottom, pad_left, pad_right) < 0:
        raise BlurB200Error(_lib.EINVAL, "negative pad")
    t, b = min(pad_top, rows - 1), min(pad_bottom, rows - 1)
    l, r = min(pad_left, cols - 1), min(pad_right, cols - 1)
    out = np.empty((rows + t + b, cols + l + r, ch), np.uint8)
    sz = (C.c_int * 2)(rows, cols)
    check(lib.blur_b200_reflect_101_u8_host(img.ctypes.data, out.ctypes.data, pad_top, pad_bottom, pad_left,
                                            pad_right, sz, ch))
    return out if img.ndim == 3 else out[..., 0]


_TYPE_CODES = {"torch.uint8": 0, "torch.uint16": 1, "torch.int32": 2, "torch.float32": 3,
               "uint8": 0, "uint16": 1, "int32": 2, "float32": 3}


def _type_code(t) -> int:
    """Element type code of a tensor / array, or of the first of a list or tuple of planes."""
    if isinstance(t, (list, tuple)) and len(t):
        t = t[0]
    try:
        return _TYPE_CODES[str(t.dtype)]
    except (KeyError, AttributeError):
        what = getattr(t, "dtype", type(t).__name__)
        raise BlurB200Error(_lib.EINVAL, f"element type {what} is not one of uint8 / uint16 / int32 / float32") from None


def _elem_bytes(t) -> int:
    """Bytes per element of a torch tensor (element_size()) or of anything with an itemsize (numpy, cupy)."""
    if hasattr(t, "element_size"):
        return int(t.element_size())
    if hasattr(t, "itemsize"):
        return int(t.itemsize)
    raise BlurB200Error(_lib.EINVAL, f"cannot tell the element size of a {type(t).__name__}")


def deinterleave_BGR(d_interleaved, d_planes, total: int, stream: int | None = None) -> None:
    """deinterleave_BGR<T,U>, Utils.hpp:159-184: d_planes[c][x] = (U)(d_interleaved[3x+c] + round), round = 0.5f only
    for floating -> integral.  T, U from the tensors' dtypes (uint8 / uint16 / int32 / float32); the blur's own
    instantiation is <uint8_t,float>.  d_planes: CUDA tensor [3, total], or a list / tuple of three plane tensors."""
    check(lib.blur_b200_deinterleave_typed(_ptr(d_interleaved), _ptr(d_planes[0]), _ptr(d_planes[1]), _ptr(d_planes[2]), total,
                                           _type_code(d_interleaved), _type_code(d_planes),
                                           _stream() if stream is None else stream))


def interleave_BGR(d_planes, d_interleaved, total: int, stream: int | None = None) -> None:
    """interleave_BGR<T,U>, Utils.hpp:186-210: d_interleaved[3x+c] = (T)(d_planes[c][x] + round) (+0.5f and
    truncation for float -> integral)."""
    check(lib.blur_b200_interleave_typed(_ptr(d_planes[0]), _ptr(d_planes[1]), _ptr(d_planes[2]), _ptr(d_interleaved), total,
                                         _type_code(d_planes), _type_code(d_interleaved),
                                         _stream() if stream is None else stream))


def flip_block(d_in, d_out, w: int, h: int, stream: int | None = None, channels: int = 1) -> None:
    """flip_block<T,C>(in, out, w, h): in is h x w pixels of C elements, out is w x h (Source.cpp:540,562 call
    it as <float,1>)."""
    check(lib.blur_b200_flip_block_pixels(_ptr(d_in), _ptr(d_out), w, h, channels * _elem_bytes(d_in),
                                          _stream() if stream is None else stream))


# ---- DFT_image ----------------------------------------------------------------------------------

def dft_image(image: np.ndarray, nsmooth: float = 0.1, want_f32: bool = False):
    """pocketfft_2D built with `#define DFT_image` (Source.cpp:240-252, 268-276): returns the
    uint8 spectrum image 20*log10(|Re F| + 1e-5) (fft-shifted, cropped to the input size), and
    with want_f32 also the float spectrum of the padded planes [3, H, W]."""
    img = _as_image(np.ascontiguousarray(image).copy())
    rows, cols = img.shape[:2]
    padded = (C.c_int * 2)()
    border = (C.c_int * 4)()
    check(lib.blur_b200_dft_image_sizes(rows, cols, float(nsmooth), padded, border))
    out = np.empty_like(img)
    spec = np.empty((3, padded[0], padded[1]), np.float32) if want_f32 else None
    check(lib.blur_b200_dft_image_host(img.ctypes.data, out.ctypes.data, rows, cols, float(nsmooth),
                                       spec.ctypes.data if want_f32 else None, padded))
    return (out, spec) if want_f32 else out


# ---- dispatcher ---------------------------------------------------------------------------------

def Test(image: np.ndarray, flag: int = 0, nsmooth: float = 5) -> None:
    """Test(cv::Mat& image, int flag, double nsmooth), Source.cpp:574-608: 5 pocketfft_1D,
    4 fastboxblur (ksize = nsmooth*nsmooth, 2 passes), 3 pffft_, 2 pocketfft_2D.  Any other flag is the
    reference's OpenCV comparison line (`cv::GaussianBlur(image, image, Size(0, 0), nsmooth)`, :596-603):
    host OpenCV through `cv2` when it is importable, an error otherwise -- it is a comparison baseline,
    not part of the accelerated path."""
    if flag == 5:
        pocketfft_1D(image, nsmooth)
    elif flag == 4:
        img = _as_image(image)
        fastboxblur(img, img.shape[1], img.shape[0], 3, int(nsmooth * nsmooth), 2)
    elif flag == 3:
        pffft_(image, nsmooth)
    elif flag == 2:
        pocketfft_2D(image, nsmooth)
    else:
        try:
            import cv2
        except ImportError:
            raise BlurB200Error(_lib.EINVAL, "flag 1 is the reference's OpenCV baseline (cv::GaussianBlur); cv2 is not importable here") from None
        cv2.GaussianBlur(image, (0, 0), nsmooth, dst=image)


def kernel_launches() -> int:
    return int(lib.blur_b200_kernel_launches())
