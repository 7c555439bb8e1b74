"""Loader for the in-tree CUDA library (blur_algorithms_b200/libblur_b200.so).

There is no fallback: if the shared library is missing or does not load, importing the package
fails loudly.  Build it with `python -c "import __graft_entry__ as g; g.build()"` or
`make -C blur_algorithms_b200/csrc`.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BLUR_B200_LIB") or os.path.join(_HERE, "libblur_b200.so")   # env: experiment builds only

OK, EINVAL, EPAD, ESIZE, ECUDA, ENOMEM, ESCRATCH = 0, 1, 2, 3, 4, 5, 6
NOALLOC_SCRATCH_BYTES = 32 << 20   # BLUR_B200_NOALLOC_SCRATCH_BYTES of include/blur_b200.h
NYQUIST_EXACT, NYQUIST_PFFFT = 0, 1


class BlurB200Error(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__(what)
        self.code = code


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA extension is not built (run __graft_entry__.build()). "
            "blur_algorithms_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, u32, f64 = C.c_void_p, C.c_int, C.c_uint32, C.c_double
    pi32 = C.POINTER(C.c_int)
    sig = {
        "blur_b200_strerror": (C.c_char_p, [i32]),
        "blur_b200_last_cuda_error": (C.c_char_p, []),
        "blur_b200_is_valid_size": (i32, [i32]),
        "blur_b200_nearest_transform_size": (i32, [i32]),
        "blur_b200_gaussian_window": (i32, [f64, i32]),
        "blur_b200_get_gaussian": (i32, [vp, f64, i32, i32]),
        "blur_b200_gauss_geometry": (i32, [i32, i32, f64, pi32]),
        "blur_b200_gauss_shard_geometry": (i32, [i32, i32, f64, i32, i32, i32, C.POINTER(C.c_int64)]),
        "blur_b200_gauss_plan_create": (i32, [C.POINTER(vp), i32, i32, f64, i32, i32]),
        "blur_b200_gauss_plan_create_sharded": (i32, [C.POINTER(vp), i32, i32, f64, i32, i32, i32, i32]),
        "blur_b200_gauss_plan_destroy": (None, [vp]),
        "blur_b200_gauss_plan_info": (i32, [vp, C.POINTER(C.c_int64)]),
        "blur_b200_gauss_plan_uses_direct": (i32, [vp]),
        "blur_b200_gauss_plan_set_direct": (i32, [vp, i32]),
        "blur_b200_gauss_run_device": (i32, [vp, vp, vp, vp]),
        "blur_b200_gauss_run_host": (i32, [vp, vp, vp]),
        "blur_b200_gauss_run_host_batch": (i32, [vp, C.POINTER(vp), C.POINTER(vp), i32]),
        "blur_b200_gauss_u8c3": (i32, [vp, i32, i32, f64, i32]),
        "blur_b200_release_cached": (None, []),
        "blur_b200_gauss_row_pass": (i32, [vp, vp, vp, vp]),
        "blur_b200_gauss_col_pass": (i32, [vp, vp, vp, vp]),
        "blur_b200_gauss_row_pass_peer": (i32, [vp, vp, C.POINTER(vp), i32, vp]),
        "blur_b200_gauss_row_pass_push": (i32, [vp, vp, C.POINTER(vp), i32, i32, vp]),
        "blur_b200_boxfft_klen": (i32, [i32, i32, f64]),
        "blur_b200_boxfft_plan_create": (i32, [C.POINTER(vp), i32, i32, f64, i32, i32]),
        "blur_b200_boxfft_u8c3": (i32, [vp, i32, i32, f64, i32]),
        "blur_b200_gauss_halo_info": (i32, [vp, i32, i32, C.POINTER(C.c_int64)]),
        "blur_b200_gauss_halo_row_pass": (i32, [vp, i32, i32, vp, vp, vp]),
        "blur_b200_gauss_halo_push": (i32, [vp, i32, i32, vp, vp, vp, vp]),
        "blur_b200_gauss_halo_col_pass": (i32, [vp, i32, i32, vp, vp, vp]),
        "blur_b200_gauss_quirk_buffers": (i32, [vp, C.POINTER(vp), C.POINTER(C.c_int64), C.POINTER(vp), C.POINTER(C.c_int64)]),
        "blur_b200_gauss_quirk_local": (i32, [vp, vp, vp]),
        "blur_b200_gauss_quirk_finish": (i32, [vp, vp]),
        "blur_b200_gauss_staged_supported": (i32, [vp, i32]),
        "blur_b200_gauss_row_pass_staged": (i32, [vp, vp, C.POINTER(vp), i32, i32, i32, i32, vp]),
        "blur_b200_gauss_stage_wait": (i32, [vp, i32, vp]),
        "blur_b200_gauss_col_pass_stage": (i32, [vp, vp, vp, i32, i32, i32, vp]),
        "blur_b200_gauss_rowpass_f32": (i32, [vp, vp, vp, vp]),
        "blur_b200_fastboxblur_device": (i32, [vp, i32, i32, i32, i32, i32, i32, vp, vp]),
        "blur_b200_fastboxblur_device_noalloc": (i32, [vp, i32, i32, i32, i32, i32, i32, vp]),
        "blur_b200_fastboxblur": (i32, [vp, i32, i32, i32, i32, i32]),
        "blur_b200_fastboxblur_batch_host": (i32, [vp, i32, i32, i32, i32, i32, i32]),
        "blur_b200_reflect_101_u8": (i32, [vp, vp, i32, i32, i32, i32, pi32, i32, vp]),
        "blur_b200_reflect_101_u8_host": (i32, [vp, vp, i32, i32, i32, i32, pi32, i32]),
        "blur_b200_deinterleave_u8_f32": (i32, [vp, vp, vp, vp, u32, vp]),
        "blur_b200_interleave_f32_u8": (i32, [vp, vp, vp, vp, u32, vp]),
        "blur_b200_flip_block_f32": (i32, [vp, vp, i32, i32, vp]),
        "blur_b200_deinterleave_u8_f32_host": (i32, [vp, vp, vp, vp, u32]),
        "blur_b200_interleave_f32_u8_host": (i32, [vp, vp, vp, vp, u32]),
        "blur_b200_flip_block_f32_host": (i32, [vp, vp, i32, i32]),
        "blur_b200_deinterleave_typed": (i32, [vp, vp, vp, vp, u32, i32, i32, vp]),
        "blur_b200_interleave_typed": (i32, [vp, vp, vp, vp, u32, i32, i32, vp]),
        "blur_b200_flip_block_pixels": (i32, [vp, vp, i32, i32, i32, vp]),
        "blur_b200_deinterleave_typed_host": (i32, [vp, vp, vp, vp, u32, i32, i32]),
        "blur_b200_interleave_typed_host": (i32, [vp, vp, vp, vp, u32, i32, i32]),
        "blur_b200_flip_block_pixels_host": (i32, [vp, vp, i32, i32, i32]),
        "blur_b200_dft_image_device": (i32, [vp, vp, i32, i32, f64, vp, pi32, vp]),
        "blur_b200_dft_image_host": (i32, [vp, vp, i32, i32, f64, vp, pi32]),
        "blur_b200_dft_image_sizes": (i32, [i32, i32, f64, pi32, pi32]),
        "blur_b200_kernel_launches": (C.c_uint64, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)   # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()
EXPORTED = [
    "blur_b200_strerror", "blur_b200_last_cuda_error", "blur_b200_is_valid_size",
    "blur_b200_nearest_transform_size", "blur_b200_gaussian_window", "blur_b200_get_gaussian",
    "blur_b200_gauss_geometry", "blur_b200_gauss_shard_geometry", "blur_b200_gauss_plan_create", "blur_b200_gauss_plan_create_sharded",
    "blur_b200_gauss_plan_destroy", "blur_b200_gauss_plan_info", "blur_b200_gauss_plan_uses_direct", "blur_b200_gauss_plan_set_direct", "blur_b200_gauss_run_device",
    "blur_b200_gauss_run_host", "blur_b200_gauss_run_host_batch", "blur_b200_gauss_u8c3", "blur_b200_release_cached", "blur_b200_gauss_row_pass",
    "blur_b200_gauss_col_pass", "blur_b200_gauss_row_pass_peer", "blur_b200_gauss_row_pass_push", "blur_b200_boxfft_klen", "blur_b200_boxfft_plan_create", "blur_b200_boxfft_u8c3",
    "blur_b200_gauss_halo_info", "blur_b200_gauss_halo_row_pass", "blur_b200_gauss_halo_push", "blur_b200_gauss_halo_col_pass",
    "blur_b200_gauss_quirk_buffers", "blur_b200_gauss_quirk_local", "blur_b200_gauss_quirk_finish",
    "blur_b200_gauss_staged_supported",
    "blur_b200_gauss_row_pass_staged", "blur_b200_gauss_stage_wait", "blur_b200_gauss_col_pass_stage", "blur_b200_gauss_rowpass_f32", "blur_b200_fastboxblur_device",
    "blur_b200_fastboxblur_device_noalloc", "blur_b200_fastboxblur", "blur_b200_fastboxblur_batch_host", "blur_b200_reflect_101_u8", "blur_b200_reflect_101_u8_host",
    "blur_b200_deinterleave_u8_f32", "blur_b200_interleave_f32_u8", "blur_b200_flip_block_f32",
    "blur_b200_deinterleave_u8_f32_host", "blur_b200_interleave_f32_u8_host", "blur_b200_flip_block_f32_host",
    "blur_b200_deinterleave_typed", "blur_b200_interleave_typed", "blur_b200_flip_block_pixels",
    "blur_b200_deinterleave_typed_host", "blur_b200_interleave_typed_host", "blur_b200_flip_block_pixels_host",
    "blur_b200_dft_image_device", "blur_b200_dft_image_host", "blur_b200_dft_image_sizes",
    "blur_b200_kernel_launches",
]


def check(code: int) -> None:
    if code == OK:
        return
    msg = lib.blur_b200_strerror(code).decode()
    if code == ECUDA:
        msg += ": " + lib.blur_b200_last_cuda_error().decode()
    raise BlurB200Error(code, msg)
