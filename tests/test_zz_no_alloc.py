"""blur_b200_fastboxblur_device_noalloc: the FastBoxBlur entry for callers that are not allowed to allocate
GPU memory — a filter node that lives inside another application's memory pool and is handed nothing but its
input and output surfaces, or a real-time loop where cudaMalloc (device-wide synchronisation, unbounded
latency) is banned once the loop runs.  The ordinary entry makes the caller bring a scratch buffer as large
as the image batch; this one uses a fixed 32 MiB array built into the library instead.

What is checked, through the C ABI (ctypes -> libblur_b200.so), FastBoxBlur's tolerance (bit-exact):
  * two calls in a row, different images and different geometry, each equal to blur_b200_fastboxblur_device
    byte for byte — the second call finds the first call's leftovers in the shared array and must not care;
  * a repeat call does not lower the GPU's free memory (nothing was allocated behind the caller's back);
  * a batch beyond the built-in array is refused with BLUR_B200_ESCRATCH and a message that names the limit.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a problem here should not
hide the parity tests.  It does not depend on anything the other tests did; it passes alone as well.
"""
import numpy as np
import pytest

from util import synth


@pytest.fixture(scope="module")
def bb():
    import blur_algorithms_b200 as m
    return m


def _both(bb, torch, imgs, ksize, passes):
    """Blur `imgs` (uint8 [batch][h][w][c]) with the ordinary entry (caller's scratch) and with the
    no-allocation entry; return both results as numpy arrays."""
    batch, h, w, c = imgs.shape
    want = torch.from_numpy(imgs).cuda()
    bb.fastboxblur_device(want, w, h, c, ksize, passes, batch, torch.empty_like(want))
    got = torch.from_numpy(imgs).cuda()
    bb.fastboxblur_device_noalloc(got, w, h, c, ksize, passes, batch)
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


@pytest.mark.gpu
def test_noalloc_box_equals_scratch_entry_twice_in_a_row(bb):
    import torch
    assert torch.cuda.is_available()
    # call 1: BASELINE config 5's shape (1080p RGB, ksize 31, 2 passes), 4 images = 24.9 MB of the 32 MiB
    a = np.stack([synth("noise", 1080, 1920, s) for s in range(4)])
    got, want = _both(bb, torch, a, 31, 2)
    assert np.array_equal(got, want)
    # call 2 right behind it: other content, other size, odd pass count (the result ends up in the shared
    # array and is copied back), single channel path of the kernels
    b = np.stack([synth("noise", 300, 400, 10 + s)[:, :, :1] for s in range(3)])
    b = np.ascontiguousarray(b)
    got, want = _both(bb, torch, b, 9, 3)
    assert np.array_equal(got, want)
    assert not np.array_equal(got, b)          # it did blur
    # call 3 = call 1 again, now that the array holds call 2's bytes; and no memory may have been taken
    d = torch.from_numpy(a).cuda()
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    bb.fastboxblur_device_noalloc(d, 1920, 1080, 3, 31, 2, 4)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after >= free_before, (free_before, free_after)
    ref = torch.from_numpy(a).cuda()
    bb.fastboxblur_device(ref, 1920, 1080, 3, 31, 2, 4, torch.empty_like(ref))
    assert torch.equal(d, ref)


def test_noalloc_box_refuses_more_than_its_array(bb):
    """The size check comes before any CUDA call, so this runs without a GPU: 6 x 1080p RGB = 37.3 MB."""
    from blur_algorithms_b200 import _lib
    assert _lib.NOALLOC_SCRATCH_BYTES == 32 << 20
    fake = 0x1000   # never dereferenced: the call must fail on the size alone
    rc = _lib.lib.blur_b200_fastboxblur_device_noalloc(fake, 1920, 1080, 3, 31, 2, 6, None)
    assert rc == _lib.ESCRATCH
    assert b"32 MiB" in _lib.lib.blur_b200_strerror(rc)
    with pytest.raises(_lib.BlurB200Error) as ei:
        _lib.check(rc)
    assert ei.value.code == _lib.ESCRATCH
    # exactly at the limit is not a size error (a null image is, which is what we pass)
    assert _lib.lib.blur_b200_fastboxblur_device_noalloc(None, 4096, 2048, 4, 31, 2, 1, None) == _lib.EINVAL
    # dimensions whose product overflows 32 bits are still caught
    assert _lib.lib.blur_b200_fastboxblur_device_noalloc(fake, 65536, 65536, 4, 3, 1, 4, None) == _lib.ESCRATCH
