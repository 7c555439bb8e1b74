// box_noalloc.cu — blur_b200_fastboxblur_device_noalloc: the device-resident FastBoxBlur for callers that
// must not allocate.
//
// Who this is for.  blur_b200_fastboxblur_device needs a second buffer as large as the image batch (the two
// passes ping-pong between the images and that buffer), and the caller has to find it somewhere.  Some
// callers cannot: a filter node running inside a host application that owns the GPU memory pool and hands the
// node only its input and output surfaces; a real-time loop in which a cudaMalloc (which can synchronise the
// whole device and takes an unbounded time) is forbidden once it has started; a process that has been told
// its memory budget is spent.  For them the library carries its own second buffer: a fixed array in the
// device-side data segment of the library image, BLUR_B200_NOALLOC_SCRATCH_BYTES large.  It comes into being
// when the CUDA runtime loads this library's code onto a GPU and stays until the process exits, on every GPU
// that does so; nothing is allocated or freed by the call itself — no cudaMalloc / cudaFree, no stream-ordered
// allocation, no library routine that allocates.
//
// What to know before using it.
//   * Limit: w*h*channels*batch <= BLUR_B200_NOALLOC_SCRATCH_BYTES (32 MiB: one 3840x2160 RGB image, or five
//     1920x1080 RGB images).  Anything larger is refused with BLUR_B200_ESCRATCH before any CUDA work.
//   * The first call on each GPU looks the array's device address up (cudaGetSymbolAddress, with that GPU
//     current) and remembers it; it may also be the call that makes the runtime load the library's kernels.
//     Make that first call during set-up, and before beginning any stream capture.
//   * There is one array per GPU and every call on that GPU uses it.  Two calls whose kernels can be in
//     flight at the same time (different streams, or different host threads on different streams) would
//     scribble over each other's intermediate image: keep calls on one GPU ordered one after the other.
//   * The array keeps whatever the previous call left in it.  That is harmless here: each pass writes every
//     byte of its destination before anything reads it, so nothing has to start at zero and nothing is cleared.
//
// The arithmetic is that of blur_b200_fastboxblur_device (same kernels, same launch path): results are
// identical byte for byte.
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <mutex>

#include "../../include/blur_b200.h"

namespace blurb200 {
__attribute__((visibility("hidden"))) void set_cuda_error(const char* what, cudaError_t e);
}

// The second buffer.  Host code may only use this name as a handle for cudaGetSymbolAddress: as a pointer it
// is a host-side shadow, not where the GPU keeps the bytes.
static __device__ __align__(256) unsigned char g_noalloc_scratch[BLUR_B200_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint8_t* g_scratch_addr[kMaxDevices] = {};   // device address of g_noalloc_scratch on GPU i, once resolved

// Address of the array on the GPU that is current for the calling thread; resolved once per GPU.
int scratch_on_current_device(uint8_t** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { blurb200::set_cuda_error("cudaGetDevice", e); return BLUR_B200_ECUDA; }
    if (dev < 0 || dev >= kMaxDevices) return BLUR_B200_EINVAL;
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            blurb200::set_cuda_error("cudaGetSymbolAddress(noalloc scratch)", e);
            return BLUR_B200_ECUDA;
        }
        g_scratch_addr[dev] = static_cast<uint8_t*>(p);
    }
    *out = g_scratch_addr[dev];
    return BLUR_B200_OK;
}

}  // namespace

extern "C" int blur_b200_fastboxblur_device_noalloc(uint8_t* d_img, int w, int h, int channels, int ksize, int passes,
                                                    int batch, void* stream)
{
    if (!d_img || w < 1 || h < 1 || channels < 1 || channels > 4 || ksize < 1 || passes < 0 || batch < 1)
        return BLUR_B200_EINVAL;
    const unsigned long long total = 1ull * static_cast<unsigned>(w) * static_cast<unsigned>(h) *
                                     static_cast<unsigned>(channels);
    if (total > BLUR_B200_NOALLOC_SCRATCH_BYTES || total * static_cast<unsigned>(batch) > BLUR_B200_NOALLOC_SCRATCH_BYTES)
        return BLUR_B200_ESCRATCH;
    uint8_t* scratch = nullptr;
    const int rc = scratch_on_current_device(&scratch);
    if (rc) return rc;
    return blur_b200_fastboxblur_device(d_img, w, h, channels, ksize, passes, batch, scratch, stream);
}
