// boxblur.cu — FastBoxBlur, one fused kernel per 2-D pass (sm_100a).
//
// fastboxblur(in, w, h, channels, ksize, passes), call site Source.cpp:587; declared semantics in
// SURVEY.md A.7 (upstream source absent): r = (ksize-1)/2, k = 2r+1, reflect-101 borders, a
// horizontal sliding-window mean rounded to uint8, then a vertical one on that uint8 result,
// rounding (2*sum + k) / (2k) == (sum + r) / k for odd k; `passes` repeats.
//
// boxblur_tile_kernel: one CTA owns a tile of TY rows x 384 bytes of the interleaved image.
//   load    rows y0-r .. y0+TY-1+r (reflect-101), bytes x0-(r+1)C .. x0+384+rC-1 (reflect-101 per pixel)
//           into shared memory with 32-bit loads (byte path on image borders)
//   h pass  one thread per tile row walks the row with C running sums (sliding accumulators); new and
//           old samples arrive as 32-bit words (funnel-shifted to their byte phase), four results
//           leave as one 32-bit word written IN PLACE behind the read position
//   v pass  one thread per 32-bit column walks down the rows with four running sums and stores the
//           rounded word straight to global memory (a warp stores 128 contiguous bytes per row)
// The horizontal pass, both of the reference's transposes and the intermediate image of one 2-D
// pass never touch HBM: algorithmic traffic is 3 B read + 3 B written per pixel and pass.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdlib>

namespace blurb200 {

namespace {

constexpr int kBoxTXB = 384;        // tile width in bytes: a multiple of 4*C for C = 1..4
constexpr int kBoxThreads = 128;

__device__ __forceinline__ int box_refl(int i, int L)
{
    if (i < 0) i = -i;
    if (i >= L) i = 2 * (L - 1) - i;
    return i < 0 ? 0 : (i >= L ? L - 1 : i);          // far outside: clamped, such samples are never used
}

// acc + byte kk of `add` - byte kk of `sub`: two dp4a (unsigned bytes x signed selector bytes)
__device__ __forceinline__ uint32_t box_acc(uint32_t acc, uint32_t add, uint32_t sub, int kk)
{
    int r;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(add), "r"(1 << (8 * kk)), "r"(static_cast<int>(acc)));
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(sub), "r"(static_cast<int>(0xFFu << (8 * kk))), "r"(r));
    return static_cast<uint32_t>(r);
}

struct BoxTileArgs {
    const uint8_t* in;
    uint8_t* out;
    int w, h, r;
    uint32_t magic;          // ceil(2^32 / k)
    int TY, rows_t, pitch;   // tile rows, staged rows (TY + 2r), shared-memory row pitch in bytes
    int O;                   // storage offset: input byte u (relative to the tile's first byte) lives at u + O
    int tiles_x;
    size_t img_stride;
};

template <int C>
__global__ void __launch_bounds__(kBoxThreads)
boxblur_tile_kernel(const __grid_constant__ BoxTileArgs a)
{
    extern __shared__ __align__(16) uint8_t sm[];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int tx = blockIdx.x % a.tiles_x, ty = blockIdx.x / a.tiles_x;
    const int x0b = tx * kBoxTXB, y0 = ty * a.TY;
    const int wb = a.w * C, rC = a.r * C;
    const uint8_t* img = a.in + blockIdx.y * a.img_stride;
    uint8_t* dst = a.out + blockIdx.y * a.img_stride;

    // ---- load: staged byte u in [-(r+1)C, TXB + rC) of tile row ry at sm[ry*pitch + u + O]
    const int ulo = -(rC + C), uhi = kBoxTXB + rC;
    for (int ry = wid; ry < a.rows_t; ry += kBoxThreads / 32) {
        const uint8_t* row = img + static_cast<size_t>(box_refl(y0 - a.r + ry, a.h)) * wb;
        uint8_t* srow = sm + ry * a.pitch;
        // u = 0 (mod 4) is word aligned on both sides (x0b, O, wb and the image base are multiples of 4).
        // Words that lie inside the image row travel as asynchronous 4-byte copies (every row of the tile
        // in flight at once, no register staging); only the reflected bytes of border tiles go one by one.
        const int w0 = -((-ulo + 3) & ~3);                     // ulo (negative) floored to a word
        const int x0p = x0b / C;
        for (int u = w0 + 4 * lane; u < uhi; u += 128) {
            const int gx = x0b + u;
            if (gx >= 0 && gx + 4 <= wb) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(srow + u + a.O))),
                             "l"(row + gx) : "memory");
            } else {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const int uu = u + kk;
                    if (uu < ulo || uu >= uhi) continue;
                    const int j = (uu + (rC + C)) / C - (a.r + 1), c = uu - j * C;      // pixel offset (floor), channel
                    srow[uu + a.O] = row[box_refl(x0p + j, a.w) * C + c];
                }
            }
        }
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- horizontal pass, in place: H[b] (b in [0, TXB)) overwrites storage byte b of the row
    const int sh_n = (rC + a.O) & 3, sh_o = (a.O - C - rC) & 3;
    for (int ry = tid; ry < a.rows_t; ry += kBoxThreads) {
        uint8_t* srow = sm + ry * a.pitch;
        uint32_t acc[C];
#pragma unroll
        for (int c = 0; c < C; ++c) {
            // window of the (virtual) outputs -C + c: samples u = -C + c + i*C, i = -r..r ; + r = rounding bias
            uint32_t s = a.r;
            for (int i = -a.r; i <= a.r; ++i) s += srow[-C + c + i * C + a.O];
            acc[c] = s;
        }
        const uint32_t* nwp = reinterpret_cast<const uint32_t*>(srow + ((rC + a.O) & ~3));
        const uint32_t* owp = reinterpret_cast<const uint32_t*>(srow + ((a.O - C - rC) & ~3));
        uint32_t* hw = reinterpret_cast<uint32_t*>(srow);
        uint32_t n_lo = nwp[0], o_lo = owp[0];
#pragma unroll 1
        for (int wq = 0; wq < kBoxTXB / 4; wq += C) {
#pragma unroll
            for (int it = 0; it < C; ++it) {
                const uint32_t n_hi = nwp[wq + it + 1], o_hi = owp[wq + it + 1];
                const uint32_t nw = __funnelshift_r(n_lo, n_hi, 8 * sh_n);
                const uint32_t ow = __funnelshift_r(o_lo, o_hi, 8 * sh_o);
                n_lo = n_hi; o_lo = o_hi;
                uint32_t outw = 0;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const int ph = (4 * it + kk) % C;
                    acc[ph] = box_acc(acc[ph], nw, ow, kk);
                    outw |= __umulhi(acc[ph], a.magic) << (8 * kk);
                }
                hw[wq + it] = outw;
            }
        }
    }
    __syncthreads();

    // ---- vertical pass: thread = one 32-bit column of H, sliding down the staged rows
    for (int wc = tid; wc < kBoxTXB / 4; wc += kBoxThreads) {
        if (x0b + 4 * wc >= wb) break;
        const uint8_t* col = sm + 4 * wc;
        uint32_t acc[4] = {static_cast<uint32_t>(a.r), static_cast<uint32_t>(a.r), static_cast<uint32_t>(a.r), static_cast<uint32_t>(a.r)};
        for (int ry = 0; ry <= 2 * a.r; ++ry) {
            const uint32_t v = *reinterpret_cast<const uint32_t*>(col + ry * a.pitch);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) acc[kk] += (v >> (8 * kk)) & 255u;
        }
        const int ny = min(a.TY, a.h - y0);
        uint8_t* o = dst + static_cast<size_t>(y0) * wb + x0b + 4 * wc;
        for (int j = 0; j < ny; ++j) {
            uint32_t outw = 0;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) outw |= __umulhi(acc[kk], a.magic) << (8 * kk);
            *reinterpret_cast<uint32_t*>(o + static_cast<size_t>(j) * wb) = outw;
            if (j + 1 < ny) {
                const uint32_t vn = *reinterpret_cast<const uint32_t*>(col + (j + 1 + 2 * a.r) * a.pitch);
                const uint32_t vo = *reinterpret_cast<const uint32_t*>(col + j * a.pitch);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) acc[kk] = box_acc(acc[kk], vn, vo, kk);
            }
        }
    }
}

template <int C>
cudaError_t launch_tile(const BoxTileArgs& a, int tiles, int batch, size_t smem, cudaStream_t st)
{
    // per-device function attribute: set on every launch (cheap), never cached process-wide
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(boxblur_tile_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    boxblur_tile_kernel<C><<<dim3(tiles, batch), kBoxThreads, smem, st>>>(a);
    return cudaGetLastError();
}

}  // namespace

// One 2-D pass (horizontal then vertical) of the box blur over `batch` images, in -> out (distinct
// buffers).  Returns cudaErrorNotSupported when the fused kernel does not cover the case (the caller
// then runs the two-kernel path): channels > 4, k == 1 or k > 4095, a byte pitch that is not a
// multiple of 4, or a window too wide for the shared-memory tile.
cudaError_t launch_box_fused_pass(const uint8_t* in, uint8_t* out, int w, int h, int C, int r, int batch, size_t img_stride,
                                  cudaStream_t st)
{
    const int k = 2 * r + 1;
    if (C < 1 || C > 4 || k < 3 || k > 4095 || batch > 65535) return cudaErrorNotSupported;
    if ((w * C) % 4 || img_stride % 4 || reinterpret_cast<uintptr_t>(in) % 4 || reinterpret_cast<uintptr_t>(out) % 4)
        return cudaErrorNotSupported;
    BoxTileArgs a{};
    a.in = in; a.out = out; a.w = w; a.h = h; a.r = r;
    a.magic = static_cast<uint32_t>(((1ull << 32) + k - 1) / k);
    const int rC = r * C;
    a.O = (rC + C + 4 + 3) & ~3;
    int pitch = (a.O + kBoxTXB + rC + 8 + 3) & ~3;
    if ((pitch / 4) % 2 == 0) pitch += 4;                 // odd word pitch: 32 rows hit 32 different banks
    a.pitch = pitch;
    a.TY = 0;
    static const int ty_env = std::getenv("BLUR_B200_BOX_TY") ? std::atoi(std::getenv("BLUR_B200_BOX_TY")) : 0;   // experiment knob
    // The kernel is occupancy-bound (r = 15, 256 x 1080p: 96-row tiles / 3 resident CTAs 6.45 ms, 64 / 4 CTAs
    // 6.03 ms, 62 / 5 CTAs 5.53 ms): take the most resident CTAs whose tile still has at least as many
    // output rows as halo rows, then even the tiles out over the image height.
    if (ty_env && static_cast<size_t>(ty_env + 2 * r) * pitch <= 110u * 1024u) a.TY = ty_env;
    for (int n = 5; n >= 2 && !a.TY && !ty_env; --n) {
        const int rows_fit = static_cast<int>((227u * 1024u / n - 1024u) / pitch);
        const int ty = std::min(rows_fit - 2 * r, h);
        if (ty >= 16 && ty >= 2 * r) {
            const int tiles = (h + ty - 1) / ty;
            a.TY = (h + tiles - 1) / tiles;
        }
    }
    for (int ty : {96, 64, 32, 16}) {
        if (a.TY) break;
        if (ty_env && ty != ty_env) continue;
        const size_t need = static_cast<size_t>(ty + 2 * r) * pitch;
        if (need <= (ty == 96 ? 74u : 110u) * 1024u) { a.TY = ty; break; }
    }
    if (!a.TY) return cudaErrorNotSupported;
    a.rows_t = a.TY + 2 * r;
    a.tiles_x = (w * C + kBoxTXB - 1) / kBoxTXB;
    a.img_stride = img_stride;
    const int tiles = a.tiles_x * ((h + a.TY - 1) / a.TY);
    const size_t smem = static_cast<size_t>(a.rows_t) * pitch;
    switch (C) {
    case 1: return launch_tile<1>(a, tiles, batch, smem, st);
    case 2: return launch_tile<2>(a, tiles, batch, smem, st);
    case 3: return launch_tile<3>(a, tiles, batch, smem, st);
    default: return launch_tile<4>(a, tiles, batch, smem, st);
    }
}

}  // namespace blurb200
