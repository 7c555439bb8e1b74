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


// ------------------------------------------------------------------------------------------------
// box2_tile_kernel (C = 3 or 4, 3 <= k <= 127, 16-byte aligned rows): the same tile scheme with the
// arithmetic packed two 16-bit sums to a register and the rounding done on the FMA pipe.
//   load    one bulk copy (TMA, cp.async.bulk) per staged row, completion on an mbarrier; bytes outside
//           the image row are mirrored afterwards from the staged row itself (border tiles only)
//   h pass  thread = tile row.  The row is read in 16-byte chunks (LDS.128, conflict free for a pitch
//           of an odd number of chunks); the new-sample stream is chunk aligned up to a byte shift
//           (funnel shift), the old-sample stream additionally by a word shift (template WO).  The
//           running sums of the four bytes of an output word live in two registers: e = (a0 | a2<<16),
//           o = (a1 | a3<<16); sum[b] = sum[b-C] + new[b] - old[b] is, for C = 3,
//           e' = o + n_e - o_e ; o' = (e.hi | e'.lo<<16) + n_o - o_o   (C = 4: lane-wise).
//           The window is primed by nw add-only iterations over the k*C bytes left of the first output.
//   round   a 15-bit sum n becomes the float 32768 + n with one PRMT (exponent 0x47, n in mantissa bits
//           8..22); fma(32768 + n, 1/k, (r + 1/2)/k - 32768/k) lies within 1e-4 of (n + r + 1/2)/k, whose
//           fractional part is at least 1/(2k) away from an integer, so its floor (add.rm with 2^23, low
//           mantissa byte) is exactly (n + r) / k.
//   v pass  thread = 32-bit column of the h result, same packed sums, result words straight to global.
constexpr int kBox2Threads = 128;

struct Box2Args {
    const uint8_t* in;
    uint8_t* out;
    int w, h, r, wb;
    int TY, rows_t, pitch, TXB, tiles_x;
    int O;            // input byte u (relative to the tile's first byte) is staged at srow[u + O], O % 16 == 0
    int E;            // h result byte b is stored at srow[b + E], E in {0, 4, 8, 12}
    int nw, ni;       // priming / main iterations of the h pass (16 output bytes each)
    int cn0, co0;     // first 16-byte chunk of the new stream (priming) and of the old stream (main loop)
    int bn, bo;       // byte shifts of the two streams
    int wo;           // word shift of the old stream (selects the template instance)
    int mcnt;         // leading bytes of the first priming chunk that lie left of the window
    float inv, cb;    // rounding constants, see above
    size_t img_stride;
};

// The ALU pipe (PRMT / LOP3 / IADD3 / SHF, 16 lanes per scheduler) is this kernel's bound, so everything
// that can be phrased as a multiply-add is: even(v) = v - 256 * odd(v), the byte packing of the rounded
// results, and every plain add go to the FMA pipe as IMAD.
__device__ __forceinline__ uint32_t box2_odd(uint32_t v) { return __byte_perm(v, 0u, 0x4341); }
__device__ __forceinline__ uint32_t box2_mad(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ uint32_t box2_round4(uint32_t e, uint32_t o, float inv, float cb)
{
    constexpr uint32_t K = 0x47000000u;
    const uint32_t f0 = __byte_perm(e, K, 0x7104), f2 = __byte_perm(e, K, 0x7324);
    const uint32_t f1 = __byte_perm(o, K, 0x7104), f3 = __byte_perm(o, K, 0x7324);
    uint32_t u0, u1, u2, u3;
    asm("{\n.reg .b64 a, b, m, c, t;\n"
        "mov.b64 m, {%6, %6};\nmov.b64 c, {%7, %7};\n"
        "mov.b64 a, {%4, %5};\nfma.rn.f32x2 a, a, m, c;\n"
        "mov.b64 b, {%8, %9};\nfma.rn.f32x2 b, b, m, c;\n"
        "mov.b64 t, {%10, %10};\n"
        "add.rm.f32x2 a, a, t;\nadd.rm.f32x2 b, b, t;\n"
        "mov.b64 {%0, %1}, a;\nmov.b64 {%2, %3}, b;\n}"
        : "=r"(u0), "=r"(u1), "=r"(u2), "=r"(u3)
        : "r"(f0), "r"(f1), "f"(inv), "f"(cb), "r"(f2), "r"(f3), "f"(8388608.f));
    // u_j = 0x4B000000 + q_j: q0 + 256 q1 + 65536 q2 + 2^24 q3 = u0 + 256 u1 + 65536 (u2 + 256 u3) - 0x4B000000 (mod 2^32)
    return box2_mad(box2_mad(u3, 256u, u2), 65536u, box2_mad(u1, 256u, u0 - 0x4B000000u));
}

// one output word: slide the four sums by one word of new (nw) and old (ow) samples.
// D = nw - ow as plain 32-bit words is sum_j (n_j - o_j) 256^j, so the even-byte differences are D - 256 * (odd differences).
template <int C>
__device__ __forceinline__ void box2_slide(uint32_t& e, uint32_t& o, uint32_t nw, uint32_t ow)
{
    const uint32_t dod = box2_odd(nw) - box2_odd(ow);
    const uint32_t D = nw - ow;
    if (C == 3) {
        const uint32_t e2 = box2_mad(dod, 0xFFFFFF00u, o + D);
        o = __byte_perm(e, e2, 0x5432) + dod;
        e = e2;
    } else {
        e = box2_mad(dod, 0xFFFFFF00u, e + D);
        o += dod;
    }
}
template <int C>
__device__ __forceinline__ void box2_prime(uint32_t& e, uint32_t& o, uint32_t nw)
{
    const uint32_t od = box2_odd(nw);
    if (C == 3) {
        const uint32_t e2 = box2_mad(od, 0xFFFFFF00u, o + nw);
        o = __byte_perm(e, e2, 0x5432) + od;
        e = e2;
    } else {
        e = box2_mad(od, 0xFFFFFF00u, e + nw);
        o += od;
    }
}

template <int C, int WO>
__device__ __forceinline__ void box2_hpass(uint8_t* srow, const Box2Args& a)
{
    uint4* row4 = reinterpret_cast<uint4*>(srow);
    const uint4* np = row4 + a.cn0;
    const int shn = 8 * a.bn, sho = 8 * a.bo;
    uint32_t e = 0, o = 0;
    uint4 nc = np[0];
    // leading bytes of the first chunk that are not part of any window
    uint32_t m[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int z = a.mcnt - 4 * j;                       // bytes of word j to clear
        m[j] = z <= 0 ? 0xFFFFFFFFu : (z >= 4 ? 0u : 0xFFFFFFFFu << (8 * z));
    }
#pragma unroll 1
    for (int i = 0; i < a.nw; ++i) {
        const uint4 nn = np[i + 1];
        const uint32_t w0 = __funnelshift_r(nc.x, nc.y, shn) & m[0], w1 = __funnelshift_r(nc.y, nc.z, shn) & m[1];
        const uint32_t w2 = __funnelshift_r(nc.z, nc.w, shn) & m[2], w3 = __funnelshift_r(nc.w, nn.x, shn) & m[3];
        m[0] = m[1] = m[2] = m[3] = 0xFFFFFFFFu;
        box2_prime<C>(e, o, w0); box2_prime<C>(e, o, w1); box2_prime<C>(e, o, w2); box2_prime<C>(e, o, w3);
        nc = nn;
    }
    np += a.nw;
    const uint4* op = row4 + a.co0;
    uint4 oc = op[0];
#pragma unroll 1
    for (int i = 0; i < a.ni; ++i) {
        const uint4 nn = np[i + 1], on = op[i + 1];
        const uint32_t nwd[4] = {__funnelshift_r(nc.x, nc.y, shn), __funnelshift_r(nc.y, nc.z, shn),
                                 __funnelshift_r(nc.z, nc.w, shn), __funnelshift_r(nc.w, nn.x, shn)};
        const uint32_t ow8[8] = {oc.x, oc.y, oc.z, oc.w, on.x, on.y, on.z, on.w};
        uint4 res;
        uint32_t* rw = reinterpret_cast<uint32_t*>(&res);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t owd = __funnelshift_r(ow8[WO + j], ow8[WO + j + 1], sho);
            box2_slide<C>(e, o, nwd[j], owd);
            rw[j] = box2_round4(e, o, a.inv, a.cb);
        }
        row4[i] = res;
        nc = nn; oc = on;
    }
}

__device__ __forceinline__ void box2_mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@!p bra W;\n}" ::"r"(bar), "r"(parity) : "memory");
}

template <int C>
__global__ void __launch_bounds__(kBox2Threads, 3)
box2_tile_kernel(const __grid_constant__ Box2Args a)
{
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar_mem;
    const int tid = threadIdx.x;
    const int tx = blockIdx.x % a.tiles_x, ty = blockIdx.x / a.tiles_x;
    const int x0b = tx * a.TXB, y0 = ty * a.TY;
    const int wb = a.wb, rC = a.r * C;
    const uint8_t* img = a.in + blockIdx.y * a.img_stride;
    uint8_t* dst = a.out + blockIdx.y * a.img_stride;
    const uint32_t bar = static_cast<uint32_t>(__cvta_generic_to_shared(&bar_mem));

    // ---- load: global bytes [g0, g1) of every staged row, one bulk copy per row
    const int g0 = max(0, x0b - a.O), g1 = min(wb, x0b + ((a.TXB + rC + 15) & ~15));
    const uint32_t nbytes = static_cast<uint32_t>(g1 - g0);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(nbytes * static_cast<uint32_t>(a.rows_t)) : "memory");
    if (tid < a.rows_t) {
        const uint8_t* src = img + static_cast<size_t>(box_refl(y0 - a.r + tid, a.h)) * wb + g0;
        const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(sm + tid * a.pitch + (g0 - (x0b - a.O))));
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(d), "l"(src), "r"(nbytes), "r"(bar) : "memory");
    }
    // one thread polls (a spinning CTA would take issue slots from the two computing next to it); the
    // barrier orders its acquire of the completed phase before everybody's reads
    if (tid == 0) box2_mbar_wait(bar, 0);
    __syncthreads();
    // mirrored bytes left / right of the image row (reflect-101 per pixel), taken from the staged row
    const int ul = -(rC + 16);
    if (x0b + ul < 0) {
        const int n = -(x0b + ul);
        for (int idx = tid; idx < a.rows_t * n; idx += kBox2Threads) {
            const int ry = idx / n, gb = x0b + ul + idx % n;        // gb < 0
            const int j = -((-gb + C - 1) / C), c = gb - j * C;
            uint8_t* srow = sm + ry * a.pitch + a.O - x0b;
            srow[gb] = srow[min(-j, a.w - 1) * C + c];
        }
    }
    if (x0b + a.TXB + rC > wb) {
        const int n = min(x0b + a.TXB, wb) + rC - wb;
        for (int idx = tid; idx < a.rows_t * n; idx += kBox2Threads) {
            const int ry = idx / n, gb = wb + idx % n;
            const int j = gb / C, c = gb - j * C;
            uint8_t* srow = sm + ry * a.pitch + a.O - x0b;
            srow[gb] = srow[max(2 * (a.w - 1) - j, 0) * C + c];
        }
    }
    if (x0b + ul < 0 || x0b + a.TXB + rC > wb) __syncthreads();

    // ---- horizontal pass, in place
    if (tid < a.rows_t) {
        uint8_t* srow = sm + tid * a.pitch;
        switch (a.wo) {
        case 0: box2_hpass<C, 0>(srow, a); break;
        case 1: box2_hpass<C, 1>(srow, a); break;
        case 2: box2_hpass<C, 2>(srow, a); break;
        default: box2_hpass<C, 3>(srow, a); break;
        }
    }
    __syncthreads();

    // ---- vertical pass
    if (4 * tid < a.TXB && x0b + 4 * tid < wb) {
        const int pw = a.pitch / 4;
        const uint32_t* col = reinterpret_cast<const uint32_t*>(sm + a.E) + tid;
        uint32_t e = 0, o = 0;
        const int k = 2 * a.r + 1;
#pragma unroll 4
        for (int ry = 0; ry < k; ++ry) box2_prime<4>(e, o, col[ry * pw]);
        const int ny = min(a.TY, a.h - y0);
        uint8_t* po = dst + static_cast<size_t>(y0) * wb + x0b + 4 * tid;
        const uint32_t* pn = col + k * pw;                     // sample entering the window of row j + 1
        const uint32_t* pc = col;                              // sample leaving it
#pragma unroll 4
        for (int j = 0; j < ny; ++j) {
            const uint32_t vn = *pn, vo = *pc;                 // staged rows 0 .. TY + 2r: always inside the tile's shared memory
            *reinterpret_cast<uint32_t*>(po) = box2_round4(e, o, a.inv, a.cb);
            box2_slide<4>(e, o, vn, vo);
            pn += pw; pc += pw; po += wb;
        }
    }
}

template <int C>
cudaError_t launch_box2(const Box2Args& a, int tiles, int batch, size_t smem, cudaStream_t st)
{
    cudaError_t e = cudaFuncSetAttribute(box2_tile_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    box2_tile_kernel<C><<<dim3(tiles, batch), kBox2Threads, smem, st>>>(a);
    return cudaGetLastError();
}

// Geometry of the packed kernel; false when the case is outside its envelope (the caller then uses
// boxblur_tile_kernel).
bool box2_plan(Box2Args& a, int w, int h, int C, int r)
{
    const int k = 2 * r + 1, wb = w * C, rC = r * C, kC = k * C;
    if ((C != 3 && C != 4) || k < 3 || k > 127 || wb % 16 || w <= r + 8 || h <= r) return false;
    a.w = w; a.h = h; a.r = r; a.wb = wb;
    a.tiles_x = (wb + 511) / 512;
    a.TXB = (((wb + a.tiles_x - 1) / a.tiles_x) + 15) & ~15;
    a.tiles_x = (wb + a.TXB - 1) / a.TXB;
    a.bn = rC & 3;
    const int S = rC - a.bn;
    a.E = S & 15;
    a.nw = (kC + 15) / 16;
    a.mcnt = 16 * a.nw - kC;
    a.O = (std::max(rC + 16, 16 * a.nw - (S - a.E)) + 15) & ~15;
    const int cN = (S - a.E + a.O) / 16;
    a.cn0 = cN - a.nw;
    const int So = a.O - a.E - C - rC;
    a.co0 = So >> 4; a.wo = (So >> 2) & 3; a.bo = So & 3;
    a.ni = (a.TXB + a.E + 15) / 16;
    int chunks = std::max(a.ni + cN + 1, a.ni + a.co0 + 2);
    chunks = std::max(chunks, (a.O + a.TXB + rC + 15) / 16 + 1);
    if (chunks % 2 == 0) ++chunks;                              // odd pitch in chunks: LDS.128 by row is conflict free
    a.pitch = 16 * chunks;
    const int rows_max = std::min(kBox2Threads, static_cast<int>(74u * 1024u / a.pitch));
    const int ty_max = rows_max - 2 * r - 1;                     // the v pass prefetches one row past the last window
    if (ty_max < 8 || ty_max < r) return false;
    const int tiles_y = (h + ty_max - 1) / ty_max;
    a.TY = (h + tiles_y - 1) / tiles_y;
    a.rows_t = a.TY + 2 * r + 1;
    const double inv = static_cast<double>(static_cast<float>(1.0 / k));
    a.inv = static_cast<float>(inv);
    a.cb = static_cast<float>((r + 0.5) / k - 32768.0 * inv);
    return true;
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
    static const bool packed_ok = !(std::getenv("BLUR_B200_BOX_PACKED") && std::atoi(std::getenv("BLUR_B200_BOX_PACKED")) == 0);
    Box2Args b{};
    if (packed_ok && img_stride % 16 == 0 && reinterpret_cast<uintptr_t>(in) % 16 == 0 && reinterpret_cast<uintptr_t>(out) % 16 == 0 &&
        box2_plan(b, w, h, C, r)) {
        b.in = in; b.out = out; b.img_stride = img_stride;
        const int tiles = b.tiles_x * ((h + b.TY - 1) / b.TY);
        const size_t smem = static_cast<size_t>(b.rows_t) * b.pitch;
        return C == 3 ? launch_box2<3>(b, tiles, batch, smem, st) : launch_box2<4>(b, tiles, batch, smem, st);
    }
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
