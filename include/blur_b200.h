/*
 * blur_b200.h — C ABI of the B200-native drop-in for the hot path of
 * michelerenzullo/Blur_algorithms (reference commit 2f873a4e).
 *
 * Plain pointers and sizes only.  Every entry point names the reference interface it
 * replaces (file:line into the reference tree).  All functions return 0 on success or a
 * BLUR_B200_E* code; blur_b200_strerror() explains it.  There is no CPU fallback: without a
 * CUDA device every compute entry point returns BLUR_B200_ECUDA.
 *
 * Pointer naming: h_* = host memory, d_* = device memory on the plan's device.
 * `stream` is a cudaStream_t passed as void* (NULL = default stream).
 */
#ifndef BLUR_B200_H
#define BLUR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    BLUR_B200_OK = 0,
    BLUR_B200_EINVAL = 1,      /* bad argument (null pointer, non-positive size, ...) */
    BLUR_B200_EPAD = 2,        /* kernel half-width exceeds min(rows,cols)-1: the reference would
                                  read out of bounds (README.md:33-38); we refuse instead */
    BLUR_B200_ESIZE = 3,       /* FFT length not plannable / exceeds shared memory (N > 28160) */
    BLUR_B200_ECUDA = 4,       /* CUDA runtime error (see blur_b200_last_cuda_error) */
    BLUR_B200_ENOMEM = 5,
    BLUR_B200_ESCRATCH = 6     /* blur_b200_fastboxblur_device_noalloc only: the image batch is larger than
                                  the library's built-in work buffer (BLUR_B200_NOALLOC_SCRATCH_BYTES) */
};

/* nyquist_mode for the Gaussian paths */
enum {
    BLUR_B200_NYQUIST_EXACT = 0,  /* every bin multiplied by its own kernel gain:
                                     pocketfft_1D (Source.cpp:362,378) and pocketfft_2D (:258) */
    BLUR_B200_NYQUIST_PFFFT = 1   /* pffft_sorted_optimized_convolution (Source.cpp:414-427):
                                     bin N/2 multiplied by the kernel's DC gain */
};

const char *blur_b200_strerror(int code);
const char *blur_b200_last_cuda_error(void);

/* ---- sizing helpers (host, exact integer semantics of the reference) --------------------- */
int blur_b200_is_valid_size(int n);                         /* isValidSize,          Utils.hpp:141-148 */
int blur_b200_nearest_transform_size(int n);                /* nearestTransformSize, Utils.hpp:150-157 */
int blur_b200_gaussian_window(double sigma, int max_width); /* gaussian_window,      Source.cpp:60-73  */
/* getGaussian, Source.cpp:75-102.  kernel must hold max(width, fft_length) floats; width 0 =
 * gaussian_window(sigma). */
int blur_b200_get_gaussian(float *kernel, double sigma, int width, int fft_length);

/* Derived sizes of the 1-D tile path (Source.cpp:434-457): out[0]=kSize out[1]=pad
 * out[2]=sizes[0] (column FFT length) out[3]=sizes[1] (row FFT length). */
int blur_b200_gauss_geometry(int rows, int cols, double sigma, int out[4]);

/* ---- Gaussian blur, 1-D tile path (pffft_ Source.cpp:429-570, pocketfft_1D :280-392) ------ */

typedef struct blur_b200_plan blur_b200_plan;

/* Plan = cached twiddles + separable kernel spectra (row and column) + device workspaces for a
 * (rows, cols, sigma, nyquist_mode) problem on `device`.  Replaces the per-call
 * pffft_new_setup / getGaussian / kernel FFT of Source.cpp:465-503. */
int blur_b200_gauss_plan_create(blur_b200_plan **plan, int rows, int cols, double sigma,
                                int nyquist_mode, int device);

/* Sharded plan for one rank of a row-stripe x column-stripe decomposition of a single image
 * (SURVEY.md §8e).  The image is rows x cols; this rank owns image rows
 * [rank*stripe_rows, min(rows, (rank+1)*stripe_rows)) in the row pass and column pairs
 * [rank*cpp, (rank+1)*cpp) in the column pass, where stripe_rows = even_ceil(rows/nranks) and
 * cpp = ceil(ceil(cols/2)/nranks) are reported by blur_b200_gauss_plan_info. */
int blur_b200_gauss_plan_create_sharded(blur_b200_plan **plan, int rows, int cols, double sigma,
                                        int nyquist_mode, int device, int rank, int nranks);
void blur_b200_gauss_plan_destroy(blur_b200_plan *plan);

/* info[0]=kSize [1]=pad [2]=n_col [3]=n_row [4]=stripe_rows [5]=cpp (column pairs per rank)
 * [6]=rank's first row [7]=rank's valid rows [8]=floats per (stripe,part) block of Q
 * [9]=kernel launches per run [10]=threads row kernel [11]=threads col kernel
 * [12]=batch row [13]=batch col [14]=smem row [15]=smem col
 * [16]=algorithm: 0 = one full-length FFT per line (the reference's tile, fft_kernels.cuh),
 *                 1 = overlap-save blocks (ols_kernels.cuh)
 * [17],[18]=overlap-save block length row/col pass [19],[20]=valid outputs per block
 * [21],[22]=blocks per line [23]=row quads per stripe */
int blur_b200_gauss_plan_info(const blur_b200_plan *plan, int64_t info[24]);

/* Host-only (no GPU needed): how a rows x cols image is cut over `nranks` GPUs and which
 * algorithm the plan will use.  out[0]=algorithm (0 full-length FFT per line, 1 overlap-save)
 * [1]=stripe_rows [2]=cpp (column pairs per rank) [3]=rank's first row [4]=rank's valid rows
 * [5]=rank's valid columns [6]=floats per (stripe,part) block of Q [7]=Q row pitch (algorithm 1).
 * Q layout exchanged between the passes (block p of the sender goes to rank p):
 *   algorithm 0: float [nranks][3][cpp][stripe_rows][2]           one all-to-all
 *   algorithm 1: float [3][nranks][stripe_rows][2*cpp]            one all-to-all per channel
 * (the reference's `resf` + flip_block, Source.cpp:536,540). */
int blur_b200_gauss_shard_geometry(int rows, int cols, double sigma, int nyquist_mode, int rank, int nranks,
                                   int64_t out[8]);

/* Whole blur, device resident: d_in and d_out are uint8 [rows][cols][3]; may alias (in place,
 * like image.data in the reference). */
int blur_b200_gauss_run_device(blur_b200_plan *plan, const uint8_t *d_in, uint8_t *d_out, void *stream);
/* 1 if the whole-image entries of this plan (run_device, run_host, run_host_batch, u8c3) use the fused
 * direct-convolution kernel instead of the two FFT passes: single-GPU plans whose kernel half width is <= 20
 * (sigma <= ~6.3; BASELINE config 1).  Same convolution (reflect-101, the float taps of getGaussian,
 * Source.cpp:75-102), evaluated tap by tap in fp32; the pass-wise entries below always use the FFT kernels. */
int blur_b200_gauss_plan_uses_direct(const blur_b200_plan *plan);
/* enable = 0: the whole-image entries of this plan use the two FFT passes from now on, like the pass-wise, sharded and
 * halo entries always do.  The direct kernel and the FFT kernels agree within 1 LSB, not bit for bit: turn the direct
 * kernel off on the single-GPU plan whenever its output is compared byte for byte with a sharded or pass-wise result of
 * a narrow kernel (half width <= 20).  enable = 1 switches it back on (BLUR_B200_ESIZE if the plan never had it).
 * No counterpart in the reference. */
int blur_b200_gauss_plan_set_direct(blur_b200_plan *plan, int enable);

/* Whole blur on host buffers (pinned or pageable); H2D + kernels + D2H.  In place allowed. */
int blur_b200_gauss_run_host(blur_b200_plan *plan, const uint8_t *h_in, uint8_t *h_out);

/* Streamed host path (the reference's `Test` loop over images, Source.cpp:574-608,
 * 625-634, blurs one host image per call): n host images of the plan's size through one plan.
 * Upload of image i+1, kernels of image i and download of image i-1 overlap (three streams,
 * double-buffered device images); inside one image the row pass follows the upload chunk by chunk
 * and finished row blocks of the column pass are downloaded while the next block is transformed.
 * Pinned host buffers are needed for the overlap (pageable ones still work).  h_out[i] may equal
 * h_in[i].  blur_b200_gauss_run_host is this call with n = 1. */
int blur_b200_gauss_run_host_batch(blur_b200_plan *plan, const uint8_t *const *h_in, uint8_t *const *h_out, int n);

/* One-shot drop-in for `pffft_(image, sigma)` / `pocketfft_1D` / `pocketfft_2D` acting on
 * image.data in place (Source.cpp:429,280,143).  The plan of the last geometry (tables and device
 * workspaces, ~7 B/px + the float32 intermediate 12 B/px) stays cached for the next call of the same
 * rows/cols/sigma/mode; blur_b200_release_cached() frees it.  Calls are serialised. */
int blur_b200_gauss_u8c3(uint8_t *h_bgr, int rows, int cols, double sigma, int nyquist_mode);
void blur_b200_release_cached(void);

/* Pieces of the sharded pipeline (single-GPU plans accept them too, with nranks = 1):
 *   row pass : d_in_stripe = this rank's rows of the image (uint8 [valid rows][cols][3])
 *              d_q         = nranks * info[8] floats in the layout of blur_b200_gauss_shard_geometry
 *                            (zero it once before the first call: padding columns are never written)
 *   col pass : d_q_chunks  = the same layout after the exchange (block s came from rank s)
 *              d_out_cols  = uint8 [rows][2*cpp][3]  this rank's column stripe of the result */
int blur_b200_gauss_row_pass(blur_b200_plan *plan, const uint8_t *d_in_stripe, float *d_q, void *stream);
int blur_b200_gauss_col_pass(blur_b200_plan *plan, const float *d_q_chunks, uint8_t *d_out_cols, void *stream);

/* Row pass with the exchange fused into its epilogue (overlap-save plans, nranks <= 8): q_peers[r]
 * is the address, mapped into THIS process (CUDA IPC / symmetric memory over NVLink), of rank r's
 * receive buffer (nranks * info[8] floats, the layout d_q_chunks of the column pass expects).  Every
 * result is stored straight into the buffer of the rank that owns its column, as block `rank`;
 * nothing is staged in local HBM and no collective runs.  The caller synchronises the ranks
 * (all row passes finished) before any column pass reads its buffer, and again before the next
 * row pass overwrites it.  Returns BLUR_B200_ESIZE for plans that use the full-length algorithm. */
int blur_b200_gauss_row_pass_peer(blur_b200_plan *plan, const uint8_t *d_in_stripe, float *const *q_peers, int npeers,
                                  void *stream);

/* Row pass pipelined with the exchange (overlap-save plans, nranks <= 8): the stripe is cut into
 * `nchunks` wave-aligned groups of row pairs; each group is transformed into the plan's local send buffer (the
 * part this rank owns goes straight into q_peers[rank]) and, while the next group is being transformed, the
 * GPU's copy engines push the finished group's blocks into the owners' receive buffers q_peers[r] (one 2-D peer
 * copy per destination and group over the NVLink peer mappings).  On return `stream` waits for every copy.
 * Same caller-side synchronisation between ranks as blur_b200_gauss_row_pass_peer. */
int blur_b200_gauss_row_pass_push(blur_b200_plan *plan, const uint8_t *d_in_stripe, float *const *q_peers, int npeers,
                                  int nchunks, void *stream);

/* Row-stripe ("halo") sharding of one image: every rank keeps a stripe of whole column-pass blocks, runs the row
 * pass on it, receives only `halo` rows of the float32 intermediate from the stripe above and below, and runs the
 * column pass on its own blocks: input AND output row-sharded, 2*halo*cols*12 bytes exchanged per GPU (65 MB
 * instead of 352 MB at 16384^2 sigma=50 on 8 GPUs), output bit-identical to the single-GPU plan.  The entries take
 * an ordinary single-GPU plan of the WHOLE image (exact mode, two-factor kernels: BLUR_B200_ESIZE otherwise);
 * rank / nranks select the stripe.  info = {row0, valid_rows, stripe capacity, halo, qpitch, floats of Q_local,
 * first block, blocks}.  Q_local (caller's, peer-mapped for the neighbours): float [3][capacity + 2*halo][qpitch].
 * Sequence: halo_row_pass -> halo_push (q_up / q_down = the neighbours' Q_local, NULL at the image edges) ->
 * cross-rank barrier -> halo_col_pass (d_out_rows: uint8 [valid_rows][cols][3]).  The halo variant of the
 * distributed flip_block (Source.cpp:540); the contract configuration (row stripes -> all-to-all -> column
 * stripes) is the sharded plan above. */
int blur_b200_gauss_halo_info(const blur_b200_plan *plan, int rank, int nranks, int64_t info[8]);
int blur_b200_gauss_halo_row_pass(blur_b200_plan *plan, int rank, int nranks, const uint8_t *d_in_stripe, float *d_q_local, void *stream);
int blur_b200_gauss_halo_push(blur_b200_plan *plan, int rank, int nranks, const float *d_q_local, float *q_up, float *q_down, void *stream);
int blur_b200_gauss_halo_col_pass(blur_b200_plan *plan, int rank, int nranks, const float *d_q_local, uint8_t *d_out_rows, void *stream);

/* pffft-mode (BLUR_B200_NYQUIST_PFFFT) SHARDED overlap-save plans: the Nyquist quirk of Source.cpp:414-427 is a
 * per-line correction; a stripe's row corrections are local, the column corrections need two small sums over all
 * rows.  Per blur on every rank: blur_b200_gauss_quirk_local(stripe) -> row pass (any exchange entry above) ->
 * the caller all-reduces (SUM) the two buffers blur_b200_gauss_quirk_buffers returns (3*rows floats, of which
 * every element is non-zero on one rank only, and nseg*3*cols ints: the sums are exact) ->
 * blur_b200_gauss_quirk_finish -> column pass.  Single-GPU plans do all of this inside the row pass. */
int blur_b200_gauss_quirk_buffers(blur_b200_plan *plan, float **d_corr_row, int64_t *n_corr_row, int **d_insum, int64_t *n_insum);
int blur_b200_gauss_quirk_local(blur_b200_plan *plan, const uint8_t *d_in_stripe, void *stream);
int blur_b200_gauss_quirk_finish(blur_b200_plan *plan, void *stream);

/* Staged exchange (two-factor overlap-save plans, nranks <= 8): the exchange hidden behind BOTH FFT passes.
 * Every rank's part (2*cpp columns) is cut into `nstages` column sub-ranges of w = 2*cpp / nstages columns.  The
 * row pass runs as nstages launches; launch k transforms, for every row of the stripe, the overlap-save blocks that
 * touch sub-range k of any rank's part, and as soon as it has finished the copy engines push sub-range k into the
 * owners' receive buffers q_peers[r] while launch k+1 computes.  blur_b200_gauss_stage_wait makes `stream` wait
 * for this rank's outgoing copies of one stage; after the caller's cross-rank barrier for that stage,
 * blur_b200_gauss_col_pass_stage transforms and interleaves sub-range k of this rank's column stripe — while the
 * later stages are still on the wire.  Receive-buffer layout for this mode: [stage][3][nranks][stripe_rows][w]
 * (same size as the unstaged layout).  blur_b200_gauss_staged_supported: BLUR_B200_OK, or BLUR_B200_ESIZE when the
 * plan / stage count cannot run staged (use blur_b200_gauss_row_pass_push then).  Replaces flip_block, Source.cpp:540.
 * store != 0: no copy engines and no send buffer — launch k's epilogue writes every sub-range straight into its owner's
 * buffer with bulk shared->global copies over NVLink (the SM-issued path sustains more all-to-all bandwidth than the
 * copy engines at 8 GPUs); blur_b200_gauss_stage_wait then waits for launch k itself.  row_ctas_per_sm /
 * col_ctas_per_sm > 0 hold a launch to that many CTAs per SM so that the column pass of sub-range k (issued on a
 * second stream) runs beside row launch k+1 while that one waits on the links; 0 = the kernels' own occupancy. */
int blur_b200_gauss_staged_supported(const blur_b200_plan *plan, int nstages);
int blur_b200_gauss_row_pass_staged(blur_b200_plan *plan, const uint8_t *d_in_stripe, float *const *q_peers, int npeers,
                                    int nstages, int store, int row_ctas_per_sm, void *stream);
int blur_b200_gauss_stage_wait(blur_b200_plan *plan, int stage, void *stream);
int blur_b200_gauss_col_pass_stage(blur_b200_plan *plan, const float *d_q_recv, uint8_t *d_out_cols, int stage, int nstages,
                                   int col_ctas_per_sm, void *stream);

/* Parity hook: the float32 row-pass result (`resf` of Source.cpp:536) as planar [3][rows][cols]. */
int blur_b200_gauss_rowpass_f32(blur_b200_plan *plan, const uint8_t *d_in, float *d_planar, void *stream);

/* ---- FFT box blur: the reference built with `#define boxblur` (Source.cpp:12) ------------------
 * pffft_ / pocketfft_1D / pocketfft_2D then convolve with a box of width kLen convolved with itself
 * (box_kernel, Source.cpp:128-140: a tent of 2*kLen-1 taps) instead of the Gaussian; kLen =
 * (int)(ns'*ns'), ns' = sqrt(min((int)nsmooth^2, min(rows-1, cols-1))) (:287-291, :437-441), pad = kLen-1.
 * The plan is an ordinary blur_b200_plan: every blur_b200_gauss_run_* entry applies.  Odd kLen >= 3
 * only (an even kLen wraps its outermost taps around the tile in the reference): BLUR_B200_EINVAL. */
int blur_b200_boxfft_klen(int rows, int cols, double nsmooth);
int blur_b200_boxfft_plan_create(blur_b200_plan **plan, int rows, int cols, double nsmooth, int nyquist_mode, int device);
int blur_b200_boxfft_u8c3(uint8_t *h_bgr, int rows, int cols, double nsmooth, int nyquist_mode);

/* ---- FastBoxBlur (fastboxblur(in,w,h,channels,ksize,passes), call site Source.cpp:587) ------
 * Declared semantics (upstream source absent): r=(ksize-1)/2, k=2r+1,
 * reflect-101, horizontal then vertical sliding sums, uint8 after each 1-D pass with
 * (2*acc+k)/(2k), `passes` repeats, in place.  `batch` images back to back. */
int blur_b200_fastboxblur_device(uint8_t *d_img, int w, int h, int channels, int ksize, int passes,
                                 int batch, uint8_t *d_scratch /* batch*w*h*channels bytes */, void *stream);
int blur_b200_fastboxblur(uint8_t *h_img, int w, int h, int channels, int ksize, int passes);
/* `batch` contiguous host images (the C5 workload; the reference calls fastboxblur once per image,
 * Source.cpp:587), streamed in groups: uploads, kernels and downloads overlap.  In place. */
int blur_b200_fastboxblur_batch_host(uint8_t *h_imgs, int w, int h, int channels, int ksize, int passes, int batch);

/* FastBoxBlur for callers that cannot allocate GPU memory (no counterpart in the reference, whose fastboxblur
 * works on host memory; added for this library's users).  Same arguments and byte-identical results as
 * blur_b200_fastboxblur_device, minus d_scratch: the second buffer is a fixed array of
 * BLUR_B200_NOALLOC_SCRATCH_BYTES that is part of the library's device image, so the call itself never
 * allocates or frees anything.  Meant for plug-in nodes that only get their input/output surfaces from a
 * host application's memory pool, and for real-time loops where cudaMalloc is off limits once running.
 *   - w*h*channels*batch must not exceed BLUR_B200_NOALLOC_SCRATCH_BYTES: BLUR_B200_ESCRATCH otherwise.
 *   - Runs on the calling thread's current GPU (d_img must live there).  The first call on each GPU resolves
 *     the array's address and may load the kernels: make it at set-up, before any stream capture begins.
 *   - One array per GPU, shared by every call: calls that can overlap on the GPU (different streams or
 *     threads) corrupt each other — issue them one after another.  Old contents are harmless; every pass
 *     writes its whole destination first, so nothing is zeroed between calls.
 *   - Each GPU that runs any kernel of this library holds the 32 MiB until the process exits. */
#define BLUR_B200_NOALLOC_SCRATCH_BYTES (32u << 20)
int blur_b200_fastboxblur_device_noalloc(uint8_t *d_img, int w, int h, int channels, int ksize, int passes,
                                         int batch, void *stream);

/* ---- layout utilities (device pointers) ------------------------------------------------------ */
/* Reflect_101<uint8_t,C>, Utils.hpp:212-243.  original_size = {rows, cols}.  Pads are clamped to
 * dim-1 like the reference; d_out must hold (rows+top+bottom)*(cols+left+right)*channels bytes
 * computed with the CLAMPED pads. */
int blur_b200_reflect_101_u8(const uint8_t *d_in, uint8_t *d_out, int top, int bottom, int left, int right,
                             const int original_size[2], int channels, void *stream);
/* same on host buffers (H2D + kernel + D2H): the signature of the reference template */
int blur_b200_reflect_101_u8_host(const uint8_t *h_in, uint8_t *h_out, int top, int bottom, int left, int right,
                                  const int original_size[2], int channels);
/* deinterleave_BGR<uint8_t,float> / interleave_BGR<uint8_t,float>, Utils.hpp:159-210 */
int blur_b200_deinterleave_u8_f32(const uint8_t *d_in, float *d_b, float *d_g, float *d_r, uint32_t total, void *stream);
int blur_b200_interleave_f32_u8(const float *d_b, const float *d_g, const float *d_r, uint8_t *d_out, uint32_t total, void *stream);
/* flip_block<float,1>(in,out,w,h): in is h x w, out is w x h (call sites Source.cpp:540,562) */
int blur_b200_flip_block_f32(const float *d_in, float *d_out, int w, int h, void *stream);
/* the same three on HOST buffers (H2D + kernel + D2H): the pointer kinds of the reference's own
 * signatures (Utils.hpp:159-160,186-187; flip_block call sites Source.cpp:367,384,540,562) */
int blur_b200_deinterleave_u8_f32_host(const uint8_t *h_in, float *h_b, float *h_g, float *h_r, uint32_t total);
int blur_b200_interleave_f32_u8_host(const float *h_b, const float *h_g, const float *h_r, uint8_t *h_out, uint32_t total);
int blur_b200_flip_block_f32_host(const float *h_in, float *h_out, int w, int h);

/* ---- the layout templates for other element types (Utils.hpp:159-210 are templates on T and U; flip_block<T,C>) ---- */
/* element type codes */
enum { BLUR_B200_T_U8 = 0, BLUR_B200_T_U16 = 1, BLUR_B200_T_I32 = 2, BLUR_B200_T_F32 = 3 };
/* deinterleave_BGR<T,U>: planes[c][x] = (U)(in[3x+c] + round), round = 0.5f iff U integral and T floating
 * (Utils.hpp:163); in_type = T, out_type = U.  Float -> integral truncates toward zero and saturates. */
int blur_b200_deinterleave_typed(const void *d_in, void *d_b, void *d_g, void *d_r, uint32_t total, int in_type, int out_type,
                                 void *stream);
/* interleave_BGR<T,U>: out[3x+c] = (T)(planes[c][x] + round), round = 0.5f iff T integral and U floating
 * (Utils.hpp:189); plane_type = U, out_type = T. */
int blur_b200_interleave_typed(const void *d_b, const void *d_g, const void *d_r, void *d_out, uint32_t total, int plane_type,
                               int out_type, void *stream);
/* flip_block<T,C>(in,out,w,h) for any pixel of pixel_bytes = C*sizeof(T) in {1,2,3,4,6,8,12,16} bytes */
int blur_b200_flip_block_pixels(const void *d_in, void *d_out, int w, int h, int pixel_bytes, void *stream);
/* host-buffer forms (H2D + kernel + D2H) */
int blur_b200_deinterleave_typed_host(const void *h_in, void *h_b, void *h_g, void *h_r, uint32_t total, int in_type, int out_type);
int blur_b200_interleave_typed_host(const void *h_b, const void *h_g, const void *h_r, void *h_out, uint32_t total, int plane_type,
                                    int out_type);
int blur_b200_flip_block_pixels_host(const void *h_in, void *h_out, int w, int h, int pixel_bytes);

/* ---- DFT_image spectrum path (pocketfft_2D with #define DFT_image, Source.cpp:240-252) ------- */
/* d_in uint8 [rows][cols][3]; d_out uint8 [rows][cols][3] spectrum image; optional d_spec_f32
 * float planar [3][H][W] of the padded size (H,W returned in padded[2]). nsmooth selects the
 * padding exactly as pocketfft_2D does. */
int blur_b200_dft_image_device(const uint8_t *d_in, uint8_t *d_out, int rows, int cols, double nsmooth,
                               float *d_spec_f32, int padded[2], void *stream);
int blur_b200_dft_image_host(const uint8_t *h_in, uint8_t *h_out, int rows, int cols, double nsmooth,
                             float *h_spec_f32, int padded[2]);
/* padded = {sizes[0], sizes[1]}, border = {top, bottom, left, right} of Source.cpp:160-176 */
int blur_b200_dft_image_sizes(int rows, int cols, double nsmooth, int padded[2], int border[4]);

/* Number of kernels this library has launched in this process (bench.py reports it). */
uint64_t blur_b200_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif
