/* p2p_b200.h -- C ABI of libp2p_b200.so: the B200 (sm_100a) hot path of ishine/pix2pixHDAudioSR.
 *
 * The reference has no FFI on this path (it is Python calling torch.fft / torch.nn); these entry
 * points are what a replacement for the bodies of
 *     models/mdct.py:486-513   MDCT4.forward          -> p2p_mdct4_analysis
 *     models/mdct.py:542-566   IMDCT4.forward         -> p2p_mdct4_synthesis
 *     autograd of the two        -> the same two kernels, roles swapped
 *     models/mdct.py:483-484 / 539-540 (twiddles)     -> p2p_mdct4_fill_tables
 * binds.  INTEGRATION.md shows the ctypes stub a reference maintainer would add.
 *
 * Conventions: every function returns 0 on success, a negative P2P_E* code otherwise and never
 * throws; p2p_last_error() gives the message for the calling thread.  All device work is enqueued
 * asynchronously on the caller's stream (a cudaStream_t passed as void*).  All buffers are owned
 * by the caller (device pointers unless the name says host); nothing is allocated, nothing
 * synchronises.  Pointers must be 4-byte aligned; 16-byte alignment enables the vector paths.
 */
#ifndef P2P_B200_H
#define P2P_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define P2P_OK 0
#define P2P_EINVAL (-1)   /* bad argument (message says which) */
#define P2P_ECUDA (-2)    /* a CUDA runtime call or launch failed */
#define P2P_ENOTSUP (-3)  /* shape outside what the kernels cover */

const char* p2p_last_error(void);
int p2p_abi_version(void);
/* Number of kernel launches this process has issued through the entry points below (bench.py's gpu_launches). */
long long p2p_kernel_launches(void);
/* Deterministic mode (also: environment P2P_DETERMINISTIC=1): every floating-point reduction runs with one writer per
 * address -- no split-K / split-pixel red.add between CTAs, single-chunk statistics, single-CTA loss sums -- so results are
 * bit-reproducible run to run, at lower parallelism on the few-tile layers.  on < 0 only queries.  Returns the old value. */
int p2p_set_deterministic(int on);

/* ---- MDCT4 / IMDCT4 ------------------------------------------------------------------------
 * M = n_fft/2 bins, L = M/2 = complex FFT length.  n_fft: power of two, 16..4096.
 * Table layout (float32 words): pre[2L] post[2L] roots[2L] perm[L as int32] xtw[2L] -> 9L words
 * (xtw: exchange twiddles of the register-FFT fast path, zero when n_fft < 256).
 */
/* Testing/benchmark knob: non-zero forces the general shared-memory kernels even where the
 * register-FFT fast path (win == n_fft, hop == n_fft/2, n_fft >= 256) applies.  Returns the old value. */
int p2p_mdct4_force_general(int on);
int p2p_mdct4_table_words(int n_fft);
/* host_tables: HOST buffer of p2p_mdct4_table_words(n_fft) 4-byte words, computed in double. */
int p2p_mdct4_fill_tables(int n_fft, void* host_tables);

/* spec[b,f,k] = scale * sum_{n<win} window[n] * x[b, f*hop - offset + n] * cos(pi/M (n+1/2+M/2)(k+1/2)),
 * reads of x outside [0,T) are zero.  x [B,T] fp32, spec [B,F,M] fp32 (contiguous).
 * MDCT4.forward: offset = start_pad, scale = 1.  IMDCT4 backward: x = grad_out, offset = crop_left,
 * scale = 4/n_fft. */
int p2p_mdct4_analysis(const float* x, float* spec, const float* window, const void* tables,
                       int B, int64_t T, int F, int n_fft, int hop, int win, int64_t offset,
                       float scale, void* stream);

/* out[b,s] = scale * sum_f window[n] * y_f[n],  n = s + crop_left - f*hop in [0,win),
 * y_f[n] = sum_k spec[b,f,k] cos(pi/M (n+1/2+M/2)(k+1/2)).  spec [B,F,M], out [B,T_out].
 * IMDCT4.forward: crop_left = win/2 (center) or 0, scale = 4/n_fft.  MDCT4 backward: spec = grad,
 * crop_left = start_pad, T_out = T, scale = 1. */
int p2p_mdct4_synthesis(const float* spec, float* out, const float* window, const void* tables,
                        int B, int F, int64_t T_out, int n_fft, int hop, int win, int64_t crop_left,
                        float scale, void* stream);


/* ---- pix2pixHD conv GAN layers ---------------------------------------------------------------
 * Replaces the torch.nn / cuDNN calls under models/networks.py:129-383 (GlobalGenerator, LocalEnhancer,
 * ResnetBlock, NLayerDiscriminator, MultiscaleDiscriminator) and the loss / optimiser arithmetic of
 * networks.py:68-110, pix2pixHD_model.py:390-398, train.py:155-184.
 *
 * Activations are NHWC ("channels last"), dtype P2P_F32 or P2P_BF16; statistics, losses, master weights,
 * weight gradients and optimiser state are always fp32.  Weights are tap-major: W[tap][row][col] with
 * tap = kh*KW + kw and (row, col) strides given per call, so Conv2d ([Cout,Cin,KH,KW] logical) and
 * ConvTranspose2d ([Cin,Cout,KH,KW] logical) share one physical layout [KH][KW][Cout][Cin].
 */
#define P2P_F32 0
#define P2P_BF16 1
#define P2P_PAD_ZERO 0
#define P2P_PAD_REFLECT 1
#define P2P_MAX_TAPS 49

/* One implicit-GEMM pass:  out[n, y*oys+oyo, x*oxs+oxo, j] (+)= bias[j] +
 *     sum_{t<ntaps} sum_{k<K} in[n, pad(y*sy + dh[t]), pad(x*sx + dw[t]), k] * W[wtap[t]*w_tap + k*w_k + j*w_j]
 * for n<N, y<gy, x<gx, j<J.  pad(): zero outside [0,H)x[0,W) or reflection (no edge repeat).
 * A stride-1/2 convolution is one pass; a stride-2 transposed convolution (and the input gradient of a
 * stride-2 convolution) is four passes, one per output parity, each with its own tap subset. */
typedef struct {
    int N, H, W, K;          /* gathered tensor: NHWC with K channels */
    int OH, OW, J;           /* full extent of the written tensor: NHWC with J channels */
    int gy, gx;              /* GEMM rows per image */
    int sy, sx;              /* gather stride */
    int oys, oyo, oxs, oxo;  /* scatter stride / offset of the written position */
    int ntaps;
    int pad_mode;
    long long w_tap, w_k, w_j;   /* weight strides in elements */
    signed char dh[P2P_MAX_TAPS], dw[P2P_MAX_TAPS];
    unsigned char wtap[P2P_MAX_TAPS];
} p2p_conv_pass;

/* out = pass(in, W) (+ bias).  in and W have dtype `dtype`, out has `out_dtype` (a bf16 layer that feeds an
 * InstanceNorm keeps its pre-normalisation output in fp32: rounding y to bf16 before subtracting the
 * per-plane mean would cost |mean|/std times the bf16 epsilon); bias fp32 or NULL. */
int p2p_conv_forward(const p2p_conv_pass* pass, const void* in, const void* weight, const float* bias,
                     void* out, int dtype, int out_dtype, void* stream);
/* Testing/benchmark knob: non-zero routes every convolution pass to the exact CUDA-core kernels even where the
 * tcgen05 tensor-core kernels (bf16, gathered channels % 16 == 0) apply; negative only queries.  Returns the old value. */
int p2p_conv_force_simple(int on);
/* Number of tcgen05 convolution kernels launched by this process so far. */
long long p2p_conv_umma_launches(void);
/* dW[wtap[t]*w_tap + k*w_k + j*w_j] += sum_{n,y,x} in[n, pad(..t..), k] * dout[n, y*oys+oyo, x*oxs+oxo, j]
 * (fp32 accumulate with red.add into the caller's fp32 buffer -- zeroed, or a live gradient buffer being
 * accumulated; safe to call once per pass of a layer). */
int p2p_conv_wgrad(const p2p_conv_pass* pass, const void* in, const void* dout, float* dweight, int dtype,
                   void* stream);
/* dbias[j] = sum over rows of dout [rows, J] (overwrites). */
int p2p_bias_grad(const void* dout, float* dbias, long long rows, int J, int dtype, void* stream);
/* Stride-1 convolutions as a TMA-fed shifted-window implicit GEMM (csrc/conv_window.cu): the padding is written once
 * (p2p_pad_nhwc: xp [N, H+pt+pb, W+pl+pr, C], zero or reflection; needs C * element size % 16 == 0), then
 *   out[n,y,x,j] (+ bias[j]) = sum_{dh<KH, dw<KW} sum_c xp[n, y+dh, x+dw, c] * W[tap, c, j],   y < Hp-KH+1, x < Wp-KW+1,
 * bf16 operands, fp32 accumulation, out fp32 or bf16 [N, Hp-KH+1, Wp-KW+1, J].  tap = dh*KW+dw, or the flipped tap when
 * `flip` (input gradients); W element (tap, c, j) at tap*w_tap + c*w_k + j*w_j with w_k == 1 or w_j == 1.
 * P2P_ENOTSUP when C % 16 != 0 or the window (128 + (KH-1)*Wp + KW-1 pixel rows) exceeds 320 rows. */
int p2p_pad_nhwc(const void* x, void* xp, int N, int H, int W, int C, int pt, int pb, int pl, int pr, int reflect,
                 int dtype, void* stream);
int p2p_conv_window(const void* xp, const void* weight, const float* bias, void* out, int N, int Hp, int Wp, int C, int J,
                    int KH, int KW, int flip, long long w_tap, long long w_k, long long w_j, int out_dtype, void* stream);
/* General window form (strided layers and transposed-convolution / strided-dgrad parity passes).  planes: nplanes (<= 4)
 * tensors [N,Hp,Wp,C] stored back to back -- p2p_pad_planes writes plane (a,b) = zero-padded x sampled at
 * (s*y + a - pt, s*x + b - pl), s in {1,2}.  Up to 4 sub-passes (output parities) share the planes in ONE launch: tap t
 * belongs to sub-pass tap_sub[t] (taps grouped by sub-pass, then sorted by plane), reads plane tap_plane[t] at
 * (y + tap_dh[t], x + tap_dw[t]) and multiplies weight tap tap_w[t] (< wtaps_total); rows (y < gy[u], x < gx[u]) of sub-pass u
 * are written to out[n, y*oys+oyo[u], x*oxs+oxo[u], :] of an [N,OHf,OWf,J] tensor. */
typedef struct {
    int nplanes, N, Hp, Wp, C;
    int OHf, OWf, J;
    int wtaps_total;
    long long w_tap, w_k, w_j;
    int nsub, oys, oxs;
    int gy[4], gx[4], oyo[4], oxo[4];
    int ntaps;
    signed char tap_sub[P2P_MAX_TAPS], tap_plane[P2P_MAX_TAPS];
    short tap_dh[P2P_MAX_TAPS], tap_dw[P2P_MAX_TAPS];
    unsigned char tap_w[P2P_MAX_TAPS];
    int wrap;   /* != 0: taps may read past their sub-pass grid (y + dh >= Hp wraps into the next rows / image of the flattened
                 * plane): the caller guarantees those pixels hold zeros.  Used by the input-gradient pass that shares the
                 * weight-gradient's copy of dout (dout at the pitch of the padded input, zero prefix before the first image) */
    long long plane_rows;   /* rows between consecutive planes; 0 = N*Hp*Wp (densely packed).  The window weight-gradient kernel
                             * wants planes of a multiple of 64 rows with a zero tail: one copy then serves both passes */
} p2p_conv_window_desc;
int p2p_pad_planes(const void* x, void* planes, int N, int H, int W, int C, int pt, int pl, int Hh, int Wh, int s,
                   long long plane_rows /* >= N*Hh*Wh; the tail of every plane is zero-filled */, int reflect /* s == 1 only */,
                   int dtype, void* stream);
/* Weight gradients in window form:  dW[tap_w[t]][j][c] (fp32 [wtaps_total][J][C], accumulated with red.add) +=
 *   sum_q aplanes[tap_a[t]][q][j] * bplanes[tap_b[t]][q + tap_dh[t]*Wp + tap_dw[t]][c]
 * over the padded pixel grid: aplanes = the output gradient laid out at the pitch of the padded input (zero where no output
 * exists), bplanes = the padded input (or its parity planes).  plane_rows % 64 == 0; taps grouped by (tap_a, tap_b); <= 32 taps.
 * dw_is_zero != 0 promises that dW holds zeros (first gradient after zero_grad): tiles with a single writer are then stored
 * instead of red.add-ed.  dw_is_zero == 2: dW holds STALE values and must be overwritten (the caller skipped its memset,
 * optimizer.zero_grad of train.py:164-184): one writer per tile is forced.  p2p_conv_wgrad_window_stored() tells whether the
 * calling thread's last call stored (1) or accumulated (0). */
int p2p_conv_wgrad_window_stored(void);
int p2p_conv_wgrad_window(const void* aplanes, int n_aplanes, const void* bplanes, int n_bplanes, long long plane_rows, int Wp,
                          int J, int C, float* dw, int wtaps_total, int ntaps, const int* tap_a, const int* tap_b,
                          const int* tap_dh, const int* tap_dw, const int* tap_w, int dw_is_zero, void* stream);
/* Debug hook: builds with -DP2P_WINDOW_TIMING write, per CTA, the MMA warp's cycles {waiting for the window, waiting for weight
 * stages, issuing, waiting for the accumulator, stages, total, tiles, -} to buf (8 x int64 per CTA); a no-op otherwise. */
int p2p_conv_window_debug_buffer(long long* buf);
int p2p_conv_window_run(const p2p_conv_window_desc* d, const void* planes, const void* weight, const float* bias, void* out,
                        int out_dtype, void* stream);
/* The same launches with InstanceNorm's statistics taken in the epilogue (replaces the separate reduction over the conv
 * output that models/networks.py:190-207's `norm_layer` implies): stats [N][J][2] fp32 is zeroed and then receives, per
 * (image, output channel), the sum and the sum of squares of (out - bias) over the written pixels.  Needs an fp32 output,
 * J % 4 == 0 and an unsplit channel loop (enough tiles to fill the chip): P2P_ENOTSUP otherwise, nothing launched. */
int p2p_conv_window_stats(const void* xp, const void* weight, const float* bias, void* out, int N, int Hp, int Wp, int C, int J,
                          int KH, int KW, int flip, long long w_tap, long long w_k, long long w_j, int out_dtype, float* stats,
                          void* stream);
int p2p_conv_window_run_stats(const p2p_conv_window_desc* d, const void* planes, const void* weight, const float* bias, void* out,
                              int out_dtype, float* stats, void* stream);
/* z = act(InstanceNorm(y)) + res written DIRECTLY as the next convolution's input: s*s planes of plane_rows rows x C bf16 in
 * the geometry of p2p_pad_planes (plane (a,b)[n,yy,xx,:] = z[n, s*yy+a-pt, s*xx+b-pl, :], zero or -- s == 1 -- reflection
 * outside, zero tail), and optionally the plain bf16 NHWC tensor z_out too.  y [N,H,W,C] of y_dtype.  stats_in [N,C,2]:
 * raw_mode 1 = the (sum, sum of squares) of (y - bias) from p2p_conv_window_*_stats, stats_out [N,C,2] then receives
 * (mean, rstd) for p2p_instnorm_act_bwd; raw_mode 0 = (mean, rstd) already.  res: bf16 with element strides per image / row /
 * pixel (channels contiguous) or NULL.  C % 8 == 0. */
int p2p_instnorm_act_planes(const void* y, const float* bias, const float* stats_in, int raw_mode, float* stats_out,
                            const void* res, long long res_sn, long long res_sh, long long res_sw, void* planes, void* z_out,
                            int N, int H, int W, int C, int pt, int pl, int Hh, int Wh, int s, long long plane_rows, int reflect,
                            int norm, int act, float slope, float eps, int y_dtype, void* stream);
/* Adjoint of reflection padding: dx[n,h,w,c] = sum of dxp over the padded positions that mirror onto (h,w).
 * dxp is [N, H+2p, W+2p, C], dx is [N,H,W,C]. */
int p2p_reflect_pad_fold(const void* dxp, void* dx, int N, int H, int W, int C, int p, int dtype, void* stream);
/* the same reading an fp32 padded gradient (what the split-K input-gradient pass accumulates) and writing `dtype` */
int p2p_reflect_pad_fold_f32(const float* dxp, void* dx, int N, int H, int W, int C, int p, int dtype, void* stream);

/* Operand packing for the thin layers (networks.py:145-154 G's 7x7 2->ngf / ngf->2 convolutions, :340-361 D's 4->ndf and
 * ->1 logit convolutions): KW*C <= 16 kernel-column x channel pairs become the 16 tensor-core channels of a KHx1 convolution.
 *   pack_in:  xpk[n, y, ox, dw*C + c]      = pad(x)[n, y, ox*stride + dw, c]          xpk: [N][H+2pad][OW][16]
 *   fold_out: y[n, oh, ow, co]             = bias[co] + sum_dw ypk[n, oh, ow + dw, dw*Cout + co]   ypk: fp32 [N][OH][Wp][16]
 *   pack_dy:  dypk[n, y, x', dw*Cout + co] = dy[n, y, x' - dw, co] (0 outside)        dypk: [N][OH][Wp][16] */
int p2p_pack_in(const void* x, void* xpk, int N, int H, int W, int C, int KW, int stride, int pad, int reflect, int OW,
                int dtype, void* stream);
int p2p_fold_out(const float* ypk, const float* bias, void* y, int N, int OH, int Wp, int OW, int KW, int Cout, int out_dtype,
                 void* stream);
/* the same with the kernel ROWS packed into the channels (taps shift along a pixel row, so the shifted-window kernel runs it):
 *   fold_out_rows: y[n, oh, ow, co] = bias[co] + sum_dh ypk[n, oh + dh, ow, dh*Cout + co]   ypk: fp32 [N][Hp][Wy][16], KH*Cout <= 16 */
int p2p_fold_out_rows(const float* ypk, const float* bias, void* y, int N, int Hp, int Wy, int OH, int OW, int KH, int Cout,
                      int out_dtype, void* stream);
int p2p_pack_dy(const void* dy, void* dypk, int N, int OH, int OW, int Cout, int KW, int Wp, int dtype, void* stream);
/* Full-patch forms (bf16 only; KH*KW*C <= Kc <= 128, Kc % 8 == 0): the thin side's whole patch becomes one channel vector,
 * so the layer is a 1x1 convolution for the window kernels.  Buffers have rows_alloc rows (a multiple of 64 for the
 * window wgrad kernel); rows beyond the tensor and channels beyond the patch are written as zero.
 *   im2col:       xcol[(n,oy,ox), (kh,kw,c)]    = pad(x)[n, oy*stride + kh, ox*stride + kw, c]
 *   col2im:       dx[n,y,x,c]                   = sum_{kh,kw} dxcol[(n,oy,ox), (kh,kw,c)] with oy*stride + kh - pad = y (zero padding)
 *   pack_dy_full: out[(n,yp,xp), (kh,kw,co)]    = dy[n, yp - kh, xp - kw, co] on the padded grid Hp = OH+KH-1, Wp = OW+KW-1 */
int p2p_im2col(const void* x, void* xcol, int N, int H, int W, int C, int KH, int KW, int stride, int pad, int reflect, int OH,
               int OW, int Kc, long long rows_alloc, void* stream);
int p2p_col2im(const void* dxcol, void* dx, int N, int H, int W, int C, int KH, int KW, int stride, int pad, int OH, int OW,
               int Kc, void* stream);
int p2p_pack_dy_full(const void* dy, void* out, int N, int OH, int OW, int Cout, int KH, int KW, int Kc, long long rows_alloc,
                     void* stream);

/* InstanceNorm2d(affine=False, eps) [+ activation] [+ residual]:  z = act((y-mean)/sqrt(var+eps)) + res.
 * act: 0 none, 1 ReLU, 2 LeakyReLU(slope).  norm=0 skips the normalisation (plain activation).
 * y: [N,HW,C] `y_dtype`; res,z: `dtype`; stats: fp32 [N,C,2] = (mean, rstd), written by fwd, read by bwd. */
int p2p_instnorm_act_fwd(const void* y, const void* res, void* z, float* stats, int N, long long HW, int C,
                         int norm, int act, float slope, float eps, int y_dtype, int dtype, void* stream);
/* dy from dz (and dres = dz is the caller's business): needs y and stats of the forward.
 * work: fp32 scratch [N,C,2] (ignored when norm == 0). */
int p2p_instnorm_act_bwd(const void* y, const float* stats, const void* dz, void* dy, float* work, int N,
                         long long HW, int C, int norm, int act, float slope, int y_dtype, int dtype, void* stream);
/* p2p_instnorm_act_bwd that writes dy directly in the layout the stride-1 gradient kernels of the conv in front of the
 * normalisation read (the adjoint of ReflectionPad2d / zero padding of models/networks.py:196-204 folded into the producer):
 * grid [rows, C], image n at pixel row n*Hh*Wh, dy[n,h,w,:] at + h*Wh + w, zeros everywhere else (Hh >= H, Wh >= W,
 * rows >= N*Hh*Wh).  C % 4 == 0 and 16-byte aligned pointers, else P2P_ENOTSUP. */
int p2p_instnorm_act_bwd_wrap(const void* y, const float* stats, const void* dz, void* grid, float* work, int N, int H,
                              int W, int C, int Hh, int Wh, long long rows, int norm, int act, float slope, int y_dtype,
                              int dtype, void* stream);
/* Caller-owned, ZEROED device scratch (ws_floats >= 2*N*C floats, cnt_words >= N*ceil(C/128) 32-bit words, for the largest
 * tensor) that lets the InstanceNorm reductions of p2p_instnorm_act_fwd / _bwd / p2p_instnorm_stats run as ONE launch each
 * (partial sums meet in ws, the last CTA finalises and re-zeroes it) instead of memset + reduction + finalise.  The calls
 * that use it must be ordered on one stream; NULL pointers unregister it. */
int p2p_instnorm_workspace(float* ws, long long ws_floats, unsigned* cnt, long long cnt_words);
/* Bytes of those two regions for tensors of up to N images x C channels (the library's only held workspace; every other
 * entry point takes caller-allocated scratch as explicit arguments). */
int p2p_workspace_bytes(int N, int C, long long* ws_bytes, long long* cnt_bytes);
/* stats [N,C,2] = (mean, rstd) of y [N,HW,C] alone: the reduction half of p2p_instnorm_act_fwd, for p2p_instnorm_act_planes. */
int p2p_instnorm_stats(const void* y, float* stats, int N, long long HW, int C, float eps, int y_dtype, void* stream);
/* Allocation-free variant of p2p_instnorm_stats.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and result; the single-launch reduction's scratch (partial sums + group
 * counters) is a file-scope device array inside the library (520 KB on every GPU that loads it, held until exit) whose
 * device address is looked up once per device with cudaGetSymbolAddress.  The call makes no cudaMalloc / cudaFree /
 * allocating library call.  Limits: N*C <= 65536, N*ceil(C/128) <= 2048, C % 4 == 0, HW < 2e9, y and stats 16-byte
 * aligned; anything beyond returns P2P_ENOTSUP with a message (use p2p_instnorm_stats there).  All calls on a device share
 * the array: the used region is cleared on `stream` before each launch, and calls that OVERLAP on the GPU (different
 * streams, concurrent graph branches) race on it - order them on one stream. */
int p2p_instnorm_stats_static(const void* y, float* stats, int N, long long HW, int C, float eps, int y_dtype, void* stream);

/* AvgPool2d(3, stride 2, padding 1, count_include_pad=False), NHWC. OH=(H-1)/2+1, OW=(W-1)/2+1. */
int p2p_avgpool3s2_fwd(const void* x, void* y, int N, int H, int W, int C, int dtype, void* stream);
int p2p_avgpool3s2_bwd(const void* dy, void* dx, int N, int H, int W, int C, int dtype, void* stream);

/* elementwise: y = tanh(x); dx = dy * (1 - y^2) */
int p2p_tanh_fwd(const void* x, void* y, long long n, int dtype, void* stream);
int p2p_tanh_bwd(const void* y, const void* dy, void* dx, long long n, int dtype, void* stream);

/* if loss: loss[0] += weight * mean((x - target)^2); if dx: dx = gscale * weight * 2 (x - target) / n
 * (LSGAN, networks.py:100-110).  gscale_dev (device fp32 scalar, may be NULL) multiplies gscale: the upstream
 * gradient never has to visit the host. */
int p2p_loss_mse_const(const void* x, float target, float weight, float* loss, void* dx, float gscale,
                       const float* gscale_dev, long long n, int dtype, void* stream);
/* if loss: loss[0] += weight * mean(|a - b|); if da: da = gscale * weight * sign(a - b) / n  (feature matching) */
int p2p_loss_l1(const void* a, const void* b, float weight, float* loss, void* da, float gscale,
                const float* gscale_dev, long long n, int dtype, void* stream);

/* torch.optim.Adam step (no weight decay, no amsgrad) over a flat fp32 buffer; also refreshes the bf16 copy
 * used by the tensor-core kernels when w_bf16 != NULL.  step is the 1-based step count. */
int p2p_adam_step(float* w, const float* g, float* m, float* v, void* w_bf16, long long n, float lr, float beta1,
                  float beta2, float eps, int step, void* stream);

/* The same update with the step count and the learning rate resident on the device (so that an optimisation step
 * captured in a CUDA graph advances its own bias correction): *step_dev is incremented, then used; the gradient is
 * multiplied by grad_scale first (1 / world size after a summing all-reduce).  scratch2: 2 floats of device scratch. */
int p2p_adam_step_dev(float* w, const float* g, float* m, float* v, void* w_bf16, long long n, const float* lr_dev,
                      float beta1, float beta2, float eps, float grad_scale, int* step_dev, float* scratch2, void* stream);

/* ---- spectrogram encode / decode around the networks (models/pix2pixHD_model.py:142-249) -----------------
 * The image the networks see is channels-last [B, M bins (H), F frames (W), C], C = 2 with explicit +/- encoding
 * (to_spectro :149-158: dB of alpha*pos+(1-alpha)*neg and of (1-alpha)*pos+alpha*neg), C = 1 otherwise (:160-162);
 * dB(x) = 20 log10(max(x, min_value)) - 20.
 * p2p_spectro_encode: spec [B,F,M] (MDCT4 output) -> logs [B,M,F,C] un-normalised, pha [B,M,F] = sign(spec), and the
 *   batch-global min / max / sum / sum of squares accumulated in stats_work (32 bytes of device scratch).
 * p2p_spectro_normalize: logs = (logs - min) / (max - min) in place (:193); when use_mask, rows [M-mask_rows, M) are
 *   replaced by noise [B,mask_rows,F,C] (or zeros when noise == NULL; :196-226); norm_out[4] = max, min, mean, std. */
int p2p_spectro_encode(const float* spec, float* logs, float* pha, void* stats_work, int B, int F, int M, int channels,
                       float alpha, float min_value, void* stream);
int p2p_spectro_normalize(float* logs, const void* stats_work, float* norm_out, const float* noise, int B, int F, int M,
                          int channels, int mask_rows, int use_mask, void* stream);
/* denormalize + explicit decode + transpose (:229-249): x is a [B,C,M,F] image with element strides (sb,sc,sh,sw) and
 * dtype `dtype`; coef[b,f,k] = gain * decode(x[b,:,k,f]); norm = {max, min} device scalars; pha [B,M,F] or NULL
 * multiplies the magnitude when channels == 1. */
int p2p_spectro_decode(const void* x, long long sb, long long sc, long long sh, long long sw, const float* norm,
                       const float* pha, float* coef, int B, int channels, int F, int M, float alpha, float min_value,
                       float gain, int dtype, void* stream);

/* ---- time-domain side of the loss path (models/pix2pixHD_model.py:251-258, 314-320, 375-387, 408-415) ------------
 * p2p_frame_window: frames[b,f,n] = window[n] * x_pad[b, f*hop + n], x_pad = audio shifted right by `offset` zeros and
 *   zero-extended -- the windowed frames MDCT2.forward(signal, return_ola=True) returns (models/mdct.py:497-503), which
 *   to_spectro stores in norm_param['frames'] (pix2pixHD_model.py:143-145, 227).
 * p2p_frames_db_fwd / _bwd: out = 20 log10(max(|gain * window[n] * frames[r,n]|, min_value)) - 20 over `rows` frames of
 *   N samples (discriminate_time_D :317-318 fused with sqrt(up_ratio-1) * window * frames of :375); window may be NULL
 *   (then only the gain applies).  _bwd: dframes = dout * d out / d frames (zero below the clamp).
 * p2p_frame_match (:408-415): weight * mean_{b, f < F-1, i < N/2}((y[b,f,N/2+i] w[i] - y[b,f+1,i] w[N/2+i])^2).  Value
 *   mode: loss != NULL (atomically accumulated: zero it first); gradient mode: dframes != NULL receives
 *   gscale * (*gscale_dev if non-NULL) * d loss / d y.  Both may be requested in one call.
 * p2p_spectro_decode_bwd: gradient of p2p_spectro_decode with channels == 2 (to_frames / to_audio in the loss path):
 *   dcoef [B,F,M] -> dx [B,M,F,2] fp32 channels-last; x, strides, norm, alpha, gain as in the forward call. */
int p2p_frame_window(const float* audio, const float* window, float* frames, int B, long long T, int F, int hop, int win,
                     long long offset, void* stream);
int p2p_frames_db_fwd(const float* frames, const float* window, float gain, float* out, long long rows, int N,
                      float min_value, void* stream);
int p2p_frames_db_bwd(const float* frames, const float* window, float gain, const float* dout, float* dframes,
                      long long rows, int N, float min_value, void* stream);
int p2p_frame_match(const float* frames, const float* window, float weight, float* loss, float* dframes, float gscale,
                    const float* gscale_dev, int B, int F, int N, void* stream);
int p2p_spectro_decode_bwd(const void* x, long long sb, long long sc, long long sh, long long sw, const float* norm,
                           const float* dcoef, float* dx, int B, int F, int M, float alpha, float gain, int dtype,
                           void* stream);

/* ---- MDCT2 / IMDCT2: the DCT-II lapped transform (models/mdct.py:352-454 with dct/dct_native.py:7-67) -------------
 *   DCT_2N_native(u)[k]  = (2/N) sum_n u[n] cos(pi (2n+1) k / (2N));   IDCT_2N_native(X)[n] = X[0] + 2 sum_{k>=1} X[k] cos(.)
 * p2p_mdct2_analysis:  spec[b,f,k] = scale * c0_k * DCT_2N(window * x_pad[b, f*hop : f*hop+win], zero-extended to n_fft)[k]
 * p2p_mdct2_synthesis: out[b, s - crop_left] = scale * sum_f window[n] * (2 sum_k c0_k spec[b,f,k] cos(pi (2n+1) k / (2 n_fft))),
 *                      n = s - f*hop in [0, win), s in [crop_left, crop_left + T_out)
 * with c0_0 = c0 and c0_k = 1 for k > 0.  MDCT2.forward = analysis(scale 1, c0 1); IMDCT2.forward = synthesis(scale 1/2,
 * c0 1/2); their autograd adjoints are synthesis(1/n_fft, 1) and analysis(n_fft/2, 1/2).  x_pad is x shifted right by
 * `offset` zeros and zero-extended.  x [B,T], spec [B,F,n_fft], window [win], all fp32 device pointers; tables: device
 * copy of p2p_mdct2_fill_tables(n_fft) (p2p_mdct2_table_words(n_fft) 4-byte words, computed in double on the host).
 * n_fft: power of two in [4, 4096]; 1 <= hop <= win <= n_fft.  win == n_fft, hop == n_fft/2, n_fft in 64..1024 runs the
 * register-FFT kernels (two frames per complex FFT, TMA-prefetched spans); p2p_mdct2_force_general(1) (a testing knob,
 * returns the old value) forces the general shared-memory kernels there too. */
int p2p_mdct2_force_general(int on);
int p2p_mdct2_table_words(int n_fft);
int p2p_mdct2_fill_tables(int n_fft, void* host_tables);
int p2p_mdct2_analysis(const float* x, float* spec, const float* window, const void* tables, int B, int64_t T, int F,
                       int n_fft, int hop, int win, int64_t offset, float scale, float c0, void* stream);
int p2p_mdct2_synthesis(const float* spec, float* out, const float* window, const void* tables, int B, int F, int64_t T_out,
                        int n_fft, int hop, int win, int64_t crop_left, float scale, float c0, void* stream);

/* ---- device-side resampling for the data path (data/audio_dataset.py:56-58 -> torchaudio.functional.resample) ---------
 * out[b, i*new_freq + j] = sum_{k < 2*width+orig} bank[j, k] * x_pad[b, i*orig + k], x_pad = x shifted right by `width`
 * zeros and zero-extended; orig / new_freq already divided by their gcd; bank [new_freq, 2*width+orig] fp32 on the
 * device (the windowed-sinc filter bank of torchaudio's _get_sinc_resample_kernel, built by the caller);
 * T_out <= ceil(new_freq * T / orig). */
int p2p_resample(const float* x, const float* bank, float* out, int B, long long T, long long T_out, int orig, int new_freq,
                 int width, void* stream);

#ifdef __cplusplus
}
#endif
#endif
