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

#ifdef __cplusplus
}
#endif
#endif
