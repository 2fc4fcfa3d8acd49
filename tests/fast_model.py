"""numpy model of the algorithm the CUDA kernels implement (fold -> DCT-IV by an M/2-point complex
FFT -> unfold), used by tests to check the kernel maths on CPU and to document index formulas.
Not an oracle: the oracle is oracle/mdct4_oracle.py (the reference's N-point-FFT algorithm)."""
import numpy as np


def tables(n_fft):
    """pre[n] = e^{-j pi (n+1/4)/M}, post[k] = e^{-j pi k/M}, n,k < L = M/2."""
    M = n_fft // 2
    L = M // 2
    n = np.arange(L, dtype=np.float64)
    return np.exp(-1j * np.pi * (n + 0.25) / M), np.exp(-1j * np.pi * n / M)


def dct4(v):
    """DCT-IV (unnormalised) along the last axis, length M, via one M/2-point complex FFT."""
    M = v.shape[-1]
    L = M // 2
    pre, post = tables(2 * M)
    n = np.arange(L)
    z = (v[..., 2 * n] + 1j * v[..., M - 1 - 2 * n]) * pre
    Y = np.fft.fft(z, axis=-1) * post
    X = np.empty(v.shape, np.float64)
    X[..., 2 * n] = Y.real
    X[..., M - 1 - 2 * n] = -Y.imag
    return X


def fold(u):
    """windowed frame u [..., N] -> v [..., M] with MDCT(u) = DCT4(v) (App. A.3)."""
    N = u.shape[-1]
    M = N // 2
    i = np.arange(M // 2)
    v = np.empty(u.shape[:-1] + (M,), np.float64)
    v[..., i] = -u[..., 3 * M // 2 - 1 - i] - u[..., 3 * M // 2 + i]
    v[..., M // 2 + i] = u[..., i] - u[..., M - 1 - i]
    return v


def unfold(t):
    """t = DCT4(X) [..., M] -> y [..., N] = C^T X (App. A.4)."""
    M = t.shape[-1]
    i = np.arange(M // 2)
    y = np.empty(t.shape[:-1] + (2 * M,), np.float64)
    y[..., 3 * M // 2 - 1 - i] = -t[..., i]
    y[..., 3 * M // 2 + i] = -t[..., i]
    y[..., i] = t[..., M // 2 + i]
    y[..., M - 1 - i] = -t[..., M // 2 + i]
    return y


def analysis(x, window, n_fft, hop, F, offset, scale=1.0):
    """The analysis kernel's contract: out[b,f,:] = scale * C (w . x[b, f*hop - offset + n]),
    reads outside [0,T) are zero."""
    x = np.asarray(x, np.float64)
    B, T = x.shape
    win = len(window)
    idx = np.arange(F)[:, None] * hop + np.arange(win)[None, :] - offset
    ok = (idx >= 0) & (idx < T)
    fr = np.where(ok, x[:, np.clip(idx, 0, T - 1)], 0.0) * np.asarray(window, np.float64)
    u = np.zeros((B, F, n_fft))
    u[..., :win] = fr
    return scale * dct4(fold(u))


def synthesis(spec, window, n_fft, hop, T_out, crop_left, scale):
    """The synthesis kernel's contract: out[b,s] = scale * sum_f w[n] y_f[n], n = s + crop_left - f*hop."""
    spec = np.asarray(spec, np.float64)
    B, F, M = spec.shape
    win = len(window)
    y = unfold(dct4(spec))[..., :win] * np.asarray(window, np.float64)
    full = np.zeros((B, max((F - 1) * hop + win, crop_left + T_out)))
    for f in range(F):
        full[:, f * hop: f * hop + win] += y[:, f]
    return scale * full[:, crop_left: crop_left + T_out]
