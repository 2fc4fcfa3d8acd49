"""CPU oracle for MDCT2 / IMDCT2 and DCT_2N_native / IDCT_2N_native.  TEST INFRASTRUCTURE ONLY: only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may import this file.

float64 numpy restatement of the reference, step by step:
  * dct_2n            <- /root/reference/dct/dct_native.py:16-33   (zero-extend to 2N, rfft, first N bins, * 1/N,
                                                                    * [2cos, 2sin](pi k / 2N) and add: discrete_spectral_transform.py:28-41)
  * idct_2n           <- /root/reference/dct/dct_native.py:44-67   (X * 2e^{j pi k/2N}, zero-extend to 2N, irfft of
                                                                    its first N+1 bins, first N samples, * N)
  * mdct2_forward     <- /root/reference/models/mdct.py:377-398    (pad, unfold, window, zero-extend, DCT)
  * imdct2_forward    <- /root/reference/models/mdct.py:432-454    (IDCT / 2, truncate, window, fold, crop)
  * dct_2n_closed_form / idct_2n_closed_form: independent O(N^2) cosine sums.

Parity pinning: tests/golden/mdct2_ref.npz was produced by importing the reference itself in the build container
(tests/golden/make_golden_mdct2.py; outputs AND autograd gradients); tests/test_oracle_mdct2.py checks this file against
it and against the known answers of test/DCT_test.ipynb cells 34-35 (DCT_2N_native(arange(1,17)), IDCT(DCT(x)) = 2x).
"""
import numpy as np

from .mdct4_oracle import frame_signal, overlap_add, pad_amounts


def dct_2n(x):
    x = np.asarray(x, np.float64)
    N = x.shape[-1]
    xp = np.concatenate([x, np.zeros_like(x)], axis=-1)                       # dct_native.py:21
    y = np.fft.rfft(xp, axis=-1)[..., :N] / N                                 # :23-24
    a = np.arange(N, dtype=np.float64) * (np.pi / (2 * N))
    return y.real * (2 * np.cos(a)) + y.imag * (2 * np.sin(a))                # :30-33 (expk = [2cos, 2sin])


def idct_2n(X):
    X = np.asarray(X, np.float64)
    N = X.shape[-1]
    a = np.arange(N, dtype=np.float64) * (np.pi / (2 * N))
    z = X * (2 * np.cos(a) + 2j * np.sin(a))                                  # :52
    zp = np.concatenate([z, np.zeros_like(z)], axis=-1)                       # :54
    # torch_fft_api.irfft(onesided=False, signal_sizes=[2N]) calls torch.fft.irfft(z, 2N) (torch_fft_api.py:85): bins
    # 0..N of the 2N-long row are read as a half spectrum (bin N is padding, i.e. zero)
    return np.fft.irfft(zp[..., :N + 1], n=2 * N, axis=-1)[..., :N] * N       # :60-61


def _cos(N):
    n = np.arange(N, dtype=np.float64)[None, :]
    k = np.arange(N, dtype=np.float64)[:, None]
    return np.cos(np.pi * (2 * n + 1) * k / (2 * N))                          # [k, n]


def dct_2n_closed_form(x):
    x = np.asarray(x, np.float64)
    return (2.0 / x.shape[-1]) * x @ _cos(x.shape[-1]).T


def idct_2n_closed_form(X):
    X = np.asarray(X, np.float64)
    return 2.0 * X @ _cos(X.shape[-1]) - X[..., :1]


def mdct2_forward(x, window, n_fft, hop, center=True, signal_len=None, return_ola=False):
    """mdct.py:377-398.  x [..., T] -> float64 [..., F, n_fft] (and the windowed frames [..., F, win])"""
    x = np.asarray(x)
    win = len(window)
    if signal_len is None:
        signal_len = x.shape[0]                                               # len(signal), SURVEY H8a
    sp, ep = pad_amounts(signal_len, hop, center)
    fr = frame_signal(x, sp, ep, win, hop).astype(np.float64) * np.asarray(window, np.float64)
    ola = fr
    if n_fft > win:
        fr = np.pad(fr, [(0, 0)] * (fr.ndim - 1) + [(0, n_fft - win)])
    X = dct_2n(fr)
    return (X, ola) if return_ola else X


def imdct2_forward(spec, window, n_fft, hop, center=True, out_length=None):
    """mdct.py:432-454.  spec [B, F, n_fft] -> float64 [B, 1, 1, T_out]"""
    spec = np.asarray(spec)
    assert spec.ndim == 3 and spec.shape[-1] == n_fft
    win = len(window)
    y = (idct_2n(spec) / 2.0)[..., :win] * np.asarray(window, np.float64)
    out = overlap_add(y, hop)
    if center:
        out = out[..., win // 2: -win // 2]
    out = out[:, None, None, :]
    return out if out_length is None else out[..., :out_length]
