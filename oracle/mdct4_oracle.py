"""CPU oracle for the MDCT4 / IMDCT4 hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this file.  The product path (pix2pixhdaudiosr_b200.*) never does and
fails loudly when the CUDA library is missing.

This is a float64 numpy restatement of the reference algorithm, step by step:
  * kbdwin            <- /root/reference/util/util.py:186-193
  * mdct4_forward     <- /root/reference/models/mdct.py:486-513 (twiddles 483-484)
  * imdct4_forward    <- /root/reference/models/mdct.py:542-566 (twiddles 539-540)
  * the two adjoints (what torch autograd yields through those lines)
  * mdct4_closed_form / imdct4_closed_form: independent O(N^2) cosine sums (App. A.1-2)

Third-party arithmetic under the reference path: torch.fft.fft (pocketfft on CPU),
torch.kaiser_window, F.pad, Tensor.unfold, F.fold from torch 2.11.0.  numpy.fft is the same
pocketfft algorithm.

Parity pinning: tests/golden/mdct4_*.npz were produced by importing the reference itself in the
build container (tests/golden/make_golden.py); tests/test_oracle.py checks this file against
them, against README.md:112-118 (shape [257,512], round-trip MSE) and the closed form.
"""
import numpy as np


def kaiser_window_symmetric(n, beta):
    """torch.kaiser_window(n, periodic=False, beta): I0(beta*sqrt(1-((2i/(n-1))-1)^2))/I0(beta)."""
    if n == 1:
        return np.ones(1)
    i = np.arange(n, dtype=np.float64)
    r = 2.0 * i / (n - 1) - 1.0
    return np.i0(beta * np.sqrt(np.clip(1.0 - r * r, 0.0, None))) / np.i0(beta)


def kbdwin(N, beta=12.0, dtype=np.float32):
    """util/util.py:186-193.  The reference computes in fp32 (torch default dtype); pass
    dtype=np.float64 for the notebook's double window (test/DCT_test.ipynb cell 30)."""
    assert N % 2 == 0, "N must be even"
    w = kaiser_window_symmetric(N // 2 + 1, beta * np.pi).astype(dtype)
    half = np.sqrt(np.cumsum(w, dtype=dtype) / w.sum(dtype=dtype))[:-1]
    return np.concatenate([half, half[::-1]]).astype(dtype)


def pad_amounts(signal_len, hop, center):
    """mdct.py:488-496.  signal_len is len(signal): T for 1-D input, the BATCH size for 2-D."""
    start_pad = hop if center else 0
    r = signal_len % hop
    end_pad = start_pad + (hop - r if r else 0)
    return start_pad, end_pad


def n_frames(T, start_pad, end_pad, win, hop):
    return (T + start_pad + end_pad - win) // hop + 1


def _twiddles_fwd(n_fft):
    exp1 = np.exp(-1j * np.pi / n_fft * np.arange(0, n_fft, dtype=np.float64))           # :483
    exp2 = np.exp(-1j * (np.pi / (2 * n_fft) + np.pi / 4) * np.arange(1, n_fft, 2, dtype=np.float64))  # :484
    return exp1, exp2


def frame_signal(x, start_pad, end_pad, win, hop):
    """pad (constant) + unfold(-1, win, hop): mdct.py:497-500."""
    x = np.asarray(x)
    pad = [(0, 0)] * (x.ndim - 1) + [(start_pad, end_pad)]
    xp = np.pad(x, pad)
    F = (xp.shape[-1] - win) // hop + 1
    idx = np.arange(F)[:, None] * hop + np.arange(win)[None, :]
    return xp[..., idx]


def mdct4_forward(x, window, n_fft, hop, center=True, signal_len=None):
    """mdct.py:486-513.  x [..., T] (fp32 as the reference multiplies by the window in fp32, :503),
    returns float64 [..., F, n_fft/2]."""
    x = np.asarray(x)
    win = len(window)
    if signal_len is None:
        signal_len = x.shape[0]  # len(signal), the reference quirk
    sp, ep = pad_amounts(signal_len, hop, center)
    fr = frame_signal(x, sp, ep, win, hop)
    fr = (fr.astype(np.float32) * np.asarray(window, np.float32)).astype(np.float64) \
        if x.dtype == np.float32 and np.asarray(window).dtype == np.float32 \
        else fr.astype(np.float64) * np.asarray(window, np.float64)
    if n_fft > win:
        fr = np.pad(fr, [(0, 0)] * (fr.ndim - 1) + [(0, n_fft - win)])
    exp1, exp2 = _twiddles_fwd(n_fft)
    s = np.fft.fft(fr * exp1, axis=-1)[..., : n_fft // 2]
    return np.real(exp2 * s)


def imdct4_frames(spec, n_fft):
    """mdct.py:547-549: y[f, n] for n < n_fft (before truncation/window)."""
    spec = np.asarray(spec, np.float64)
    exp1 = np.exp(-1j * (np.pi / (2 * n_fft) + np.pi / 4) * np.arange(1, n_fft, 2, dtype=np.float64))  # :539
    exp2 = np.exp(-1j * np.pi / (2 * n_fft) * np.arange(0, 2 * n_fft, 2, dtype=np.float64))             # :540
    s = np.fft.fft(exp1 * spec, n=n_fft, axis=-1)
    return np.real(s * exp2)


def overlap_add(frames, hop):
    """F.fold(kernel=(1,win), stride=(1,hop)) on [..., F, win] -> [..., (F-1)*hop+win] (mdct.py:559-560)."""
    F, win = frames.shape[-2:]
    out = np.zeros(frames.shape[:-2] + ((F - 1) * hop + win,), np.float64)
    for f in range(F):
        out[..., f * hop: f * hop + win] += frames[..., f, :]
    return out


def imdct4_forward(spec, window, n_fft, hop, center=True, out_length=None):
    """mdct.py:542-566.  spec [B, F, n_fft/2] -> float64 [B, 1, 1, T_out]."""
    spec = np.asarray(spec)
    assert spec.ndim == 3 and spec.shape[-1] == n_fft // 2
    win = len(window)
    y = imdct4_frames(spec, n_fft)[..., :win] * np.asarray(window, np.float64)
    out = 4.0 / n_fft * overlap_add(y, hop)
    if center:
        out = out[..., win // 2: -win // 2]   # same Python slice expression as mdct.py:564
    out = out[:, None, None, :]
    return out if out_length is None else out[..., :out_length]


# ---- adjoints (autograd of the two forwards) ---------------------------

def mdct4_backward(g, window, n_fft, hop, T, start_pad):
    """dL/dx for MDCT4 given g = dL/dX [..., F, n_fft/2]: crop(OLA(w * C^T g)), no 4/N scale."""
    win = len(window)
    y = imdct4_frames(g, n_fft)[..., :win] * np.asarray(window, np.float64)
    ola = overlap_add(y, hop)
    out = ola[..., start_pad: start_pad + T]
    if out.shape[-1] < T:   # frames did not cover the tail (dropped partial frame): zero gradient there
        out = np.pad(out, [(0, 0)] * (out.ndim - 1) + [(0, T - out.shape[-1])])
    return out


def imdct4_backward(g, window, n_fft, hop, F, center=True):
    """dL/dX for IMDCT4 given g = dL/dout [B, T_out]: (4/N) * C * (w * frame(zero-re-pad(g)))."""
    win = len(window)
    g = np.asarray(g, np.float64)
    full = (F - 1) * hop + win
    left = win // 2 if center else 0
    buf = np.zeros(g.shape[:-1] + (full,), np.float64)
    n = min(g.shape[-1], full - left)
    buf[..., left: left + n] = g[..., :n]
    idx = np.arange(F)[:, None] * hop + np.arange(win)[None, :]
    fr = buf[..., idx] * np.asarray(window, np.float64)
    if n_fft > win:
        fr = np.pad(fr, [(0, 0)] * (fr.ndim - 1) + [(0, n_fft - win)])
    exp1, exp2 = _twiddles_fwd(n_fft)
    s = np.fft.fft(fr * exp1, axis=-1)[..., : n_fft // 2]
    return 4.0 / n_fft * np.real(exp2 * s)


# ---- independent closed forms -----------------------------------------------------------------

def cos_matrix(n_fft):
    M = n_fft // 2
    n = np.arange(n_fft, dtype=np.float64)[None, :]
    k = np.arange(M, dtype=np.float64)[:, None]
    return np.cos(np.pi / M * (n + 0.5 + M / 2) * (k + 0.5))     # [M, N]


def mdct4_closed_form(frames_windowed, n_fft):
    """X[k] = sum_n u[n] cos(pi/M (n+1/2+M/2)(k+1/2)); frames zero-extended to n_fft."""
    u = np.asarray(frames_windowed, np.float64)
    if u.shape[-1] < n_fft:
        u = np.pad(u, [(0, 0)] * (u.ndim - 1) + [(0, n_fft - u.shape[-1])])
    return u @ cos_matrix(n_fft).T


def imdct4_closed_form(spec, n_fft):
    return np.asarray(spec, np.float64) @ cos_matrix(n_fft)
