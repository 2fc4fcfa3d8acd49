"""Geometry of the golden cases in tests/golden/mdct4_ref.npz (see make_golden.py)."""
# name: (n_fft, hop, win, center, window is kbd?, 1-D input?)
CASES = {
    "c1_small": (1024, 512, 1024, True, True, False),
    "n256": (256, 128, 256, True, True, False),
    "n512_nocenter": (512, 256, 512, False, True, False),
    "n2048": (2048, 1024, 2048, True, True, True),
    "n4096": (4096, 2048, 4096, True, True, False),
    "hop_quarter": (512, 128, 512, True, True, False),
    "win_lt_nfft": (512, 192, 384, True, False, False),
    "batch_quirk": (16, 8, 16, True, True, False),
    "n64": (64, 32, 64, True, True, False),
}


def relerr(a, b):
    import numpy as np
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
