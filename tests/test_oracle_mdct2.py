"""Pins oracle/mdct2_oracle.py to the reference's own MDCT2 / IMDCT2 / DCT_2N_native / IDCT_2N_native outputs
(tests/golden/mdct2_ref.npz, made by importing /root/reference: tests/golden/make_golden_mdct2.py) and to the known
answers of test/DCT_test.ipynb cells 34-35."""
import os

import numpy as np
import pytest

from cases import relerr
from oracle import mdct2_oracle as O2
from oracle import mdct4_oracle as O4

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# name: (n_fft, hop, win, center, kbd)  -- geometry of make_golden_mdct2.CASES
CASES2 = {
    "model": (128, 64, 128, True, True), "recipe": (512, 256, 512, True, True), "nocenter": (256, 128, 256, False, True),
    "hop_quarter": (256, 64, 256, True, True), "win_lt_nfft": (256, 96, 192, True, False), "n1024": (1024, 512, 1024, True, True),
    "n16": (16, 8, 16, True, True),
}


@pytest.fixture(scope="module")
def ref():
    return np.load(os.path.join(ROOT, "tests", "golden", "mdct2_ref.npz"))


def window_of(case):
    n_fft, hop, win, center, kbd = CASES2[case]
    return O4.kbdwin(win) if kbd else np.ones(win, np.float32)


def test_known_answers(ref):
    x = np.arange(1, 17, dtype=np.float64)
    assert relerr(O2.dct_2n(x), ref["known/dct_arange"]) < 1e-13                 # DCT_test.ipynb cell 34
    assert np.allclose(O2.idct_2n(O2.dct_2n(x)), 2 * x, atol=1e-12)               # cell 35: IDCT(DCT(x)) = 2x
    assert np.allclose(ref["known/idct_dct_arange"], 2 * x, atol=1e-12)
    assert abs(O2.dct_2n(x)[0] - 17.0) < 1e-12                                    # (2/N) * sum = 2 * mean


@pytest.mark.parametrize("N", [8, 64, 512])
def test_dct_pair_matches_reference_and_closed_form(ref, N):
    x, X = ref[f"dct{N}/x"], ref[f"dct{N}/X"]
    assert relerr(O2.dct_2n(x), ref[f"dct{N}/dct"]) < 1e-12
    assert relerr(O2.idct_2n(X), ref[f"dct{N}/idct"]) < 1e-12
    assert relerr(O2.dct_2n_closed_form(x), ref[f"dct{N}/dct"]) < 1e-11
    assert relerr(O2.idct_2n_closed_form(X), ref[f"dct{N}/idct"]) < 1e-11
    # adjoints (what the CUDA autograd pair implements): D^T g = (idct(g) + g[0]) / N ; E^T g = N * dct(g), bin 0 halved
    g = ref[f"dct{N}/g"]
    assert relerr((O2.idct_2n(g) + g[..., :1]) / N, ref[f"dct{N}/dct_gx"]) < 1e-11
    e = N * O2.dct_2n(g)
    e[..., 0] *= 0.5
    assert relerr(e, ref[f"dct{N}/idct_gX"]) < 1e-11


@pytest.mark.parametrize("case", list(CASES2))
def test_mdct2_matches_reference(ref, case):
    n_fft, hop, win, center, kbd = CASES2[case]
    w = window_of(case)
    x = ref[case + "/x"]
    spec, frames = O2.mdct2_forward(x, w, n_fft, hop, center, return_ola=True)
    assert spec.shape == ref[case + "/spec"].shape
    assert relerr(spec, ref[case + "/spec"]) < 2e-6                # the reference transforms in fp32 here
    assert relerr(frames, ref[case + "/frames"]) < 1e-6
    y = O2.imdct2_forward(ref[case + "/spec"], w, n_fft, hop, center)
    assert y.shape == ref[case + "/y"].shape
    assert relerr(y, ref[case + "/y"]) < 2e-6


# ---- host side of csrc/mdct2.cu: the tables the C ABI fills, and the kernels' index formulas restated in numpy -----
def _tables(N):
    import ctypes
    from pix2pixhdaudiosr_b200 import _lib
    lib = _lib.lib()
    words = lib.p2p_mdct2_table_words(N)
    assert words == 6 * N
    buf = np.zeros(words, np.int32)
    assert lib.p2p_mdct2_fill_tables(N, ctypes.c_void_p(buf.ctypes.data)) == 0
    f = buf.view(np.float32)
    tw = f[0:2 * N:2] + 1j * f[1:2 * N:2]
    roots = f[2 * N:4 * N:2] + 1j * f[2 * N + 1:4 * N:2]
    return tw.astype(np.complex128), roots.astype(np.complex128), buf[4 * N:5 * N].copy(), buf[5 * N:6 * N].copy()


def _dif_fft(z, roots):
    """the in-place radix-4 (+ final radix-2) DIF passes of csrc/smem_fft.cuh: natural-order in, permuted out"""
    z = z.copy()
    L = len(z)
    S = L
    while S >= 4:
        q = S // 4
        step = L // S
        for blk in range(L // S):
            for j in range(q):
                p = blk * S + j
                a0, a1, a2, a3 = z[p], z[p + q], z[p + 2 * q], z[p + 3 * q]
                s02, d02, s13, d13 = a0 + a2, a0 - a2, a1 + a3, a1 - a3
                z[p] = s02 + s13
                z[p + q] = (d02 - 1j * d13) * roots[step * j]
                z[p + 2 * q] = (s02 - s13) * roots[step * 2 * j]
                z[p + 3 * q] = (d02 + 1j * d13) * roots[step * 3 * j]
        S = q
    if S == 2:
        for r in range(L // 2):
            a0, a1 = z[2 * r], z[2 * r + 1]
            z[2 * r], z[2 * r + 1] = a0 + a1, a0 - a1
    return z


@pytest.mark.parametrize("N", [4, 8, 64, 512])
def test_kernel_algorithm_with_the_library_tables(N):
    """paired-frame Makhoul DCT-II / its inverse exactly as mdct2_analysis_kernel / mdct2_synthesis_kernel index them,
    on the tables p2p_mdct2_fill_tables produces (host-only call: no GPU needed)"""
    tw, roots, perm, inv = _tables(N)
    assert sorted(perm) == list(range(N)) and all(inv[perm[p]] == p for p in range(N))
    assert np.abs(tw - np.exp(-1j * np.pi * np.arange(N) / (2 * N))).max() < 1e-7
    rng = np.random.default_rng(N)
    uA, uB = rng.standard_normal(N), rng.standard_normal(N)
    idx = np.array([2 * p if p < N // 2 else 2 * (N - 1 - p) + 1 for p in range(N)])
    Z = _dif_fft(uA[idx] + 1j * uB[idx], roots)
    assert np.abs(Z[inv] - np.fft.fft(uA[idx] + 1j * uB[idx])).max() < 1e-5 * N        # position inv[k] holds bin k
    k = np.arange(N)
    a, m = Z[inv[k]], Z[inv[(N - k) % N]]
    va, vb = a + np.conj(m), -1j * (a - np.conj(m))
    XA, XB = (va * tw).real / N, (vb * tw).real / N
    assert relerr(XA, O2.dct_2n(uA)) < 1e-5 and relerr(XB, O2.dct_2n(uB)) < 1e-5
    # inverse: h[k] = tw[k] (X[k] + j X[N-k]) / 2, h[0] = c0 X[0]; FFT(hA + j hB) = vA + j vB; y[idx(n)] = 2 v[n]
    def h(X, c0):
        out = 0.5 * tw * (X + 1j * X[(N - k) % N])
        out[0] = c0 * X[0]
        return out
    V = _dif_fft(h(XA, 0.5) + 1j * h(XB, 0.5), roots)
    yA, yB = np.zeros(N), np.zeros(N)
    for p in range(N):
        n = perm[p]
        yA[idx[n]], yB[idx[n]] = 2 * V[p].real, 2 * V[p].imag
    assert relerr(yA, O2.idct_2n(XA)) < 1e-5 and relerr(yB, O2.idct_2n(XB)) < 1e-5
    assert relerr(yA, 2 * uA) < 1e-5                                                    # IDCT(DCT(u)) = 2u
