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
