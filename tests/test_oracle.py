"""CPU: the oracle restatement against the reference's own outputs (golden fixtures made by importing
/root/reference, tests/golden/make_golden.py), the README known answers and an independent closed form."""
import numpy as np
import pytest
import torch

import fast_model as FM
from cases import CASES, relerr
from oracle import mdct4_oracle as O


def _geom(golden, name):
    n_fft, hop, win, center, _, one_d = CASES[name]
    x, w = golden[name + "/x"], golden[name + "/window"]
    sp, ep = O.pad_amounts(x.shape[0], hop, center)
    return n_fft, hop, win, center, x, w, sp, ep


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_forward_and_adjoints_match_reference(golden, name):
    n_fft, hop, win, center, x, w, sp, ep = _geom(golden, name)
    spec = golden[name + "/spec"]
    assert relerr(O.mdct4_forward(x, w, n_fft, hop, center), spec) < 1e-14
    gx = O.mdct4_backward(golden[name + "/gs"], w, n_fft, hop, x.shape[-1], sp)
    assert relerr(gx, golden[name + "/gx"]) < 5e-7          # reference grad comes back fp32
    s3 = spec if spec.ndim == 3 else spec[None]
    assert relerr(O.imdct4_forward(s3, w, n_fft, hop, center), golden[name + "/y"]) < 1e-14
    gs = O.imdct4_backward(golden[name + "/gy"][:, 0, 0, :], w, n_fft, hop, s3.shape[1], center)
    assert relerr(gs, golden[name + "/gspec"]) < 1e-14


@pytest.mark.parametrize("name", ["n64", "batch_quirk", "n256", "win_lt_nfft"])
def test_oracle_matches_closed_form(golden, name):
    n_fft, hop, win, center, x, w, sp, ep = _geom(golden, name)
    x2 = x if x.ndim == 2 else x[None]
    fr = O.frame_signal(x2, sp, ep, win, hop).astype(np.float64) * w.astype(np.float64)
    spec = O.mdct4_closed_form(fr, n_fft)
    assert relerr(spec, golden[name + "/spec"]) < 1e-6       # reference multiplies the window in fp32
    y = O.imdct4_closed_form(golden[name + "/spec"], n_fft)
    assert relerr(y, O.imdct4_frames(golden[name + "/spec"], n_fft)) < 1e-12


@pytest.mark.parametrize("name", list(CASES))
def test_fast_algorithm_model_matches_reference(golden, name):
    """The fold + M/2-point-FFT formulation the kernels implement equals the reference's N-point FFT."""
    n_fft, hop, win, center, x, w, sp, ep = _geom(golden, name)
    spec = golden[name + "/spec"]
    x2 = x if x.ndim == 2 else x[None]
    s3 = spec if spec.ndim == 3 else spec[None]
    F = s3.shape[1]
    crop = win // 2 if center else 0
    assert relerr(FM.analysis(x2, w, n_fft, hop, F, sp), s3) < 1e-6
    y = golden[name + "/y"][:, 0, 0, :]
    assert relerr(FM.synthesis(s3, w, n_fft, hop, y.shape[-1], crop, 4.0 / n_fft), y) < 1e-12
    gs = golden[name + "/gs"]
    gs = gs if gs.ndim == 3 else gs[None]
    gx = golden[name + "/gx"]
    assert relerr(FM.synthesis(gs, w, n_fft, hop, x.shape[-1], sp, 1.0), gx if gx.ndim == 2 else gx[None]) < 5e-7
    assert relerr(FM.analysis(golden[name + "/gy"][:, 0, 0, :], w, n_fft, hop, F, crop, 4.0 / n_fft),
                  golden[name + "/gspec"]) < 1e-12


def test_config1_known_answers(golden_config1):
    """README.md:112-118 / DCT_test.ipynb cell 32: [257, 512] frames x bins; round trip reconstructs."""
    c = golden_config1
    assert list(c["spec_shape"]) == [16, 257, 512]
    assert list(c["one_d_shape"]) == [257, 512]
    assert list(c["y_shape"]) == [16, 1, 1, 131072]
    torch.manual_seed(1234)
    x = (0.1 * torch.randn(16, 130816)).numpy()
    w = O.kbdwin(1024)
    assert np.abs(w - c["kbd1024"]).max() < 5e-7
    spec = O.mdct4_forward(x[::5], c["kbd1024"], 1024, 512, True, signal_len=16)
    assert spec.shape == (4, 257, 512)
    assert relerr(spec[:, ::37, :], c["spec_rows"]) < 1e-14
    y = O.imdct4_forward(spec, c["kbd1024"], 1024, 512, True)
    assert y.shape == (4, 1, 1, 131072)
    assert relerr(y[:, 0, 0, ::101], c["y_sample"][::5]) < 1e-13
    mse = np.mean((x[::5] - y[:, 0, 0, :130816]) ** 2)
    assert mse < 1e-12 and float(c["mse"]) < 1e-12


def test_double_window_round_trip_is_exact():
    """README.md:118 quotes 4.89e-32 with a float64 KBD window (notebook cell 30)."""
    rng = np.random.default_rng(0)
    x = 0.1 * rng.standard_normal((2, 130816))
    w = O.kbdwin(1024, dtype=np.float64)
    assert np.abs(w[:512] ** 2 + w[512:] ** 2 - 1).max() < 1e-14
    y = O.imdct4_forward(O.mdct4_forward(x, w, 1024, 512, True, signal_len=130816), w, 1024, 512, True)
    assert np.mean((x - y[:, 0, 0, :130816]) ** 2) < 1e-28


def test_kbdwin_princen_bradley():
    for n in (16, 256, 1024, 4096):
        w = O.kbdwin(n).astype(np.float64)
        assert np.abs(w[: n // 2] ** 2 + w[n // 2:] ** 2 - 1).max() < 1e-6


def test_frame_count_quirk():
    """end_pad comes from len(signal) = B for 2-D input."""
    assert O.pad_amounts(8, 8, True) == (8, 8)
    assert O.n_frames(100, 8, 8, 16, 8) == 13
    assert O.pad_amounts(100, 8, True) == (8, 12)
    assert O.n_frames(100, 8, 12, 16, 8) == 14
