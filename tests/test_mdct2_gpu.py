"""GPU parity of MDCT2 / IMDCT2 / DCT_2N_native / IDCT_2N_native through the C ABI: against the
reference's own outputs and autograd gradients (tests/golden/mdct2_ref.npz), against the float64 oracle on seeded
inputs, and through size-independent properties (perfect reconstruction, IDCT(DCT(x)) = 2x, adjointness) at full size.
Tolerance: 1e-5 relative (fp32 kernels vs the float64 oracle), as for MDCT4."""
import os

import numpy as np
import pytest
import torch

from cases import relerr
from oracle import mdct2_oracle as O2
from pix2pixhdaudiosr_b200.dct.dct_native import DCT_2N_native, IDCT_2N_native
from pix2pixhdaudiosr_b200.models.mdct import IMDCT2, MDCT2
from pix2pixhdaudiosr_b200.util.util import kbdwin
from test_oracle_mdct2 import CASES2

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-5


@pytest.fixture(scope="module")
def ref():
    return np.load(os.path.join(ROOT, "tests", "golden", "mdct2_ref.npz"))


def modules(case):
    n_fft, hop, win, center, kbd = CASES2[case]
    window = kbdwin if kbd else None
    return (MDCT2(n_fft, hop, win, window, center, device="cuda", dct_op=DCT_2N_native()),
            IMDCT2(n_fft, hop, win, window, center, device="cuda", idct_op=IDCT_2N_native()))


@pytest.mark.parametrize("case", list(CASES2))
def test_forward_and_autograd_match_reference(ref, case):
    m, im = modules(case)
    x = torch.from_numpy(ref[case + "/x"]).cuda().requires_grad_(True)
    spec, frames = m(x, True)
    assert tuple(spec.shape) == ref[case + "/spec"].shape and tuple(frames.shape) == ref[case + "/frames"].shape
    assert relerr(spec.detach().cpu().numpy(), ref[case + "/spec"]) < TOL
    assert relerr(frames.cpu().numpy(), ref[case + "/frames"]) < 1e-6
    (gx,) = torch.autograd.grad(spec, x, torch.from_numpy(ref[case + "/gs"]).cuda())
    assert relerr(gx.cpu().numpy(), ref[case + "/gx"]) < TOL
    spec_in = torch.from_numpy(ref[case + "/spec"]).cuda().requires_grad_(True)
    y = im(spec_in)
    assert tuple(y.shape) == ref[case + "/y"].shape
    assert relerr(y.detach().cpu().numpy(), ref[case + "/y"]) < TOL
    (gspec,) = torch.autograd.grad(y, spec_in, torch.from_numpy(ref[case + "/gy"]).cuda())
    assert relerr(gspec.cpu().numpy(), ref[case + "/gspec"]) < TOL


@pytest.mark.parametrize("N", [8, 64, 512])
def test_dct_modules_match_reference(ref, N):
    x = torch.from_numpy(ref[f"dct{N}/x"]).float().cuda().requires_grad_(True)
    X = torch.from_numpy(ref[f"dct{N}/X"]).float().cuda().requires_grad_(True)
    g = torch.from_numpy(ref[f"dct{N}/g"]).float().cuda()
    y = DCT_2N_native()(x)
    yi = IDCT_2N_native()(X)
    assert relerr(y.detach().cpu().numpy(), ref[f"dct{N}/dct"]) < TOL
    assert relerr(yi.detach().cpu().numpy(), ref[f"dct{N}/idct"]) < TOL
    (gx,) = torch.autograd.grad(y, x, g)
    (gX,) = torch.autograd.grad(yi, X, g)
    assert relerr(gx.cpu().numpy(), ref[f"dct{N}/dct_gx"]) < TOL
    assert relerr(gX.cpu().numpy(), ref[f"dct{N}/idct_gX"]) < TOL


def test_known_answers(ref):
    x = torch.arange(1, 17, dtype=torch.float32).cuda()
    d = DCT_2N_native()(x)
    assert relerr(d.cpu().numpy(), ref["known/dct_arange"]) < TOL                      # DCT_test.ipynb cell 34
    assert relerr(IDCT_2N_native()(d).cpu().numpy(), 2 * np.arange(1, 17)) < TOL       # cell 35


@pytest.mark.parametrize("n_fft,hop,win,B,T", [(2048, 1024, 2048, 2, 20000), (4096, 2048, 4096, 1, 30000), (512, 512, 512, 3, 5000),
                                               (64, 16, 48, 5, 777), (4, 2, 4, 2, 50)])
def test_against_oracle_on_seeded_inputs(n_fft, hop, win, B, T):
    torch.manual_seed(n_fft + hop)
    x = 0.1 * torch.randn(B, T)
    w = kbdwin(win).float() if win % 2 == 0 else torch.ones(win)
    m = MDCT2(n_fft, hop, win, w, True, device="cuda")
    im = IMDCT2(n_fft, hop, win, w, True, device="cuda")
    spec = m(x.cuda())
    want = O2.mdct2_forward(x.numpy(), w.numpy(), n_fft, hop, True)
    assert tuple(spec.shape) == want.shape
    assert relerr(spec.cpu().numpy(), want) < TOL
    y = im(spec)
    ywant = O2.imdct2_forward(want, w.numpy(), n_fft, hop, True)
    assert tuple(y.shape) == ywant.shape
    assert relerr(y.cpu().numpy(), ywant) < TOL


def test_full_size_properties():
    """config-1-sized batch (16 x 130 816, the recipe's n_fft 512 / hop 256): perfect reconstruction with the KBD window,
    linearity, and <MDCT2 x, g> = <x, MDCT2^T g> (the adjoint the backward kernel implements)"""
    torch.manual_seed(1234)
    B, T, n_fft, hop = 16, 130816, 512, 256
    m = MDCT2(n_fft, hop, n_fft, kbdwin, True, device="cuda")
    im = IMDCT2(n_fft, hop, n_fft, kbdwin, True, device="cuda")
    x = (0.1 * torch.randn(B, T)).cuda().requires_grad_(True)
    spec = m(x)
    assert tuple(spec.shape) == (B, 512, 512)           # len(signal) = B = 16 drives end_pad (mdct.py:379-386)
    y = im(spec)[:, 0, 0, :T]
    assert float((y - x).abs().max()) < 1e-5 * float(x.abs().max()) * 10
    assert float(((y - x) ** 2).mean()) < 1e-12
    x2 = (0.1 * torch.randn(B, T)).cuda()
    lin = m(2.0 * x.detach() + x2) - 2.0 * spec.detach() - m(x2)
    assert float(lin.abs().max()) < 1e-4 * float(spec.detach().abs().max())
    g = torch.randn_like(spec)
    (gx,) = torch.autograd.grad(spec, x, g)
    lhs = float((spec.detach().double() * g.double()).sum())
    rhs = float((x.detach().double() * gx.double()).sum())
    assert abs(lhs - rhs) < 1e-5 * max(abs(lhs), abs(rhs), 1.0)


def test_error_behaviour():
    with pytest.raises(AssertionError):
        MDCT2(256, 128, 512, kbdwin, device="cuda")                       # win > n_fft (mdct.py:371)
    im = IMDCT2(256, 128, 256, kbdwin, device="cuda")
    with pytest.raises(AssertionError):
        im(torch.zeros(2, 3, 128).cuda())                                 # last dim must be n_fft (mdct.py:434)
    with pytest.raises(AssertionError):
        im(torch.zeros(3, 256).cuda())                                    # 3-D only (mdct.py:433)
    with pytest.raises((NotImplementedError, ValueError, RuntimeError)):
        MDCT2(384, 192, 384, None, device="cuda")(torch.zeros(2, 4000).cuda())   # not a power of two


@pytest.mark.parametrize("n_fft", [64, 128, 256, 512, 1024])
@pytest.mark.parametrize("B,T", [(1, 0), (3, 20003), (2, 65536), (5, 33 * 512 + 4)])
def test_register_fft_path_matches_oracle_and_general_kernels(n_fft, B, T):
    """canonical geometry (win = n_fft, hop = n_fft/2): the register-FFT kernels (frame pairs, ragged last tiles, odd frame
    counts, unaligned lengths, signals shorter than one tile) against the float64 oracle and the general kernels, forward
    and both autograd adjoints"""
    from pix2pixhdaudiosr_b200 import _lib
    T = T or 2 * n_fft + 6
    torch.manual_seed(n_fft + T)
    hop = n_fft // 2
    x = 0.1 * torch.randn(B, T)
    w = kbdwin(n_fft).float()
    m = MDCT2(n_fft, hop, n_fft, w, True, device="cuda")
    im = IMDCT2(n_fft, hop, n_fft, w, True, device="cuda")

    def run():
        xc = x.cuda().requires_grad_(True)
        spec = m(xc)
        g = torch.randn(spec.shape, generator=torch.Generator().manual_seed(7)).cuda()
        (gx,) = torch.autograd.grad(spec, xc, g)
        sc = spec.detach().clone().requires_grad_(True)
        y = im(sc)
        gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(8)).cuda()
        (gs,) = torch.autograd.grad(y, sc, gy)
        return [t.detach().cpu().numpy() for t in (spec, gx, y, gs)]

    fast = run()
    old = _lib.lib().p2p_mdct2_force_general(1)
    try:
        general = run()
    finally:
        _lib.lib().p2p_mdct2_force_general(old)
    want = O2.mdct2_forward(x.numpy(), w.numpy(), n_fft, hop, True)
    assert fast[0].shape == want.shape
    assert relerr(fast[0], want) < TOL
    assert relerr(fast[2], O2.imdct2_forward(fast[0].astype(np.float64), w.numpy(), n_fft, hop, True)) < TOL
    for a, b in zip(fast, general):
        assert a.shape == b.shape and relerr(a, b) < TOL
