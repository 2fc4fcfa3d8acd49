"""GPU parity: MDCT4 / IMDCT4 (through the C ABI) against the reference's golden outputs and the
CPU oracle.  Tolerance: 1e-5 relative (max-abs error over max-abs value), BASELINE.json north_star."""
import numpy as np
import pytest
import torch

from cases import CASES, relerr
from oracle import mdct4_oracle as O
from pix2pixhdaudiosr_b200.models.mdct import IMDCT4, MDCT4
from pix2pixhdaudiosr_b200.util.util import kbdwin

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _mods(name):
    n_fft, hop, win, center, kbd, one_d = CASES[name]
    w = kbdwin if kbd else None
    return (MDCT4(n_fft, hop, win, w, center, device="cuda"), IMDCT4(n_fft, hop, win, w, center, device="cuda"))


@pytest.mark.parametrize("name", list(CASES))
def test_forward_and_backward_match_reference(golden, name):
    m, im = _mods(name)
    x = torch.from_numpy(golden[name + "/x"]).cuda().requires_grad_(True)
    spec = m(x)
    ref = golden[name + "/spec"]
    assert spec.dtype == torch.float32 and tuple(spec.shape) == ref.shape
    assert relerr(spec.detach().cpu().numpy(), ref) < TOL
    (gx,) = torch.autograd.grad(spec, x, torch.from_numpy(golden[name + "/gs"]).cuda().float())
    assert relerr(gx.cpu().numpy(), golden[name + "/gx"]) < TOL

    s3 = torch.from_numpy(ref if ref.ndim == 3 else ref[None]).cuda().float().requires_grad_(True)
    y = im(s3)
    yref = golden[name + "/y"]
    assert tuple(y.shape) == yref.shape and y.dim() == 4          # [B,1,1,T] (mdct.py:560)
    assert relerr(y.detach().cpu().numpy(), yref) < TOL
    (gs,) = torch.autograd.grad(y, s3, torch.from_numpy(golden[name + "/gy"]).cuda().float())
    assert relerr(gs.cpu().numpy(), golden[name + "/gspec"]) < TOL


def test_config1_round_trip(golden_config1):
    """BASELINE config 1: 16 x 130816, n_fft 1024, hop 512, KBD, center -> [16,257,512]; MSE <= 1e-12."""
    torch.manual_seed(1234)
    x = 0.1 * torch.randn(16, 130816)
    m = MDCT4(1024, 512, 1024, kbdwin, True, device="cuda")
    im = IMDCT4(1024, 512, 1024, kbdwin, True, device="cuda")
    spec = m(x.cuda())
    assert tuple(spec.shape) == (16, 257, 512)
    assert tuple(m(x[0].cuda()).shape) == (257, 512)              # README.md:114
    assert relerr(spec[::5, ::37, :].cpu().numpy(), golden_config1["spec_rows"]) < TOL
    y = im(spec)
    assert tuple(y.shape) == (16, 1, 1, 131072)
    assert relerr(y[:, 0, 0, ::101].cpu().numpy(), golden_config1["y_sample"]) < TOL
    mse = float(((x.cuda() - y[:, 0, 0, :130816]) ** 2).mean())
    assert mse <= 1e-12, mse


@pytest.mark.parametrize("n_fft", [16, 32, 128, 256, 512, 1024, 2048, 4096])
def test_against_oracle_random(n_fft):
    rng = np.random.default_rng(n_fft)
    hop, B, T = n_fft // 2, 3, 5 * n_fft + 17
    x = rng.standard_normal((B, T)).astype(np.float32)
    w = O.kbdwin(n_fft)
    m = MDCT4(n_fft, hop, None, torch.from_numpy(w), True, device="cuda")
    im = IMDCT4(n_fft, hop, None, torch.from_numpy(w), True, device="cuda")
    spec = m(torch.from_numpy(x).cuda())
    ref = O.mdct4_forward(x, w, n_fft, hop, True)
    assert relerr(spec.cpu().numpy(), ref) < TOL
    y = im(spec)
    assert relerr(y.cpu().numpy(), O.imdct4_forward(ref, w, n_fft, hop, True)) < TOL


def test_large_batch_properties():
    """BASELINE-size inputs (sweep config: B=256, T=130816): round trip, linearity, adjointness."""
    torch.manual_seed(0)
    B, T = 256, 130816
    m = MDCT4(1024, 512, 1024, kbdwin, True, device="cuda")
    im = IMDCT4(1024, 512, 1024, kbdwin, True, out_length=T, device="cuda")
    x = torch.randn(B, T, device="cuda")
    z = torch.randn(B, T, device="cuda")
    sx, sz = m(x), m(z)
    assert tuple(sx.shape) == (B, 257, 512)
    # B % hop == 0 and T % hop != 0 -> the reference drops the partial frame (quirk H8a)
    assert tuple(m(x.repeat(2, 1)[:512, :1024 * 8 + 5]).shape) == (512, 17, 512)   # 18 with true T-based padding
    lin = m(2.0 * x - 3.0 * z)
    assert relerr(lin.cpu().numpy(), (2.0 * sx - 3.0 * sz).cpu().numpy()) < TOL
    y = im(sx)
    covered = 255 * 512                    # samples covered by two frames
    assert float(((y[:, 0, 0, :covered] - x[:, :covered]) ** 2).mean()) < 1e-11
    # <m(x), s> == <x, m^T(s)>
    s = torch.randn_like(sx)
    xg = x.clone().requires_grad_(True)
    (g,) = torch.autograd.grad(m(xg), xg, s)
    lhs = float((sx.double() * s.double()).sum())
    rhs = float((x.double() * g.double()).sum())
    assert abs(lhs - rhs) <= 1e-5 * abs(lhs)


def test_ragged_and_edge_inputs():
    m = MDCT4(256, 128, 256, kbdwin, True, device="cuda")
    im = IMDCT4(256, 128, 256, kbdwin, True, device="cuda")
    assert tuple(m(torch.zeros(0, 1000, device="cuda")).shape) == (0, 8, 128)      # empty batch
    x = torch.randn(2, 3, 1000)                                                   # extra leading dims, CPU input
    s = m(x)
    assert s.is_cuda and tuple(s.shape) == (2, 3, 9, 128)
    w = O.kbdwin(256)
    ref = O.mdct4_forward(x.numpy().reshape(6, 1000), w, 256, 128, True, signal_len=2)
    assert relerr(s.reshape(6, 9, 128).cpu().numpy(), ref) < TOL
    one = im(torch.randn(1, 1, 128, device="cuda"))                                # a single frame
    assert tuple(one.shape) == (1, 1, 1, 0)
    with pytest.raises(RuntimeError):        # shorter than one window: unfold's error in the reference
        MDCT4(256, 128, 256, kbdwin, False, device="cuda")(torch.zeros(256, 100, device="cuda"))


def test_reflect_pad_mode_and_gradcheck_like():
    x = torch.randn(2, 2000, device="cuda", requires_grad=True)
    m = MDCT4(256, 128, 256, kbdwin, True, pad_mode="reflect", device="cuda")
    s = m(x)
    xp = np.pad(x.detach().cpu().numpy(), [(0, 0), (128, 128 + 128 - 2)], mode="reflect")
    ref = O.mdct4_forward(xp, O.kbdwin(256), 256, 128, False, signal_len=256)
    assert relerr(s.detach().cpu().numpy(), ref) < TOL
    s.sum().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all()


@pytest.mark.parametrize("n_fft", [256, 512, 1024, 2048, 4096])
def test_fast_and_general_paths_agree(n_fft):
    """The register-FFT fast path and the general shared-memory kernels compute the same thing."""
    from pix2pixhdaudiosr_b200 import _lib
    torch.manual_seed(n_fft)
    hop = n_fft // 2
    x = torch.randn(5, 37 * hop + 123, device="cuda")
    m = MDCT4(n_fft, hop, n_fft, kbdwin, True, device="cuda")
    im = IMDCT4(n_fft, hop, n_fft, kbdwin, True, device="cuda")
    fast_s = m(x)
    fast_y = im(fast_s)
    old = _lib.lib().p2p_mdct4_force_general(1)
    try:
        gen_s = m(x)
        gen_y = im(fast_s)
    finally:
        _lib.lib().p2p_mdct4_force_general(old)
    assert relerr(fast_s.cpu().numpy(), gen_s.cpu().numpy()) < 2e-6
    assert relerr(fast_y.cpu().numpy(), gen_y.cpu().numpy()) < 2e-6
    # unaligned signal pointer / odd T -> the fast path's non-TMA staging
    xo = torch.randn(3, 20 * hop + 7, device="cuda")[:, 1:]
    w = O.kbdwin(n_fft)
    ref = O.mdct4_forward(xo.cpu().numpy(), w, n_fft, hop, True)
    assert relerr(m(xo).cpu().numpy(), ref) < TOL
