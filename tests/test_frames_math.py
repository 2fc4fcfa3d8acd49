"""CPU statement of the closed-form gradients csrc/frames.cu implements (frame-match loss, dB map, spectrogram decode),
checked against torch autograd through the oracle's restatement of pix2pixHD_model.py:229-237, 314-320, 408-415.
The GPU tests (tests/test_time_gpu.py) check the kernels themselves against the same oracle functions."""
import math

import numpy as np
import torch

from oracle import mdct4_oracle as M4
from oracle import pix2pixhd_oracle as O


def frame_match_closed_form(y, w, weight):
    """frame_match_kernel: every element of dy receives at most one term"""
    B, F, N = y.shape
    half = N // 2
    d = y[:, :-1, half:] * w[:half] - y[:, 1:, :half] * w[half:]
    cnt = B * (F - 1) * half
    loss = weight * (d ** 2).sum() / cnt
    dy = np.zeros_like(y)
    dy[:, :-1, half:] = 2 * weight / cnt * d * w[:half]          # first member of the pair (f, f+1)
    dy[:, 1:, :half] = -2 * weight / cnt * d * w[half:]          # second member
    return loss, dy


def test_frame_match_gradient():
    rng = np.random.default_rng(0)
    y = rng.standard_normal((3, 6, 32))
    w = M4.kbdwin(32).astype(np.float64)
    loss, dy = frame_match_closed_form(y, w, 10.0)
    yt = torch.from_numpy(y).float().requires_grad_(True)
    want = O.match_loss(yt, dict(win_length=32, lambda_mat=10.0))
    (g,) = torch.autograd.grad(want, yt)
    assert abs(loss - float(want.detach())) < 1e-5 * float(want.detach())
    assert np.abs(dy - g.numpy()).max() < 1e-5 * np.abs(dy).max()
    assert np.all(dy[:, 0, :16] == 0) and np.all(dy[:, -1, 16:] == 0)       # the two half frames no pair touches


def test_db_map_gradient():
    """frames_db_bwd_kernel: d/dy 20 log10(max(|g y|, amin)) = (20 / ln 10) / y above the clamp, 0 below"""
    y = torch.tensor([1e-9, -3e-8, 2e-7, -0.5, 3.0], dtype=torch.float64, requires_grad=True)
    gain, amin = 1.7, 1e-7
    out = 20 * torch.log10(torch.clamp((gain * y).abs(), min=amin)) - 20
    (g,) = torch.autograd.grad(out.sum(), y)
    v = gain * y.detach().numpy()
    closed = np.where(np.abs(v) >= amin, (20 / math.log(10)) * gain * np.sign(v) / np.abs(v), 0.0)
    assert np.allclose(closed, g.numpy(), rtol=1e-12)


def test_decode_gradient():
    """spectro_decode_bwd_kernel: d coef / d x_c = +-gain/(2 alpha - 1) * 0.05 ln10 (max - min) sign(x_c) * 10 * 10^(0.05 s_c)"""
    torch.manual_seed(1)
    x = (torch.rand(2, 2, 5, 7, dtype=torch.float64) * 2 - 1).requires_grad_(True)
    norm = {'max': torch.tensor(-3.0, dtype=torch.float64), 'min': torch.tensor(-120.0, dtype=torch.float64)}
    cfg = dict(alpha=0.6, min_value=1e-7)
    coef = O.decode_coefficients(x, norm, cfg)                          # [B, F, M]
    g = torch.randn_like(coef)
    (gx,) = torch.autograd.grad(coef, x, g)
    xd = x.detach().numpy()
    span = float(norm['max'] - norm['min'])
    e = 10.0 * 10.0 ** (0.05 * (np.abs(xd) * span + float(norm['min'])))
    c = 1.0 / (2 * 0.6 - 1) * 0.05 * math.log(10) * span
    gk = g.numpy().transpose(0, 2, 1)                                   # [B, M, F]
    closed = np.stack([gk * c * e[:, 0] * np.sign(xd[:, 0]), -gk * c * e[:, 1] * np.sign(xd[:, 1])], axis=1)
    assert np.abs(closed - gx.numpy()).max() < 1e-10 * np.abs(closed).max()
