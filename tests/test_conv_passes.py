"""CPU: the implicit-GEMM pass builders (pix2pixhdaudiosr_b200/ops.py) reproduce torch's conv2d /
conv_transpose2d forward, input gradient and weight gradient for every layer flavour of networks.py."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import pass_emulator as E
from pix2pixhdaudiosr_b200 import ops

# (Cin, Cout, k, stride, pad, reflect?)  -- 7x7 reflect, 3x3 reflect, 3x3 s2, 4x4 s2 p2, 4x4 s1 p2
CONVS = [(2, 3, 7, 1, 3, True), (3, 3, 3, 1, 1, True), (3, 4, 3, 2, 1, False), (4, 5, 4, 2, 2, False), (3, 2, 4, 1, 2, False)]


def _nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous().numpy().astype(np.float64)


@pytest.mark.parametrize("cfg", CONVS)
@pytest.mark.parametrize("hw", [(9, 8), (6, 11)])
def test_conv_passes(cfg, hw):
    Cin, Cout, k, s, p, reflect = cfg
    H, W = hw
    torch.manual_seed(0)
    x = torch.randn(2, Cin, H, W, dtype=torch.float64, requires_grad=True)
    w = torch.randn(Cout, Cin, k, k, dtype=torch.float64, requires_grad=True)
    b = torch.randn(Cout, dtype=torch.float64)
    xin = F.pad(x, (p, p, p, p), mode="reflect") if reflect else x
    y = F.conv2d(xin, w, b, stride=s, padding=0 if reflect else p)
    gy = torch.randn_like(y)
    gx, gw = torch.autograd.grad(y, (x, w), gy)
    mode = ops.PAD_REFLECT if reflect else ops.PAD_ZERO
    wphys = w.detach().permute(2, 3, 0, 1).contiguous().numpy().reshape(-1)

    (P,) = ops.conv_fwd_passes(2, H, W, Cin, Cout, k, k, s, p, mode)
    out = np.full((2, P.OH, P.OW, Cout), np.nan)
    E.run_pass(P, _nhwc(x.detach()), wphys, b.numpy(), out)
    assert np.allclose(out, _nhwc(y.detach()), atol=1e-10)

    passes, (DH, DW) = ops.conv_dgrad_passes(2, H, W, Cin, Cout, k, k, s, p, mode)
    buf = np.full((2, DH, DW, Cin), np.nan)
    for Pd in passes:
        E.run_pass(Pd, _nhwc(gy), wphys, None, buf)
    assert not np.isnan(buf).any()
    dx = E.fold_reflect(buf, p) if reflect else buf
    assert np.allclose(dx, _nhwc(gx), atol=1e-10)

    dw = np.zeros_like(wphys)
    E.run_wgrad(P, _nhwc(x.detach()), _nhwc(gy), dw)
    assert np.allclose(dw.reshape(k, k, Cout, Cin), gw.permute(2, 3, 0, 1).numpy(), atol=1e-10)


@pytest.mark.parametrize("hw", [(4, 5), (3, 3)])
def test_conv_transpose_passes(hw):
    Cin, Cout, k, s, p, op = 3, 2, 3, 2, 1, 1          # networks.py:154, 203
    H, W = hw
    torch.manual_seed(1)
    x = torch.randn(2, Cin, H, W, dtype=torch.float64, requires_grad=True)
    w = torch.randn(Cin, Cout, k, k, dtype=torch.float64, requires_grad=True)
    b = torch.randn(Cout, dtype=torch.float64)
    y = F.conv_transpose2d(x, w, b, stride=s, padding=p, output_padding=op)
    gy = torch.randn_like(y)
    gx, gw = torch.autograd.grad(y, (x, w), gy)
    wphys = w.detach().permute(2, 3, 1, 0).contiguous().numpy().reshape(-1)

    passes, (OH, OW) = ops.convT_fwd_passes(2, H, W, Cin, Cout, k, k, s, p, op)
    assert (OH, OW) == tuple(y.shape[2:]) and len(passes) == 4
    assert sorted(len(P.dh[:P.ntaps]) for P in passes) == [1, 2, 2, 4]      # 9 taps over the 4 output parities
    out = np.full((2, OH, OW, Cout), np.nan)
    for P in passes:
        E.run_pass(P, _nhwc(x.detach()), wphys, b.numpy(), out)
    assert np.allclose(out, _nhwc(y.detach()), atol=1e-10)

    (Pd,) = ops.convT_dgrad_passes(2, H, W, Cin, Cout, k, k, s, p, op)
    dx = np.full((2, H, W, Cin), np.nan)
    E.run_pass(Pd, _nhwc(gy), wphys, None, dx)
    assert np.allclose(dx, _nhwc(gx), atol=1e-10)

    dw = np.zeros_like(wphys)
    for P in passes:
        E.run_wgrad(P, _nhwc(x.detach()), _nhwc(gy), dw)
    assert np.allclose(dw.reshape(k, k, Cout, Cin), gw.permute(2, 3, 1, 0).numpy(), atol=1e-10)


def test_reflect_fold_is_adjoint_of_reflect_pad():
    torch.manual_seed(2)
    x = torch.randn(1, 2, 5, 6, dtype=torch.float64, requires_grad=True)
    for p in (1, 3):
        xp = F.pad(x, (p, p, p, p), mode="reflect")
        g = torch.randn_like(xp)
        (gx,) = torch.autograd.grad(xp, x, g)
        assert np.allclose(E.fold_reflect(_nhwc(g), p), _nhwc(gx), atol=1e-12)


def test_transposed_image_discriminator_keeps_the_reference_state_dict():
    """MultiscaleDiscriminator(transposed_image=True) stores every conv weight [KW][KH][Cout][Cin] behind the same
    logical parameter: same keys / shapes / values in state_dict, physical buffer contiguous and transposed"""
    import torch
    from pix2pixhdaudiosr_b200.models import networks
    torch.manual_seed(3)
    a = networks.define_D(2, 8, 3, 'instance', False, 2, False, gpu_ids=[])
    b = networks.define_D(2, 8, 3, 'instance', False, 2, False, gpu_ids=[], transposed_image=True)
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa) == list(sb) and all(sa[k].shape == sb[k].shape for k in sa)
    b.load_state_dict(sa)
    for (ka, ma), (kb, mb) in zip(a.named_modules(), b.named_modules()):
        if isinstance(ma, networks.Conv2d):
            assert mb.tswap and not ma.tswap
            assert torch.equal(ma.weight.data, mb.weight.data)
            pa, pb = ma.physical_weight(), mb.physical_weight()
            assert pa.is_contiguous() and pb.is_contiguous()
            assert torch.equal(pb, pa.transpose(0, 1))


def test_time_domain_discriminator_layers_take_the_window_form_only_when_transposed():
    """time_D sees [frames = 128, n_fft = 1024] maps: with the long axis fastest the halo (K-1) * padded width exceeds the
    shifted-window kernel's 312 rows; transposed (what define_D(..., transposed_image=True) runs) every layer fits"""
    import torch
    from pix2pixhdaudiosr_b200 import ops
    for (H, W, Cin, Cout, stride) in [(1024, 128, 64, 128, 2), (513, 65, 128, 256, 2), (257, 33, 256, 512, 1)]:
        fwd_t = ops.conv_fwd_passes(8, H, W, Cin, Cout, 4, 4, stride, 2, ops.PAD_ZERO)
        fwd_w = ops.conv_fwd_passes(8, W, H, Cin, Cout, 4, 4, stride, 2, ops.PAD_ZERO)
        if stride == 1:
            assert ops._window_ok(torch.bfloat16, Cin, Cout, 4, 4, H + 4, W + 4, False) or not ops.USE_WINDOW
            assert not ops._window_ok(torch.bfloat16, Cin, Cout, 4, 4, W + 4, H + 4, False)
        else:
            assert ops._plan_window_passes(fwd_t, torch.bfloat16) is not None
            assert ops._plan_window_passes(fwd_w, torch.bfloat16) is None
        dg_t = ops.conv_dgrad_passes(8, H, W, Cin, Cout, 4, 4, stride, 2, ops.PAD_ZERO)
        assert dg_t
