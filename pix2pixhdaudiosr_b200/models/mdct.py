"""Drop-in `MDCT4` / `IMDCT4` (reference: models/mdct.py:461-513 and :516-566) running on the fused
sm_100a kernels of csrc/mdct4.cu through the C ABI (include/p2p_b200.h).

Same constructor arguments, accepted ranks, frame-count arithmetic (including the reference's
`len(signal)` quirk, mdct.py:488), `[B,1,1,T]` IMDCT result and autograd.  Differences, all stated
in DESIGN.md: results are float32 (the reference silently promotes to float64 through its
complex128 twiddles); there is no CPU path — tensors are moved to `device` as the reference does
(mdct.py:503, 547) and a missing CUDA library or a CPU `device` raises.
"""
import ctypes

import torch
import torch.nn as nn
import torch.nn.functional as F_

from .. import _lib


class _Plan:
    """Per-module geometry plus device-resident window and twiddle tables (built lazily per device)."""

    def __init__(self, n_fft, hop, win, window_cpu):
        self.n_fft, self.hop, self.win = int(n_fft), int(hop), int(win)
        self.window_cpu = window_cpu.detach().to("cpu", torch.float32).contiguous()
        self._dev = {}

    def on(self, device):
        key = (device.type, device.index)
        if key not in self._dev:
            lib = _lib.lib()
            words = lib.p2p_mdct4_table_words(self.n_fft)
            if words < 0:
                _lib.check(words)
            host = torch.empty(words, dtype=torch.int32)
            _lib.check(lib.p2p_mdct4_fill_tables(self.n_fft, ctypes.c_void_p(host.data_ptr())))
            self._dev[key] = (self.window_cpu.to(device), host.to(device))
        return self._dev[key]


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what}: the B200 path needs a CUDA tensor/device, got {t.device} (no CPU fallback)")


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _analysis(x, plan, n_frames, offset, scale):
    _require_cuda(x, "mdct4 analysis")
    x = x.to(torch.float32).contiguous()
    B, T = x.shape
    window, tables = plan.on(x.device)
    spec = torch.empty((B, n_frames, plan.n_fft // 2), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().p2p_mdct4_analysis(
            x.data_ptr(), spec.data_ptr(), window.data_ptr(), tables.data_ptr(), B, T, n_frames,
            plan.n_fft, plan.hop, plan.win, offset, scale, _stream()))
    return spec


def _synthesis(spec, plan, t_out, crop_left, scale):
    _require_cuda(spec, "mdct4 synthesis")
    spec = spec.to(torch.float32).contiguous()
    B, n_frames, _ = spec.shape
    window, tables = plan.on(spec.device)
    out = torch.empty((B, t_out), dtype=torch.float32, device=spec.device)
    with torch.cuda.device(spec.device):
        _lib.check(_lib.lib().p2p_mdct4_synthesis(
            spec.data_ptr(), out.data_ptr(), window.data_ptr(), tables.data_ptr(), B, n_frames, t_out,
            plan.n_fft, plan.hop, plan.win, crop_left, scale, _stream()))
    return out


class _AnalysisFn(torch.autograd.Function):
    """spec = scale * C W frame(x); adjoint = the synthesis kernel."""

    @staticmethod
    def forward(ctx, x, plan, n_frames, offset, scale):
        ctx.plan, ctx.offset, ctx.scale, ctx.T = plan, offset, scale, x.shape[-1]
        return _analysis(x, plan, n_frames, offset, scale)

    @staticmethod
    def backward(ctx, g):
        return _SynthesisFn.apply(g, ctx.plan, ctx.T, ctx.offset, ctx.scale), None, None, None, None


class _SynthesisFn(torch.autograd.Function):
    """out = scale * crop(OLA(W C^T spec)); adjoint = the analysis kernel."""

    @staticmethod
    def forward(ctx, spec, plan, t_out, crop_left, scale):
        ctx.plan, ctx.crop_left, ctx.scale, ctx.F = plan, crop_left, scale, spec.shape[1]
        return _synthesis(spec, plan, t_out, crop_left, scale)

    @staticmethod
    def backward(ctx, g):
        return _AnalysisFn.apply(g, ctx.plan, ctx.F, ctx.crop_left, ctx.scale), None, None, None, None


def _make_window(window, win_length, device):
    # mdct.py:470-478 / 526-534
    if window is None:
        window = torch.ones
    if callable(window):
        win_length = int(win_length)
        window = window(win_length)
    else:
        win_length = len(window)
    return window.to(device) if _device_usable(device) else window, win_length


def _device_usable(device):
    d = torch.device(device)
    return d.type != "cuda" or torch.cuda.is_available()


class MDCT4(nn.Module):
    def __init__(self, n_fft=2048, hop_length=None, win_length=None, window=None, center=True,
                 pad_mode='constant', device='cuda') -> None:
        super().__init__()
        self.n_fft = n_fft
        self.pad_mode = pad_mode
        self.device = device
        self.hop_length = hop_length
        self.center = center
        self.window, self.win_length = _make_window(window, win_length, device)
        assert self.win_length <= self.n_fft, 'Window lenth %d should be no more than fft length %d' % (self.win_length, self.n_fft)
        assert self.hop_length <= self.win_length, 'You hopped more than one frame'
        self._plan = _Plan(self.n_fft, self.hop_length, self.win_length, self.window)

    @property
    def exp1(self):  # mdct.py:483, kept for callers that read it
        return torch.exp(-1j * torch.pi / self.n_fft * torch.arange(0, self.n_fft, 1, dtype=torch.float64))

    @property
    def exp2(self):  # mdct.py:484
        return torch.exp(-1j * (torch.pi / (2 * self.n_fft) + torch.pi / 4) * torch.arange(1, self.n_fft, 2, dtype=torch.float64))

    def pads(self, signal_len):
        """(start_pad, end_pad) exactly as mdct.py:488-496; signal_len = len(signal)."""
        start_pad = self.hop_length if self.center else 0
        additional_len = signal_len % self.hop_length
        end_pad = start_pad
        if additional_len:
            end_pad = start_pad + self.hop_length - additional_len
        return start_pad, end_pad

    def forward(self, signal):
        # len(signal) is the first dimension (the batch size for 2-D input) -- reference behaviour.
        start_pad, end_pad = self.pads(int(len(signal)))
        lead, T = signal.shape[:-1], signal.shape[-1]
        x = signal.to(self.device, torch.float32).reshape(-1, T)
        if self.pad_mode != 'constant':
            x = F_.pad(x, (start_pad, end_pad), mode=self.pad_mode)   # torch pads; kernel frames
            offset, T_eff = 0, T + start_pad + end_pad
        else:
            offset, T_eff = start_pad, T
        padded = T + start_pad + end_pad
        if padded < self.win_length:
            raise RuntimeError('maximum size for tensor at dimension %d is %d but size is %d'
                               % (signal.dim() - 1, padded, self.win_length))   # what unfold raises
        n_frames = (padded - self.win_length) // self.hop_length + 1
        assert x.shape[-1] == T_eff
        spec = _AnalysisFn.apply(x, self._plan, n_frames, offset, 1.0)
        return spec.reshape(*lead, n_frames, self.n_fft // 2)


class IMDCT4(nn.Module):
    def __init__(self, n_fft=2048, hop_length=None, win_length=None, window=None, center=True,
                 pad_mode='constant', out_length=None, device='cuda') -> None:
        super().__init__()
        self.n_fft = n_fft
        self.pad_mode = pad_mode
        self.device = device
        self.hop_length = hop_length
        self.center = center
        self.out_length = out_length
        self.window, self.win_length = _make_window(window, win_length, device)
        assert self.win_length <= self.n_fft, 'Window lenth %d should be no more than fft length %d' % (self.win_length, self.n_fft)
        assert self.hop_length <= self.win_length, 'You hopped more than one frame'
        self._plan = _Plan(self.n_fft, self.hop_length, self.win_length, self.window)

    @property
    def exp1(self):  # mdct.py:539
        return torch.exp(-1j * (torch.pi / (2 * self.n_fft) + torch.pi / 4) * torch.arange(1, self.n_fft, 2, dtype=torch.float64))

    @property
    def exp2(self):  # mdct.py:540
        return torch.exp(-1j * torch.pi / (2 * self.n_fft) * torch.arange(0, 2 * self.n_fft, 2, dtype=torch.float64))

    def out_geometry(self, n_frames):
        """(crop_left, T_out) of mdct.py:559-566 for a [B, n_frames, n_fft/2] input."""
        full = (n_frames - 1) * self.hop_length + self.win_length
        idx = range(full)
        if self.center:
            idx = idx[self.win_length // 2: -self.win_length // 2]
        if self.out_length is not None:
            idx = idx[:self.out_length]
        return (idx.start if len(idx) else 0), len(idx)

    def forward(self, signal):
        assert signal.dim() == 3, 'Only tensors shaped in BHW are supported, got tensor of shape %s' % (str(signal.size()))
        assert signal.size()[-1] == self.n_fft // 2, 'The last dim of input tensor should match the n_fft. Expected %d ,got %d' % (self.n_fft, signal.size()[-1])
        spec = signal.to(self.device, torch.float32)
        crop_left, t_out = self.out_geometry(spec.shape[1])
        out = _SynthesisFn.apply(spec, self._plan, t_out, crop_left, 4.0 / self.n_fft)
        return out.reshape(spec.shape[0], 1, 1, t_out)


# ----------------------------------------------------------------------------------------------------------------
# MDCT2 / IMDCT2 (reference: models/mdct.py:352-454 with dct/dct_native.py:7-67) on csrc/mdct2.cu
# ----------------------------------------------------------------------------------------------------------------
class _Plan2(_Plan):
    """geometry + device-resident window and DCT-II tables (p2p_mdct2_fill_tables)"""

    def on(self, device):
        key = (device.type, device.index)
        if key not in self._dev:
            lib = _lib.lib()
            words = lib.p2p_mdct2_table_words(self.n_fft)
            if words < 0:
                _lib.check(words)
            host = torch.empty(words, dtype=torch.int32)
            _lib.check(lib.p2p_mdct2_fill_tables(self.n_fft, ctypes.c_void_p(host.data_ptr())))
            self._dev[key] = (self.window_cpu.to(device), host.to(device))
        return self._dev[key]


def _analysis2(x, plan, n_frames, offset, scale, c0):
    _require_cuda(x, "mdct2 analysis")
    x = x.to(torch.float32).contiguous()
    B, T = x.shape
    window, tables = plan.on(x.device)
    spec = torch.empty((B, n_frames, plan.n_fft), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().p2p_mdct2_analysis(
            x.data_ptr(), spec.data_ptr(), window.data_ptr(), tables.data_ptr(), B, T, n_frames,
            plan.n_fft, plan.hop, plan.win, offset, scale, c0, _stream()))
    return spec


def _synthesis2(spec, plan, t_out, crop_left, scale, c0):
    _require_cuda(spec, "mdct2 synthesis")
    spec = spec.to(torch.float32).contiguous()
    B, n_frames, _ = spec.shape
    window, tables = plan.on(spec.device)
    out = torch.empty((B, t_out), dtype=torch.float32, device=spec.device)
    with torch.cuda.device(spec.device):
        _lib.check(_lib.lib().p2p_mdct2_synthesis(
            spec.data_ptr(), out.data_ptr(), window.data_ptr(), tables.data_ptr(), B, n_frames, t_out,
            plan.n_fft, plan.hop, plan.win, crop_left, scale, c0, _stream()))
    return out


class _Analysis2Fn(torch.autograd.Function):
    """spec = scale * diag(c0) D W frame(x), D = DCT_2N; adjoint = synthesis(scale / n_fft, 2 c0 ... ) -- see csrc/mdct2.cu:
    the synthesis kernel computes 2 sum_k c0_k X[k] cos(.), so D^T diag(c0) g = (1/n_fft) * synthesis-sum with the same c0."""

    @staticmethod
    def forward(ctx, x, plan, n_frames, offset, scale, c0):
        ctx.plan, ctx.offset, ctx.scale, ctx.c0, ctx.T = plan, offset, scale, c0, x.shape[-1]
        return _analysis2(x, plan, n_frames, offset, scale, c0)

    @staticmethod
    def backward(ctx, g):
        return _Synthesis2Fn.apply(g, ctx.plan, ctx.T, ctx.offset, ctx.scale / ctx.plan.n_fft, ctx.c0), None, None, None, None, None


class _Synthesis2Fn(torch.autograd.Function):
    """out = scale * crop(OLA(W E diag(c0) spec)), E[n,k] = 2 cos(.); adjoint = analysis(scale * n_fft, c0)
    (E^T u = n_fft * DCT_2N(u))."""

    @staticmethod
    def forward(ctx, spec, plan, t_out, crop_left, scale, c0):
        ctx.plan, ctx.crop_left, ctx.scale, ctx.c0, ctx.F = plan, crop_left, scale, c0, spec.shape[1]
        return _synthesis2(spec, plan, t_out, crop_left, scale, c0)

    @staticmethod
    def backward(ctx, g):
        return _Analysis2Fn.apply(g, ctx.plan, ctx.F, ctx.crop_left, ctx.scale * ctx.plan.n_fft, ctx.c0), None, None, None, None, None


class MDCT2(MDCT4):
    """models/mdct.py:352-398.  `dct_op` is accepted for signature compatibility (the reference asserts it is callable);
    the transform itself is always DCT_2N_native's (dct/dct_native.py:7-33), fused with framing and windowing.
    forward(signal, return_ola=False) -> [B, F, n_fft] (and the windowed frames [B, F, win] when return_ola)."""

    def __init__(self, n_fft=2048, hop_length=None, win_length=None, window=None, center=True, pad_mode='constant',
                 device='cuda', dct_op=None) -> None:
        super().__init__(n_fft, hop_length, win_length, window, center, pad_mode, device)
        assert dct_op is None or callable(dct_op)
        self.dct = dct_op
        self._plan = _Plan2(self.n_fft, self.hop_length, self.win_length, self.window)

    def forward(self, signal, return_ola=False):
        start_pad, end_pad = self.pads(int(len(signal)))
        lead, T = signal.shape[:-1], signal.shape[-1]
        x = signal.to(self.device, torch.float32).reshape(-1, T)
        if self.pad_mode != 'constant':
            x = F_.pad(x, (start_pad, end_pad), mode=self.pad_mode)
            offset = 0
        else:
            offset = start_pad
        padded = T + start_pad + end_pad
        if padded < self.win_length:
            raise RuntimeError('maximum size for tensor at dimension %d is %d but size is %d'
                               % (signal.dim() - 1, padded, self.win_length))
        n_frames = (padded - self.win_length) // self.hop_length + 1
        spec = _Analysis2Fn.apply(x, self._plan, n_frames, offset, 1.0, 1.0).reshape(*lead, n_frames, self.n_fft)
        if not return_ola:
            return spec
        from .. import ops
        window, _ = self._plan.on(x.device)
        frames = ops.frame_window(x.detach(), window, n_frames, self.hop_length, offset)
        return spec, frames.reshape(*lead, n_frames, self.win_length)


class IMDCT2(IMDCT4):
    """models/mdct.py:401-454: IDCT_2N_native(X) / 2 -> window -> overlap-add -> crop; result [B,1,1,T]."""

    def __init__(self, n_fft=2048, hop_length=None, win_length=None, window=None, center=True, pad_mode='constant',
                 out_length=None, device='cuda', idct_op=None) -> None:
        super().__init__(n_fft, hop_length, win_length, window, center, pad_mode, out_length, device)
        assert idct_op is None or callable(idct_op)
        self.idct = idct_op
        self._plan = _Plan2(self.n_fft, self.hop_length, self.win_length, self.window)

    def forward(self, signal):
        assert signal.dim() == 3, 'Only tensors shaped in BHW are supported, got tensor of shape %s' % (str(signal.size()))
        assert signal.size()[-1] == self.n_fft, 'The last dim of input tensor should match the n_fft. Expected %d ,got %d' % (self.n_fft, signal.size()[-1])
        spec = signal.to(self.device, torch.float32)
        crop_left, t_out = self.out_geometry(spec.shape[1])
        out = _Synthesis2Fn.apply(spec, self._plan, t_out, crop_left, 0.5, 0.5)
        return out.reshape(spec.shape[0], 1, 1, t_out)
