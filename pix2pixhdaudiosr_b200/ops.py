"""Host side of the conv-GAN kernels: implicit-GEMM pass descriptors (pure host logic, unit-tested on CPU)
and torch.autograd.Function wrappers that call the C ABI (include/p2p_b200.h).  Internal activation layout is
NHWC ([N,H,W,C] contiguous); dtype float32 (exact path) or bfloat16 (tensor-core path)."""
import ctypes
import os

import torch

from . import _lib
from ._lib import ConvPass

PAD_ZERO, PAD_REFLECT = 0, 1
_DT = {torch.float32: 0, torch.bfloat16: 1}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError(f"the B200 path needs CUDA tensors, got {t.device} (no CPU fallback)")


# ---------------------------------------------------------------------------------------------------
# pass descriptors
# ---------------------------------------------------------------------------------------------------
def _make_pass(N, H, W, K, OH, OW, J, gy, gx, sy, sx, oys, oyo, oxs, oxo, taps, pad_mode, w_tap, w_k, w_j):
    """taps: list of (dh, dw, wtap)."""
    assert 1 <= len(taps) <= _lib.MAX_TAPS, len(taps)
    P = ConvPass()
    P.N, P.H, P.W, P.K, P.OH, P.OW, P.J = N, H, W, K, OH, OW, J
    P.gy, P.gx, P.sy, P.sx = gy, gx, sy, sx
    P.oys, P.oyo, P.oxs, P.oxo = oys, oyo, oxs, oxo
    P.ntaps, P.pad_mode = len(taps), pad_mode
    P.w_tap, P.w_k, P.w_j = w_tap, w_k, w_j
    for i, (dh, dw, wt) in enumerate(taps):
        assert -128 <= dh < 128 and -128 <= dw < 128
        P.dh[i], P.dw[i], P.wtap[i] = dh, dw, wt
    return P


def conv_out_size(H, k, stride, pad):
    return (H + 2 * pad - k) // stride + 1


def conv_fwd_passes(N, H, W, Cin, Cout, KH, KW, stride, pad, pad_mode):
    """y[n,oh,ow,co] = sum x[n, oh*s - pad + kh, ow*s - pad + kw, ci] W[kh,kw,co,ci]   (weight [KH][KW][Cout][Cin])"""
    OH, OW = conv_out_size(H, KH, stride, pad), conv_out_size(W, KW, stride, pad)
    taps = [(kh - pad, kw - pad, kh * KW + kw) for kh in range(KH) for kw in range(KW)]
    return [_make_pass(N, H, W, Cin, OH, OW, Cout, OH, OW, stride, stride, 1, 0, 1, 0, taps, pad_mode,
                       Cout * Cin, 1, Cin)]


def conv_dgrad_passes(N, H, W, Cin, Cout, KH, KW, stride, pad, pad_mode):
    """dx from dy [N,OH,OW,Cout].  Zero padding: writes dx [N,H,W,Cin].  Reflect padding: writes the gradient
    of the PADDED input [N,H+2p,W+2p,Cin]; the caller folds it (p2p_reflect_pad_fold).
    Returns (passes, (DH, DW)) with (DH, DW) the spatial extent written."""
    OH, OW = conv_out_size(H, KH, stride, pad), conv_out_size(W, KW, stride, pad)
    if pad_mode == PAD_REFLECT:
        DH, DW, p = H + 2 * pad, W + 2 * pad, 0         # coordinates of the padded input
    else:
        DH, DW, p = H, W, pad
    passes = []
    for ph in range(stride):
        for pw in range(stride):
            gy, gx = (DH - ph + stride - 1) // stride, (DW - pw + stride - 1) // stride
            if gy <= 0 or gx <= 0:
                continue
            taps = [((ph + p - kh) // stride, (pw + p - kw) // stride, kh * KW + kw)
                    for kh in range(KH) for kw in range(KW)
                    if (ph + p - kh) % stride == 0 and (pw + p - kw) % stride == 0]
            if not taps:
                return None, (DH, DW)       # a parity class nothing maps to: caller uses the zero-fill route
            passes.append(_make_pass(N, OH, OW, Cout, DH, DW, Cin, gy, gx, 1, 1, stride, ph, stride, pw, taps,
                                     PAD_ZERO, Cout * Cin, Cin, 1))
    return passes, (DH, DW)


def convT_out_size(H, k, stride, pad, out_pad):
    return (H - 1) * stride - 2 * pad + k + out_pad


def convT_fwd_passes(N, H, W, Cin, Cout, KH, KW, stride, pad, out_pad):
    """y[n,oh,ow,co] = sum_{ih*s - pad + kh = oh} x[n,ih,iw,ci] Wt[kh,kw,co,ci]; one pass per output parity."""
    OH, OW = convT_out_size(H, KH, stride, pad, out_pad), convT_out_size(W, KW, stride, pad, out_pad)
    passes = []
    for ph in range(stride):
        for pw in range(stride):
            gy, gx = (OH - ph + stride - 1) // stride, (OW - pw + stride - 1) // stride
            taps = [((ph + pad - kh) // stride, (pw + pad - kw) // stride, kh * KW + kw)
                    for kh in range(KH) for kw in range(KW)
                    if (ph + pad - kh) % stride == 0 and (pw + pad - kw) % stride == 0]
            if not taps:
                return None, (OH, OW)
            passes.append(_make_pass(N, H, W, Cin, OH, OW, Cout, gy, gx, 1, 1, stride, ph, stride, pw, taps,
                                     PAD_ZERO, Cout * Cin, 1, Cin))
    return passes, (OH, OW)


def convT_dgrad_passes(N, H, W, Cin, Cout, KH, KW, stride, pad, out_pad):
    """dx[n,ih,iw,ci] = sum dy[n, ih*s - pad + kh, ...] Wt[kh,kw,co,ci]: a strided gather over dy."""
    OH, OW = convT_out_size(H, KH, stride, pad, out_pad), convT_out_size(W, KW, stride, pad, out_pad)
    taps = [(kh - pad, kw - pad, kh * KW + kw) for kh in range(KH) for kw in range(KW)]
    return [_make_pass(N, OH, OW, Cout, H, W, Cin, H, W, stride, stride, 1, 0, 1, 0, taps, PAD_ZERO,
                       Cout * Cin, Cin, 1)]


# ---------------------------------------------------------------------------------------------------
# raw launches
# ---------------------------------------------------------------------------------------------------
# optional per-launch CUDA-event timing of the convolution passes (bench.py's roofline leg; off by default)
_TIMING = None
_FLOP_RATIO = 1.0     # algorithmic / launched FLOPs of the launches being recorded (channel-padded or packed thin layers)


_EVENT_EXTERNAL = False


def start_conv_timing(external=False):
    """external=True: the events become record nodes of a CUDA graph being captured (timed after the graph's replay: the
    kernels then run back to back, with no host launch gaps inside the event pairs)"""
    global _TIMING, _EVENT_EXTERNAL
    _TIMING = []
    _EVENT_EXTERNAL = bool(external)


def pause_conv_timing():
    """stop recording (keeps what was recorded for stop_conv_timing)"""
    global _TIMING, _TIMING_HELD
    _TIMING_HELD, _TIMING = _TIMING, None


_TIMING_HELD = None


def stop_conv_timing():
    """-> list of (kind, tensor_core?, flops, milliseconds) for every convolution pass since start_conv_timing()"""
    global _TIMING, _TIMING_HELD
    if _TIMING is None and _TIMING_HELD is not None:
        _TIMING, _TIMING_HELD = _TIMING_HELD, None
    rec, _TIMING = _TIMING, None
    torch.cuda.synchronize()
    return [(kind, umma, flops, e0.elapsed_time(e1), shape) for kind, umma, flops, e0, e1, shape in rec]


def _timed(kind, P, launch, umma=None):
    if _TIMING is None:
        launch()
        return
    lib = _lib.lib()
    e0 = torch.cuda.Event(enable_timing=True, external=_EVENT_EXTERNAL)
    e1 = torch.cuda.Event(enable_timing=True, external=_EVENT_EXTERNAL)
    n0 = lib.p2p_conv_umma_launches()
    e0.record()
    launch()
    e1.record()
    flops = 2.0 * P.N * P.gy * P.gx * P.J * P.ntaps * P.K * _FLOP_RATIO
    shape = dict(M=P.N * P.gy * P.gx, J=P.J, K=P.K, taps=P.ntaps, H=P.H, W=P.W, s=P.sy, os=P.oys)
    _TIMING.append((kind + ("-window" if umma else ""), bool(umma) or lib.p2p_conv_umma_launches() > n0, flops, e0, e1, shape))


def _run_wgrad(P, x, dout, dw):
    _timed("wgrad", P, lambda: _lib.check(_lib.lib().p2p_conv_wgrad(ctypes.byref(P), _p(x), _p(dout), _p(dw), _DT[x.dtype], _stream())))


def _bias_grad(dout, J):
    db = torch.empty(J, dtype=torch.float32, device=dout.device)
    _lib.check(_lib.lib().p2p_bias_grad(_p(dout), _p(db), dout.numel() // J, J, _DT[dout.dtype], _stream()))
    return db


def _weight_for(x, w32, w16):
    """physical [KH,KW,Cout,Cin] weight in the activation dtype"""
    if x.dtype == torch.float32:
        return w32
    if w16 is None:
        raise RuntimeError("bf16 activations need the bf16 weight shadow (call sync_shadows())")
    return w16


# ---------------------------------------------------------------------------------------------------
# layer implementations (plain functions on NHWC tensors; the autograd Functions below compose them)
# ---------------------------------------------------------------------------------------------------
ACT_NONE, ACT_RELU, ACT_LRELU = 0, 1, 2


def _run_pass(P, x, w, bias, out, kind="fwd"):
    _timed(kind, P, lambda: _lib.check(_lib.lib().p2p_conv_forward(ctypes.byref(P), _p(x), _p(w), _p(bias), _p(out),
                                                                   _DT[x.dtype], _DT[out.dtype], _stream())))


class ConvCfg:
    """static description of one conv / transposed-conv layer call"""
    __slots__ = ("transposed", "stride", "pad", "pad_mode", "out_pad", "tswap")

    def __init__(self, transposed=False, stride=1, pad=0, pad_mode=PAD_ZERO, out_pad=0, tswap=False):
        self.transposed, self.stride, self.pad, self.pad_mode, self.out_pad = transposed, stride, pad, pad_mode, out_pad
        # tswap: the layer runs on the spatially TRANSPOSED image (conv(x^T, w^T) = conv(x, w)^T for square kernels with
        # symmetric stride / padding); its weight is stored [KW][KH][Cout][Cin], so the kernels see an ordinary layer
        self.tswap = tswap


def _phys(weight, cfg):
    """logical weight -> physical [KH][KW][Cout][Cin] view, (Cout, Cin, KH, KW)"""
    if cfg.transposed:
        Cin, Cout, KH, KW = weight.shape
        w = weight.permute(2, 3, 1, 0)
    elif cfg.tswap:
        Cout, Cin, KH, KW = weight.shape
        assert KH == KW
        w = weight.permute(3, 2, 0, 1)
    else:
        Cout, Cin, KH, KW = weight.shape
        w = weight.permute(2, 3, 0, 1)
    assert w.is_contiguous(), "conv weights must be stored [KH][KW][Cout][Cin]"
    return w, Cout, Cin, KH, KW


def _dw_logical(dwp, cfg):
    """physical weight gradient [K0][K1][Cout][Cin] -> the logical [Cout, Cin, KH, KW] view autograd expects"""
    if dwp is None:
        return None
    return dwp.permute(2, 3, 1, 0) if cfg.tswap else dwp.permute(2, 3, 0, 1)


def _ceil16(c):
    return (c + 15) // 16 * 16


def _pad_last(t, c):
    """zero-pad the last (channel) dimension to c"""
    if t.shape[-1] == c:
        return t
    out = torch.zeros(t.shape[:-1] + (c,), dtype=t.dtype, device=t.device)
    out[..., :t.shape[-1]] = t
    return out


def _thin(x, Cin, Cout):
    """bf16 layers with 1-4 channels on one side (the 7x7 edge layers, D's first and logit layers) are run on the
    tensor cores with that side zero-padded to 16 channels -- (padded Cin, padded Cout) or None"""
    if x.dtype != torch.bfloat16 or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return None
    if Cin % 16 == 0 and Cout % 16 == 0:
        return None
    return _ceil16(Cin), _ceil16(Cout)


def _pad_weight(wq, Cout_p, Cin_p):
    KH, KW, Cout, Cin = wq.shape
    if (Cout, Cin) == (Cout_p, Cin_p):
        return wq
    out = torch.zeros((KH, KW, Cout_p, Cin_p), dtype=wq.dtype, device=wq.device)
    out[:, :, :Cout, :Cin] = wq
    return out


# ---- stride-1 layers: TMA-fed shifted-window kernel (csrc/conv_window.cu) ------------------------------------------
USE_WINDOW = True
_WIN_ROWS_MAX = 312


class _Shape:
    """what _timed needs to count FLOPs for a window launch"""

    def __init__(self, N, OH, OW, J, taps, K, H, W):
        self.N, self.gy, self.gx, self.J, self.ntaps, self.K, self.H, self.W = N, OH, OW, J, taps, K, H, W
        self.sy = self.oys = 1


def _window_ok(dtype, C, J, KH, KW, Hp, Wp, mn_major):
    if not USE_WINDOW or dtype != torch.bfloat16 or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return False
    if C % 16 != 0 or (mn_major and J % 8 != 0) or Hp < KH or Wp < KW:
        return False
    if 128 + (KH - 1) * Wp + (KW - 1) > _WIN_ROWS_MAX:
        return False
    return (Hp - KH + 1) * (Wp - KW + 1) >= 0.5 * Hp * Wp        # rows in the padding margin are computed and dropped


def pad_nhwc(x, pt, pb, pl, pr, reflect):
    N, H, W, C = x.shape
    xp = torch.empty((N, H + pt + pb, W + pl + pr, C), dtype=x.dtype, device=x.device)
    _lib.check(_lib.lib().p2p_pad_nhwc(_p(x), _p(xp), N, H, W, C, pt, pb, pl, pr, int(reflect), _DT[x.dtype], _stream()))
    return xp


def _plan_window_passes(passes, dtype):
    """common window geometry of the zero-padded passes of one layer call (gather stride 1 or 2; up to 4 output
    parities), or None when they do not fit the shifted-window kernel"""
    P0 = passes[0]
    if dtype != torch.bfloat16 or len(passes) > 4 or P0.sy != P0.sx or P0.sy not in (1, 2) or P0.K % 16 != 0:
        return None
    mn_major = P0.w_j == 1
    if not ((P0.w_k == 1 and P0.w_j % 8 == 0) or (mn_major and P0.w_k % 8 == 0 and P0.J % 8 == 0)) or P0.w_tap % 8 != 0:
        return None
    s = P0.sy
    for P in passes:
        if (P.pad_mode != PAD_ZERO or (P.sy, P.sx) != (s, s) or (P.oys, P.oxs) != (P0.oys, P0.oxs)
                or (P.w_tap, P.w_k, P.w_j, P.K, P.J, P.OH, P.OW, P.H, P.W) !=
                (P0.w_tap, P0.w_k, P0.w_j, P0.K, P0.J, P0.OH, P0.OW, P0.H, P0.W)):
            return None
    pt = max(0, -min(P.dh[t] for P in passes for t in range(P.ntaps)))
    pl = max(0, -min(P.dw[t] for P in passes for t in range(P.ntaps)))
    subs, Hh, Wh = [], 0, 0
    for P in passes:
        taps = []
        for t in range(P.ntaps):
            a, b = P.dh[t] + pt, P.dw[t] + pl
            taps.append(((a % s) * s + (b % s), a // s, b // s, P.wtap[t]))
        taps.sort(key=lambda q: q[0])
        subs.append(taps)
        Hh, Wh = max(Hh, P.gy + max(q[1] for q in taps)), max(Wh, P.gx + max(q[2] for q in taps))
    halo = max(q[1] * Wh + q[2] for taps in subs for q in taps)
    useful = sum(P.gy * P.gx for P in passes)
    if 128 + halo > _WIN_ROWS_MAX or useful < 0.5 * len(passes) * Hh * Wh or sum(len(t) for t in subs) > _lib.MAX_TAPS:
        return None
    return dict(s=s, pt=pt, pl=pl, Hh=Hh, Wh=Wh, subs=subs)


def _planes_key(pl, N, C):
    """(pt, pl, Hh, Wh, rows, reflect, s, C): the layout of a layer's padded input planes (what _window_wgrad compares)"""
    return (pl["pt"], pl["pl"], pl["Hh"], pl["Wh"], _rows64(N * pl["Hh"] * pl["Wh"]), False, pl["s"], C)


def _window_passes(passes, x, wq, bias, out, wtaps_total, kind, aux=None, xin=None, stats=None):
    """run all passes of a layer call as ONE launch of the shifted-window kernel; False (nothing launched) if they do not fit.
    xin: (key, planes) written by the producer of x (skips the padding copy); aux: list that receives (key, planes) for the
    weight-gradient pass; stats: [N,J,2] fp32 buffer for the epilogue's InstanceNorm sums (returns "stats" when they were taken)"""
    if not USE_WINDOW or not passes or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return False
    pl = _plan_window_passes(passes, x.dtype)
    if pl is None:
        return False
    N, H, W, C = x.shape
    s, P0 = pl["s"], passes[0]
    key = _planes_key(pl, N, C)
    rows = key[4]
    if xin is not None and xin[0] == key:
        planes = xin[1]
    else:
        x = x.contiguous()
        planes = torch.empty((s * s, rows, C), dtype=x.dtype, device=x.device)
        _lib.check(_lib.lib().p2p_pad_planes(_p(x), _p(planes), N, H, W, C, pl["pt"], pl["pl"], pl["Hh"], pl["Wh"], s,
                                             rows, 0, _DT[x.dtype], _stream()))
    if aux is not None:
        aux.append((key, planes))
    d = _lib.WindowDesc()
    d.nplanes, d.N, d.Hp, d.Wp, d.C = s * s, N, pl["Hh"], pl["Wh"], C
    d.plane_rows = rows
    d.OHf, d.OWf, d.J = P0.OH, P0.OW, P0.J
    d.wtaps_total, d.w_tap, d.w_k, d.w_j = wtaps_total, P0.w_tap, P0.w_k, P0.w_j
    d.nsub, d.oys, d.oxs = len(passes), P0.oys, P0.oxs
    t = 0
    for u, (P, taps) in enumerate(zip(passes, pl["subs"])):
        d.gy[u], d.gx[u], d.oyo[u], d.oxo[u] = P.gy, P.gx, P.oyo, P.oxo
        for (plane, dh, dw, wt) in taps:
            d.tap_sub[t], d.tap_plane[t], d.tap_dh[t], d.tap_dw[t], d.tap_w[t] = u, plane, dh, dw, wt
            t += 1
    d.ntaps = t
    shape = _Shape(N, 1, sum(P.gy * P.gx * P.ntaps for P in passes), P0.J, 1, C, pl["Hh"], pl["Wh"])
    run = out
    few = _few_tiles(N * pl["Hh"] * pl["Wh"], P0.J, len(passes), (C + 63) // 64 * s * s)
    if out.dtype != torch.float32 and few:
        run = torch.empty(out.shape, dtype=torch.float32, device=out.device)
    take = stats is not None and not few and run.dtype == torch.float32 and P0.J % 4 == 0 and not deterministic()
    _timed(kind, shape, lambda: _lib.check(_lib.lib().p2p_conv_window_run_stats(ctypes.byref(d), _p(planes), _p(wq), _p(bias), _p(run),
                                                                                _DT[run.dtype], _p(stats) if take else None,
                                                                                _stream())), umma=True)
    if run is not out:
        out.copy_(run)
    return "stats" if take else True


def deterministic():
    return _lib.lib().p2p_set_deterministic(-1) == 1


_WRAP_BUFFERS = {}

# ---- weight gradients on a side stream -----------------------------------------------------------------------------------------
# Within one layer's backward the input gradient feeds the next layer down, the weight gradient feeds nothing until the
# optimiser runs.  train.train_step therefore lets the window weight-gradient kernels (tensor-core bound) run on a side
# stream while the main stream goes on with the next layer's InstanceNorm backward (memory bound) and input gradient:
#   wgrad_overlap_begin() ... loss.backward() ... wgrad_overlap_join()
# Everything a side-stream launch reads is kept alive until the join (the caching allocator must not hand the memory to a
# main-stream tensor while the side stream still reads it -- also at graph-capture time), and the shared gradient-grid
# buffers become a ring whose entries are re-used only after the weight-gradient pass that read them (event).
WGRAD_OVERLAP = os.environ.get("P2P_WGRAD_OVERLAP", "1") == "1"


class _SideWgrad:
    active = False
    streams = {}
    used = set()        # devices whose side stream took work since wgrad_overlap_begin()
    keep = []


def wgrad_overlap_begin():
    _SideWgrad.active = WGRAD_OVERLAP
    _SideWgrad.used = set()
    # nothing is in flight on the side stream here (the last join made the main stream wait for it), and an event of an earlier
    # CUDA-graph capture must not be waited on inside another one: forget the ring's events
    for key, ring in _WRAP_BUFFERS.items():
        if isinstance(key, tuple) and key and key[0] == "ring":
            for slot in ring["slots"]:
                slot[1] = None


def wgrad_overlap_join():
    if _SideWgrad.active:
        for dev in _SideWgrad.used:       # (a side stream that took no work is not part of a running capture: do not touch it)
            torch.cuda.current_stream(dev).wait_stream(_SideWgrad.streams[dev])
    _SideWgrad.active = False
    _SideWgrad.used = set()
    _SideWgrad.keep = []


def _flat_tensors(objs):
    for o in objs:
        if torch.is_tensor(o):
            yield o
        elif isinstance(o, (tuple, list)):
            yield from _flat_tensors(o)


class _wgrad_side:
    """context: launches inside go to the weight-gradient side stream (after everything already enqueued on the current
    stream); `reads` are kept alive until wgrad_overlap_join(); `ring` entries get the event that frees them for re-use"""

    def __init__(self, enable, reads=(), ring=None):
        self.on = bool(enable) and _SideWgrad.active
        self.reads, self.ring, self.ctx = reads, ring, None

    def __enter__(self):
        if not self.on:
            return self
        dev = torch.cuda.current_device()
        side = _SideWgrad.streams.get(dev)
        if side is None:
            side = _SideWgrad.streams[dev] = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        _SideWgrad.used.add(dev)
        _SideWgrad.keep.extend(_flat_tensors(self.reads))
        self.side = side
        self.ctx = torch.cuda.stream(side)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.on:
            if self.ring is not None:
                ev = torch.cuda.Event()
                ev.record(self.side)
                self.ring[1] = ev
            self.ctx.__exit__(*exc)
        return False



def _wrap_buffer(device, dtype, N, Hh, Wh, KH, KW, J):
    """the shared gradient-grid buffer of one geometry: P64 zero rows | rows64(N*Hh*Wh) grid rows | P64 zero rows, P64 >=
    (KH-1)*Wh + KW-1 -> (buffer, P64, rows, ring entry or None)"""
    rows = _rows64(N * Hh * Wh)
    P64 = _rows64((KH - 1) * Wh + KW - 1)
    capturing = torch.cuda.is_current_stream_capturing()
    if _SideWgrad.active:
        # three buffers per geometry, used in turn; an entry is re-used only after the side-stream pass that read it
        key = ("ring", device, dtype, P64, rows, J, capturing)
        ring = _WRAP_BUFFERS.get(key)
        if ring is None:
            ring = _WRAP_BUFFERS[key] = {"next": 0, "slots": [[torch.zeros((P64 + rows + P64, J), dtype=dtype, device=device), None]
                                                              for _ in range(3)]}
        slot = ring["slots"][ring["next"]]
        ring["next"] = (ring["next"] + 1) % len(ring["slots"])
        if slot[1] is not None:
            torch.cuda.current_stream().wait_event(slot[1])
            slot[1] = None
        return slot[0], P64, rows, slot
    key = (device, dtype, P64, rows, J)
    buf = _WRAP_BUFFERS.get(key)
    if buf is None or capturing != _WRAP_BUFFERS.get(("capturing", key), False):
        # zero prefix | dy grid | zero suffix; the margins are written once, the grid (and its zero tail) by every call.
        # The buffer is reused by later calls of the same geometry: the two gradient passes that read it are enqueued on the
        # same stream before the next layer overwrites it.  (One buffer per capture state: a CUDA graph keeps its own.)
        buf = torch.zeros((P64 + rows + P64, J), dtype=dtype, device=device)
        _WRAP_BUFFERS[key] = buf
        _WRAP_BUFFERS[("capturing", key)] = capturing
    return buf, P64, rows, None


def _dy_wrap(dy, Hh, Wh, KH, KW):
    """dy [N,OH,OW,J] laid out at the pitch of the padded input ([N,Hh,Wh] grid, dy at its top-left, zeros elsewhere) in a
    buffer of P64 + rows64 + P64 rows whose first and last P64 >= (KH-1)*Wh + KW-1 rows are zero -> (key, buffer, P64)"""
    N, OH, OW, J = dy.shape
    buf, P64, rows, slot = _wrap_buffer(dy.device, dy.dtype, N, Hh, Wh, KH, KW, J)
    _lib.check(_lib.lib().p2p_pad_planes(_p(dy), _p(buf[P64:]), N, OH, OW, J, 0, 0, Hh, Wh, 1, rows, 0, _DT[dy.dtype], _stream()))
    return (Hh, Wh, rows, J), buf, P64, slot


def _dy_wrap_plan(x, weight, cfg, need_dx, need_dw, need_db):
    """(Hh, Wh, KH, KW, Cout) when _conv_backward will take its shared-dy branch for this layer call (stride-1 window input
    gradient + window weight gradient from ONE copy of dy at the padded input's pitch), else None.  The producer of dy
    (norm_act_backward_wrap) then writes that copy itself."""
    if not (FUSE_BWD_WRAP and need_dx) or need_db or cfg.transposed or cfg.stride != 1 or not USE_WINDOW:
        return None
    if x.dtype != torch.bfloat16:
        return None
    _, Cout, Cin, KH, KW = _phys(weight, cfg)
    if _colable(x, cfg, Cin, Cout, KH, KW) is not None or _packable(x, cfg, Cin, Cout, KH, KW) is not None:
        return None
    if _thin(x, Cin, Cout) is not None or Cout % 4 != 0:
        return None
    N, H, W, _ = x.shape
    p = cfg.pad
    q = (KH - 1) if cfg.pad_mode == PAD_REFLECT else (KH - 1 - p)
    if not (KH == KW and q >= 0 and p > 0 and _window_ok(x.dtype, Cout, Cin, KH, KW, H + 2 * p, W + 2 * p, True)):
        return None
    return H + 2 * p, W + 2 * p, KH, KW, Cout


def _dgrad_wrap(dy_shared, wq, dtype, N, H, W, p, KH, KW, Cout, Cin, reflect):
    """input gradient of a stride-1 convolution from the shared dy copy: output on the padded grid (reflection: folded by the
    caller) or directly [N,H,W,Cin] (zero padding)"""
    (Hh, Wh, rows, J), buf, P64 = dy_shared[:3]
    P = (KH - 1) * Wh + (KW - 1)
    off = P64 - P + (0 if reflect else p * Wh + p)
    d = _lib.WindowDesc()
    d.nplanes, d.N, d.Hp, d.Wp, d.C = 1, N, Hh, Wh, Cout
    gy, gx = (Hh, Wh) if reflect else (H, W)
    d.OHf, d.OWf, d.J = gy, gx, Cin
    d.wtaps_total, d.w_tap, d.w_k, d.w_j = KH * KW, Cout * Cin, Cin, 1
    d.nsub, d.oys, d.oxs, d.wrap = 1, 1, 1, 1
    d.gy[0], d.gx[0], d.oyo[0], d.oxo[0] = gy, gx, 0, 0
    t = 0
    for a in range(KH):
        for b in range(KW):
            d.tap_sub[t], d.tap_plane[t], d.tap_dh[t], d.tap_dw[t] = 0, 0, a, b
            d.tap_w[t] = (KH - 1 - a) * KW + (KW - 1 - b)
            t += 1
    d.ntaps = t
    few = _few_tiles(N * Hh * Wh, Cin, 1, (Cout + 63) // 64)
    out = torch.empty((N, gy, gx, Cin), dtype=torch.float32 if few else dtype, device=buf.device)
    planes = buf[off:]
    _timed("dgrad", _Shape(N, H, W, Cin, KH * KW, Cout, Hh, Wh),          # algorithmic FLOPs: the unpadded H x W, not the padded grid
           lambda: _lib.check(_lib.lib().p2p_conv_window_run(ctypes.byref(d), _p(planes), _p(wq), None, _p(out), _DT[out.dtype],
                                                            _stream())), umma=True)
    # reflection: the caller's fold kernel reads the fp32 split-K result directly (no conversion pass)
    return out if (out.dtype == dtype or reflect) else out.to(dtype)


def _window_wgrad(passes, x, dy, dwp, wtaps_total, xpad=None, fresh=False, dy_shared=None, dy_planes=None):
    """weight gradient of a layer call (its forward-type passes) as ONE launch of the window wgrad kernel:
    dW_t = sum over the padded grid of dout (laid out at the input's pitch) x shifted input window.  False if it does not fit."""
    if not USE_WINDOW or not passes or x.dtype != torch.bfloat16 or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return False
    P0 = passes[0]
    s_in, s_out = P0.sy, P0.oys
    K, J = P0.K, P0.J
    if (P0.sx != s_in or P0.oxs != s_out or s_in not in (1, 2) or s_out not in (1, 2) or (s_in == 2 and s_out == 2)
            or K % 16 != 0 or J % 8 != 0 or P0.w_k != 1 or P0.w_j != K or P0.w_tap != J * K):
        return False
    reflect = P0.pad_mode == PAD_REFLECT
    if reflect and s_in != 1:
        return False
    ntaps = sum(P.ntaps for P in passes)
    if ntaps > 32:
        return False
    N, H, W, C = x.shape
    pt = max(0, -min(P.dh[t] for P in passes for t in range(P.ntaps)))
    pl = max(0, -min(P.dw[t] for P in passes for t in range(P.ntaps)))
    taps, Hh, Wh = [], 0, 0
    for P in passes:
        if (P.sy, P.sx, P.oys, P.oxs, P.pad_mode, P.K, P.J) != (s_in, s_in, s_out, s_out, P0.pad_mode, K, J):
            return False
        ap = P.oyo * s_out + P.oxo
        for t in range(P.ntaps):
            a, b = P.dh[t] + pt, P.dw[t] + pl
            taps.append((ap, (a % s_in) * s_in + (b % s_in), a // s_in, b // s_in, P.wtap[t]))
            Hh, Wh = max(Hh, P.gy + a // s_in), max(Wh, P.gx + b // s_in)
    if reflect:                      # reflection needs the margin on both sides to be smaller than the input
        Hh, Wh = max(Hh, H + 2 * pt), max(Wh, W + 2 * pl)
        if pt >= H or pl >= W:
            return False
    taps.sort(key=lambda q: (q[0], q[1]))
    halo = max(q[2] * Wh + q[3] for q in taps)
    useful = sum(P.gy * P.gx for P in passes) / float(len(passes))
    if 64 + halo > _WIN_ROWS_MAX or useful < 0.5 * Hh * Wh:
        return False
    rows = (N * Hh * Wh + 63) // 64 * 64
    lib = _lib.lib()
    if xpad is not None and xpad[0] == (pt, pl, Hh, Wh, rows, reflect, s_in, C):
        bplanes = xpad[1]                   # the forward pass's padded copy
    else:
        bplanes = torch.empty((s_in * s_in, rows, C), dtype=x.dtype, device=x.device)
        x = x.contiguous()
        _lib.check(lib.p2p_pad_planes(_p(x), _p(bplanes), N, H, W, C, pt, pl, Hh, Wh, s_in, rows, int(reflect), _DT[x.dtype], _stream()))
    if dy_shared is not None and s_out == 1 and dy_shared[0] == (Hh, Wh, rows, J):
        aplanes = dy_shared[1][dy_shared[2]: dy_shared[2] + rows]       # the copy the input-gradient pass already made
    elif dy_planes is not None and dy_planes[0] == (0, 0, Hh, Wh, rows, False, s_out, J):
        aplanes = dy_planes[1]              # stride-2 layers: the planes the input-gradient window pass padded dy into
    else:
        aplanes = torch.empty((s_out * s_out, rows, J), dtype=dy.dtype, device=dy.device)
        dy = dy.contiguous()
        _lib.check(lib.p2p_pad_planes(_p(dy), _p(aplanes), N, dy.shape[1], dy.shape[2], J, 0, 0, Hh, Wh, s_out, rows, 0, _DT[dy.dtype], _stream()))
    IntArr = ctypes.c_int * len(taps)
    ta, tb, th, tw, twt = (IntArr(*[q[i] for q in taps]) for i in range(5))
    shape = _Shape(N, 1, sum(P.gy * P.gx * P.ntaps for P in passes), J, 1, K, Hh, Wh)
    _timed("wgrad", shape, lambda: _lib.check(lib.p2p_conv_wgrad_window(
        _p(aplanes), s_out * s_out, _p(bplanes), s_in * s_in, rows, Wh, J, K, _p(dwp), wtaps_total, len(taps), ta, tb, th, tw, twt,
        int(fresh), _stream())), umma=True)
    return True


def _few_tiles(rows, J, nsub, chunks):
    """the window kernel splits its channel loop across CTAs (fp32 red.add output) when the tiles cannot fill the chip"""
    return ((rows + 127) // 128) * ((J + 255) // 256) * nsub * 2 <= 148 and chunks >= 4


def conv_window(xp, wq, bias, out_dtype, KH, KW, J, flip, w_tap, w_k, w_j, kind, shape=None, stats=None):
    """stats: optional [N,J,2] fp32 buffer; when the launch can take InstanceNorm's sums in its epilogue it does and the call
    returns (out, True), else (out, False).  Without `stats` only `out` is returned."""
    N, Hp, Wp, C = shape if shape is not None else xp.shape
    OH, OW = Hp - KH + 1, Wp - KW + 1
    few = _few_tiles(N * Hp * Wp, J, 1, (C + 63) // 64)
    run_dtype = torch.float32 if few else out_dtype
    out = torch.empty((N, OH, OW, J), dtype=run_dtype, device=xp.device)
    take = stats is not None and not few and run_dtype == torch.float32 and J % 4 == 0 and not deterministic()
    _timed(kind, _Shape(N, OH, OW, J, KH * KW, C, Hp, Wp),
           lambda: _lib.check(_lib.lib().p2p_conv_window_stats(_p(xp), _p(wq), _p(bias), _p(out), N, Hp, Wp, C, J, KH, KW, int(flip),
                                                               w_tap, w_k, w_j, _DT[run_dtype], _p(stats) if take else None,
                                                               _stream())), umma=True)
    out = out if run_dtype == out_dtype else out.to(out_dtype)
    return out if stats is None else (out, take)


# ---- thin layers (2-4 channels on one side): fold the kernel's columns into the channel dimension ---------------------
# A KHxKW convolution whose thin side has c channels with KW*c <= 16 becomes a KHx1 convolution on a tensor with
# (dw, c) packed into <= 16 channels: KW x fewer taps, no zero-padded channels multiplied on the tensor cores.
#   thin input  (G's 7x7 2->ngf layers, D's 4x4 s2 4->ndf layer):   x_pack[n, y, ox, (dw,c)] = x_pad[n, y, ox*s + dw, c]
#   thin output (G's 7x7 ngf->2 layer, D's 4x4 ->1 logit layer):    out_pack[n, y, x', (dw,co)], summed over dw afterwards
def _packable(x, cfg, Cin, Cout, KH, KW):
    if x.dtype != torch.bfloat16 or cfg.transposed or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return None
    if Cin % 16 != 0 and Cin * KW <= 16 and Cout % 16 == 0:
        return "in"
    if Cout % 16 != 0 and Cout * KW <= 16 and Cin % 16 == 0 and cfg.stride == 1:
        return "out"
    return None


def _pad_hw(x, p, reflect):
    """[N,H,W,C] -> [N,H+2p,W+2p,C] for any channel count (torch data movement; p2p_pad_nhwc needs 16-byte pixels)"""
    if p == 0:
        return x
    N, H, W, C = x.shape
    if (C * x.element_size()) % 16 == 0:
        return pad_nhwc(x, p, p, p, p, reflect)
    xc = x.permute(0, 3, 1, 2)
    xc = torch.nn.functional.pad(xc.float(), (p, p, p, p), mode="reflect" if reflect else "constant").to(x.dtype)
    return xc.permute(0, 2, 3, 1).contiguous()


def _packed_in_operands(x, wq, cfg, Cin, Cout, KH, KW, xpk=None):
    N, H, W, _ = x.shape
    s, p = cfg.stride, cfg.pad
    Hp, Wp = H + 2 * p, W + 2 * p
    OH, OW = (Hp - KH) // s + 1, (Wp - KW) // s + 1
    if xpk is None:
        xpk = torch.empty((N, Hp, OW, 16), dtype=x.dtype, device=x.device)
        _lib.check(_lib.lib().p2p_pack_in(_p(x), _p(xpk), N, H, W, Cin, KW, s, p, int(cfg.pad_mode == PAD_REFLECT), OW,
                                          _DT[x.dtype], _stream()))
    P = _make_pass(N, Hp, OW, 16, OH, OW, Cout, OH, OW, s, 1, 1, 0, 1, 0, [(dh, 0, dh) for dh in range(KH)], PAD_ZERO,
                   Cout * 16, 1, 16)
    wpk = None
    if wq is not None:      # [KH][KW][Cout][Cin] -> [KH][Cout][(dw, c) padded to 16]
        wpk = _pad_last(wq.permute(0, 2, 1, 3).reshape(KH, Cout, KW * Cin), 16).contiguous()
    return xpk, wpk, P, (OH, OW)


def _packed_out_geometry(x, cfg, Cin, Cout, KH, KW):
    N, H, W, _ = x.shape
    p = cfg.pad
    Hp, Wp = H + 2 * p, W + 2 * p
    OH, OW = Hp - KH + 1, Wp - KW + 1
    # forward / wgrad pass: rows (y, x') over [OH, Wp], gather xp[y + dh, x'], 16 packed output channels
    P = _make_pass(N, Hp, Wp, Cin, OH, Wp, 16, OH, Wp, 1, 1, 1, 0, 1, 0, [(dh, 0, dh) for dh in range(KH)], PAD_ZERO,
                   16 * Cin, 1, Cin)
    return P, (Hp, Wp, OH, OW)


def _pack_out_weight(wq, KH, KW, Cout, Cin):
    return _pad_rows(wq.reshape(KH, KW * Cout, Cin), 16)


def _pad_rows(t, r):
    if t.shape[1] == r:
        return t.contiguous()
    out = torch.zeros((t.shape[0], r, t.shape[2]), dtype=t.dtype, device=t.device)
    out[:, :t.shape[1]] = t
    return out


def _pack_dy(dy, KW, Cout, Wp):
    """dy [N,OH,OW,Cout] -> dy_pack[n, y, x', (dw,co)] = dy[n, y, x' - dw, co]"""
    N, OH, OW, _ = dy.shape
    out = torch.empty((N, OH, Wp, 16), dtype=dy.dtype, device=dy.device)
    _lib.check(_lib.lib().p2p_pack_dy(_p(dy), _p(out), N, OH, OW, Cout, KW, Wp, _DT[dy.dtype], _stream()))
    return out


ROW_PACKED = os.environ.get("P2P_ROW_PACKED", "1") == "1"


def _packed_forward(kind, x, wq, bias, cfg, Cin, Cout, KH, KW, out_dtype, aux=None, xin=None):
    if kind == "in":
        xpk, wpk, P, (OH, OW) = _packed_in_operands(x, wq, cfg, Cin, Cout, KH, KW)
        if aux is not None:
            aux.append(("packed_in", xpk))
        y = torch.empty((x.shape[0], OH, OW, Cout), dtype=out_dtype, device=x.device)
        _run_pass(P, xpk, wpk, bias, y)
        return y
    N = x.shape[0]
    P, (Hp, Wp, OH, OW) = _packed_out_geometry(x, cfg, Cin, Cout, KH, KW)
    if xin is not None:
        xp = xin[1].view(-1, Cin)           # written in this layout by the producer (plan_input_layout)
    else:
        xp = _pad_rows64(x, cfg.pad, cfg.pad_mode == PAD_REFLECT) if (Cin * x.element_size()) % 16 == 0 else \
            _pad_hw(x, cfg.pad, cfg.pad_mode == PAD_REFLECT)
    if aux is not None:
        aux.append(("packed_out", xp))
    if (ROW_PACKED and KH * Cout <= 16 and (Cin * x.element_size()) % 16 == 0
            and _window_ok(x.dtype, Cin, 16, 1, KW, Hp, Wp, False)):
        # kernel ROWS packed into the 16 output channels instead of its columns: the KW taps then shift along a pixel row
        # (halo KW-1 rows instead of (KH-1)*Wp), which the TMA-fed window kernel runs; the gather-form kernel below took
        # 168 us for G's last 7x7 layer (29 TFLOP/s)
        wr = _pad_rows(wq.permute(1, 0, 2, 3).reshape(KW, KH * Cout, Cin), 16)
        yr = conv_window(xp, wr, None, torch.float32, 1, KW, 16, False, 16 * Cin, 1, Cin, "fwd", (N, Hp, Wp, Cin))
        y = torch.empty((N, OH, OW, Cout), dtype=out_dtype, device=x.device)
        _lib.check(_lib.lib().p2p_fold_out_rows(_p(yr), _p(bias), _p(y), N, Hp, OW, OH, OW, KH, Cout, _DT[out_dtype], _stream()))
        return y
    ypk = torch.empty((N, OH, Wp, 16), dtype=torch.float32, device=x.device)
    _run_pass(P, xp, _pack_out_weight(wq, KH, KW, Cout, Cin), None, ypk)
    y = torch.empty((N, OH, OW, Cout), dtype=out_dtype, device=x.device)
    _lib.check(_lib.lib().p2p_fold_out(_p(ypk), _p(bias), _p(y), N, OH, Wp, OW, KW, Cout, _DT[out_dtype], _stream()))
    return y


def _packed_backward(kind, x, wq, dy, cfg, Cin, Cout, KH, KW, need_dx, need_dw, xpad=None):
    """-> (dx or None, dW physical [KH][KW][Cout][Cin] fp32 or None); dx only for the thin-output kind"""
    N, H, W, _ = x.shape
    dx = dwp = None
    if kind == "in":
        if need_dw:
            xpk, _, P, _ = _packed_in_operands(x, None, cfg, Cin, Cout, KH, KW,
                                               xpad[1] if xpad is not None and xpad[0] == "packed_in" else None)
            dwk = torch.zeros((KH, Cout, 16), dtype=torch.float32, device=x.device)
            _run_wgrad(P, xpk, dy, dwk)
            dwp = dwk[:, :, :KW * Cin].reshape(KH, Cout, KW, Cin).permute(0, 2, 1, 3)
        return None, dwp
    P, (Hp, Wp, OH, OW) = _packed_out_geometry(x, cfg, Cin, Cout, KH, KW)
    dypk = _pack_dy(dy, KW, Cout, Wp)
    if need_dx:
        wqk = _pack_out_weight(wq, KH, KW, Cout, Cin)
        Pd = _make_pass(N, OH, Wp, 16, Hp, Wp, Cin, Hp, Wp, 1, 1, 1, 0, 1, 0, [(-dh, 0, dh) for dh in range(KH)], PAD_ZERO,
                        16 * Cin, Cin, 1)
        buf = torch.empty((N, Hp, Wp, Cin), dtype=x.dtype, device=x.device)
        _run_pass(Pd, dypk, wqk, None, buf, "dgrad")
        p = cfg.pad
        if cfg.pad_mode == PAD_REFLECT and p > 0:
            dx = torch.empty_like(x)
            _lib.check(_lib.lib().p2p_reflect_pad_fold(_p(buf), _p(dx), N, H, W, Cin, p, _DT[x.dtype], _stream()))
        else:
            dx = buf[:, p: p + H, p: p + W, :].contiguous() if p else buf
    if need_dw:
        xp = xpad[1] if xpad is not None and xpad[0] == "packed_out" else _pad_hw(x.contiguous(), cfg.pad, cfg.pad_mode == PAD_REFLECT)
        dwk = torch.zeros((KH, 16, Cin), dtype=torch.float32, device=x.device)
        _run_wgrad(P, xp, dypk, dwk)
        dwp = dwk[:, :KW * Cout].reshape(KH, KW, Cout, Cin)
    return dx, dwp


# ---- thin layers whose whole patch fits 128 channels: 1x1 convolutions on packed operands (csrc/pack.cu) ------------------
def _colable(x, cfg, Cin, Cout, KH, KW):
    """'in': im2col of the thin input makes the layer a 1x1 convolution (forward, dW, dx);
    'out': the output gradient packed over the patch makes dW and dx 1x1 convolutions (the forward stays column-packed)"""
    if x.dtype != torch.bfloat16 or cfg.transposed or not USE_WINDOW or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return None
    if Cin % 16 != 0 and Cin * KH * KW <= 128 and Cout % 16 == 0:
        return "in"
    if Cout % 16 != 0 and Cout * KH * KW <= 128 and Cout * KW <= 16 and Cin % 16 == 0 and cfg.stride == 1:
        return "out"
    return None


def _rows64(n):
    return (n + 63) // 64 * 64


def _with_ratio(r, fn):
    global _FLOP_RATIO
    old, _FLOP_RATIO = _FLOP_RATIO, r
    try:
        return fn()
    finally:
        _FLOP_RATIO = old


def _wgrad_1x1(a2d, b2d, rows, J, C, dw, Wp):
    """dw[J][C] (fp32) += a2d[rows][J]^T b2d[rows][C]; rows % 64 == 0, tails zero"""
    one = (ctypes.c_int * 1)(0)
    shape = _Shape(1, 1, rows, J, 1, C, 1, rows)
    _timed("wgrad", shape, lambda: _lib.check(_lib.lib().p2p_conv_wgrad_window(
        _p(a2d), 1, _p(b2d), 1, rows, Wp, J, C, _p(dw), 1, 1, one, one, one, one, one, 1, _stream())), umma=True)


def _col_in_geometry(x, cfg, Cin, KH, KW):
    N, H, W, _ = x.shape
    s, p = cfg.stride, cfg.pad
    OH, OW = (H + 2 * p - KH) // s + 1, (W + 2 * p - KW) // s + 1
    return N, H, W, s, p, OH, OW, _ceil16(KH * KW * Cin)


def _col_weight(wq, Kc):
    """[KH][KW][Cout][Cin] -> [Cout][(kh,kw,c) zero-padded to Kc]"""
    KH, KW, Cout, Cin = wq.shape
    return _pad_last(wq.permute(2, 0, 1, 3).reshape(Cout, KH * KW * Cin), Kc).contiguous()


def _im2col(x, cfg, Cin, KH, KW):
    N, H, W, s, p, OH, OW, Kc = _col_in_geometry(x, cfg, Cin, KH, KW)
    rows = _rows64(N * OH * OW)
    xcol = torch.empty((rows, Kc), dtype=x.dtype, device=x.device)
    _lib.check(_lib.lib().p2p_im2col(_p(x), _p(xcol), N, H, W, Cin, KH, KW, s, p, int(cfg.pad_mode == PAD_REFLECT), OH, OW, Kc,
                                     rows, _stream()))
    return xcol


def _col_in_forward(x, wq, bias, cfg, Cin, Cout, KH, KW, out_dtype, aux, stats=None):
    """-> (y, stats_taken)"""
    N, H, W, s, p, OH, OW, Kc = _col_in_geometry(x, cfg, Cin, KH, KW)
    xcol = _im2col(x, cfg, Cin, KH, KW)
    if aux is not None:
        aux.append(("col_in", xcol))
    if stats is not None and xcol.shape[0] == N * OH * OW:      # (a zero row tail would be counted as output pixels)
        return _with_ratio(KH * KW * Cin / float(Kc),
                           lambda: conv_window(xcol, _col_weight(wq, Kc), bias, out_dtype, 1, 1, Cout, False, Cout * Kc, 1, Kc, "fwd",
                                               (N, OH, OW, Kc), stats))
    return _with_ratio(KH * KW * Cin / float(Kc),
                       lambda: conv_window(xcol, _col_weight(wq, Kc), bias, out_dtype, 1, 1, Cout, False, Cout * Kc, 1, Kc, "fwd",
                                           (N, OH, OW, Kc))), False


def _col_in_backward(x, wq, dy, cfg, Cin, Cout, KH, KW, need_dx, need_dw, xpad):
    """-> (dx or None, dW physical [KH][KW][Cout][Cin] fp32 or None); dx only with zero padding"""
    N, H, W, s, p, OH, OW, Kc = _col_in_geometry(x, cfg, Cin, KH, KW)
    ratio = KH * KW * Cin / float(Kc)
    dx = dwp = None
    if need_dw:
        xcol = xpad[1] if xpad is not None and xpad[0] == "col_in" else _im2col(x, cfg, Cin, KH, KW)
        rows = xcol.shape[0]
        if rows == N * OH * OW:
            dy2 = dy
        else:
            dy2 = torch.empty((rows, Cout), dtype=dy.dtype, device=dy.device)
            _lib.check(_lib.lib().p2p_pad_planes(_p(dy), _p(dy2), N, OH, OW, Cout, 0, 0, OH, OW, 1, rows, 0, _DT[dy.dtype], _stream()))
        dwcol = torch.zeros((Cout, Kc), dtype=torch.float32, device=x.device)
        _with_ratio(ratio, lambda: _wgrad_1x1(dy2, xcol, rows, Cout, Kc, dwcol, OW))
        dwp = dwcol[:, :KH * KW * Cin].reshape(Cout, KH, KW, Cin).permute(1, 2, 0, 3)
    if need_dx:
        dxcol = _with_ratio(ratio, lambda: conv_window(dy, _col_weight(wq, Kc), None, x.dtype, 1, 1, Kc, False, Cout * Kc, Kc, 1,
                                                       "dgrad", (N, OH, OW, Cout)))
        dx = torch.empty_like(x)
        _lib.check(_lib.lib().p2p_col2im(_p(dxcol), _p(dx), N, H, W, Cin, KH, KW, s, p, OH, OW, Kc, _stream()))
    return dx, dwp


def _col_out_backward(x, wq, dy, cfg, Cin, Cout, KH, KW, need_dx, need_dw, xpad):
    """thin-output layer (stride 1): dyfull[(n,yp,xp), (kh,kw,co)] = dy[n, yp-kh, xp-kw, co] turns dW and dx into 1x1 products"""
    N, H, W, _ = x.shape
    p = cfg.pad
    Hp, Wp = H + 2 * p, W + 2 * p
    OH, OW = Hp - KH + 1, Wp - KW + 1
    Kc = _ceil16(KH * KW * Cout)
    rows = _rows64(N * Hp * Wp)
    ratio = (KH * KW * Cout / float(Kc)) * (OH * OW / float(Hp * Wp))
    dyfull = torch.empty((rows, Kc), dtype=dy.dtype, device=dy.device)
    _lib.check(_lib.lib().p2p_pack_dy_full(_p(dy), _p(dyfull), N, OH, OW, Cout, KH, KW, Kc, rows, _stream()))
    dx = dwp = None
    if need_dx:
        wfull = _pad_rows(wq.reshape(1, KH * KW * Cout, Cin), Kc)[0]        # [Kc][Cin]
        buf = _with_ratio(ratio, lambda: conv_window(dyfull, wfull, None, x.dtype, 1, 1, Cin, False, Kc * Cin, Cin, 1, "dgrad",
                                                     (N, Hp, Wp, Kc)))
        if cfg.pad_mode == PAD_REFLECT and p > 0:
            dx = torch.empty_like(x)
            _lib.check(_lib.lib().p2p_reflect_pad_fold(_p(buf), _p(dx), N, H, W, Cin, p, _DT[x.dtype], _stream()))
        else:
            dx = buf[:, p: p + H, p: p + W, :].contiguous() if p else buf
    if need_dw:
        if xpad is not None and xpad[0] == "packed_out" and xpad[1].shape[0] == rows:
            xp = xpad[1]
        else:
            xp = _pad_rows64(x.contiguous(), p, cfg.pad_mode == PAD_REFLECT)
        dwfull = torch.zeros((Kc, Cin), dtype=torch.float32, device=x.device)
        _with_ratio(ratio, lambda: _wgrad_1x1(dyfull, xp, rows, Kc, Cin, dwfull, Wp))
        dwp = dwfull[:KH * KW * Cout].reshape(KH, KW, Cout, Cin)
    return dx, dwp


def _pad_rows64(x, p, reflect):
    """x padded by p on every side as [rows64(N*Hp*Wp)][C] with a zero tail (C * itemsize % 16 == 0)"""
    N, H, W, C = x.shape
    Hp, Wp = H + 2 * p, W + 2 * p
    rows = _rows64(N * Hp * Wp)
    xp = torch.empty((rows, C), dtype=x.dtype, device=x.device)
    _lib.check(_lib.lib().p2p_pad_planes(_p(x), _p(xp), N, H, W, C, p, p, Hp, Wp, 1, rows, int(reflect), _DT[x.dtype], _stream()))
    return xp



def plan_input_layout(shape, dtype, weight, cfg):
    """The layout (pt, pl, Hh, Wh, rows, reflect, s, C) of the padded planes conv_forward builds from an input of this shape,
    or None when the layer does not read padded planes (thin / unpadded / non-window layers).  A producer that writes its
    output in this layout (norm_act_forward_planes) saves the padding copy: pass the buffer as conv_forward(xin=(key, planes))."""
    if dtype != torch.bfloat16 or not USE_WINDOW or _lib.lib().p2p_conv_force_simple(-1) == 1:
        return None
    N, H, W, Cx = shape
    _, Cout, Cin, KH, KW = _phys(weight, cfg)
    if Cx != Cin or (Cin * 2) % 16 != 0:
        return None
    x_meta = torch.empty(0, dtype=dtype)
    if _colable(x_meta, cfg, Cin, Cout, KH, KW) == "in":
        return None
    pk = _packable(x_meta, cfg, Cin, Cout, KH, KW)
    p = cfg.pad
    if pk == "out":
        return (p, p, H + 2 * p, W + 2 * p, _rows64(N * (H + 2 * p) * (W + 2 * p)), cfg.pad_mode == PAD_REFLECT, 1, Cin) if p else None
    if pk is not None or _thin(x_meta, Cin, Cout) is not None:
        return None
    if cfg.transposed:
        passes, _ = convT_fwd_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.out_pad)
    else:
        if cfg.pad_mode == PAD_REFLECT and (p >= H or p >= W):
            return None
        if cfg.stride == 1 and _window_ok(dtype, Cin, Cout, KH, KW, H + 2 * p, W + 2 * p, False):
            return (p, p, H + 2 * p, W + 2 * p, _rows64(N * (H + 2 * p) * (W + 2 * p)), cfg.pad_mode == PAD_REFLECT, 1, Cin) if p else None
        passes = conv_fwd_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.pad_mode)
    if not passes:
        return None
    pl = _plan_window_passes(passes, dtype)
    return _planes_key(pl, N, Cin) if pl is not None else None


def _conv_forward(x, weight, bias, w16, cfg, out_dtype, aux=None, xin=None, stats=None):
    """nn.Conv2d (+ nn.ReflectionPad2d) / nn.ConvTranspose2d forward, NHWC (networks.py:145-154, 190-207, 223-246, 340-361).
    xin: (key, planes) -- the input already laid out as this layer's padded planes (plan_input_layout) by its producer;
    stats: [N,Cout,2] fp32 buffer for InstanceNorm's sums taken in the conv epilogue.  Returns (y, stats_taken)."""
    N, H, W, Cx = x.shape
    wphys, Cout, Cin, KH, KW = _phys(weight, cfg)
    assert Cx == Cin, (Cx, Cin)
    wq = _weight_for(x, wphys, w16)
    Jout = Cout
    if xin is not None and xin[0] != plan_input_layout(x.shape, x.dtype, weight, cfg):
        xin = None
    if xin is None:
        x = x.contiguous()
    if _colable(x, cfg, Cin, Cout, KH, KW) == "in":
        return _col_in_forward(x, wq, bias, cfg, Cin, Cout, KH, KW, out_dtype, aux, stats)
    pk = _packable(x, cfg, Cin, Cout, KH, KW)
    if pk is not None:
        return _packed_forward(pk, x, wq, bias, cfg, Cin, Cout, KH, KW, out_dtype, aux, xin), False
    thin = _thin(x, Cin, Cout)
    if thin is not None and thin[0] != Cin:             # the forward pass only needs the gathered side padded
        x, wq, Cin = _pad_last(x, thin[0]), _pad_weight(wq, Cout, thin[0]), thin[0]
    if cfg.transposed:
        passes, (OH, OW) = convT_fwd_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.out_pad)
        if passes is None:
            raise NotImplementedError("transposed convolution with empty output parity classes")
    else:
        if cfg.pad_mode == PAD_REFLECT and (cfg.pad >= H or cfg.pad >= W):
            raise RuntimeError("reflection padding must be smaller than the input")   # what ReflectionPad2d raises
        p = cfg.pad
        if cfg.stride == 1 and _window_ok(x.dtype, Cin, Cout, KH, KW, H + 2 * p, W + 2 * p, False):
            if p:       # padded once, in the layout the wgrad window kernel wants too (rows rounded to 64, zero tail)
                reflect = cfg.pad_mode == PAD_REFLECT
                Hp, Wp = H + 2 * p, W + 2 * p
                rows = (N * Hp * Wp + 63) // 64 * 64
                key = (p, p, Hp, Wp, rows, reflect, 1, Cin)
                if xin is not None:
                    xp = xin[1].view(rows, Cin)
                else:
                    xp = torch.empty((rows, Cin), dtype=x.dtype, device=x.device)
                    _lib.check(_lib.lib().p2p_pad_planes(_p(x), _p(xp), N, H, W, Cin, p, p, Hp, Wp, 1, rows, int(reflect),
                                                         _DT[x.dtype], _stream()))
                if aux is not None:
                    aux.append((key, xp))
            else:
                xp, Hp, Wp = x, H, W
            if stats is not None:
                return conv_window(xp, wq, bias, out_dtype, KH, KW, Jout, False, Cout * Cin, 1, Cin, "fwd", (N, Hp, Wp, Cin), stats)
            return conv_window(xp, wq, bias, out_dtype, KH, KW, Jout, False, Cout * Cin, 1, Cin, "fwd", (N, Hp, Wp, Cin)), False
        passes = conv_fwd_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.pad_mode)
        OH, OW = passes[0].OH, passes[0].OW
    y = torch.empty((N, OH, OW, Jout), dtype=out_dtype, device=x.device)
    r = _window_passes(passes, x, wq, bias, y, KH * KW, "fwd", aux if thin is None else None, xin, stats)
    if r:
        return y, r == "stats"
    x = x.contiguous()
    for P in passes:
        _run_pass(P, x, wq, bias, y)
    return y, False


def _conv_backward(x, weight, w16, dy, cfg, need_dx, need_dw, need_db, gphys=None, xpad=None, dy_pre=None):
    """gradients of conv_forward; dy has x's dtype.  Returns (dx, dw_logical, db).  With `gphys` (FusedAdam's view of
    this weight's slice of the flat gradient buffer, physical layout) the wgrad kernels accumulate straight into it
    and dw comes back None -- no zero-fill, no AccumulateGrad add."""
    N, H, W, _ = x.shape
    wphys, Cout, Cin, KH, KW = _phys(weight, cfg)
    assert dy.dtype == x.dtype and (dy.is_contiguous() or dy_pre is not None)
    if dy_pre is not None and _dy_wrap_plan(x, weight, cfg, need_dx, need_dw, need_db) is None:
        raise RuntimeError("a pre-wrapped output gradient was passed to a layer that does not take the shared-dy path")
    dx = dw = db = None
    dy_shared = None
    dy_aux = []             # (key, planes) of dy as the input-gradient window pass laid it out: reused by the weight gradient
    if need_db:
        db = _bias_grad(dy, Cout)
    col = _colable(x, cfg, Cin, Cout, KH, KW)
    if col == "in" and (not need_dx or cfg.pad_mode == PAD_ZERO or cfg.pad == 0):
        dx, dwp = _col_in_backward(x, _weight_for(x, wphys, w16) if need_dx else None, dy, cfg, Cin, Cout, KH, KW, need_dx,
                                   need_dw, xpad)
        return dx, _dw_logical(dwp, cfg), db
    if col == "out":
        dx, dwp = _col_out_backward(x, _weight_for(x, wphys, w16) if need_dx else None, dy, cfg, Cin, Cout, KH, KW, need_dx,
                                    need_dw, xpad)
        return dx, _dw_logical(dwp, cfg), db
    pk = _packable(x, cfg, Cin, Cout, KH, KW)
    if pk is not None and (pk == "out" or not need_dx):
        dx, dwp = _packed_backward(pk, x, _weight_for(x, wphys, w16) if need_dx else None, dy, cfg, Cin, Cout, KH, KW,
                                   need_dx, need_dw, xpad)
        return dx, _dw_logical(dwp, cfg), db
    pk_dw = pk == "in" and need_dw         # thin-input layer that also needs dx: dW packed, dx through the padded path
    if pk_dw:
        _, dwp = _packed_backward(pk, x, None, dy, cfg, Cin, Cout, KH, KW, False, True, xpad)
        dw = _dw_logical(dwp, cfg)
        need_dw = False
    Cin0, Cout0 = Cin, Cout
    thin = _thin(x, Cin, Cout)
    wq = _weight_for(x, wphys, w16) if (need_dx or thin is not None) else None
    if thin is not None:
        Cin, Cout = thin
        x, dy = _pad_last(x, Cin), _pad_last(dy, Cout)
        if need_dx:
            wq = _pad_weight(wq, Cout, Cin)
    if need_dx:
        if cfg.transposed:
            dx = torch.empty_like(x)
            Pd = convT_dgrad_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.out_pad)
            if not _window_passes(Pd, dy, wq, None, dx, KH * KW, "dgrad"):
                _run_pass(Pd[0], dy, wq, None, dx, "dgrad")
        else:
            q = (KH - 1) if cfg.pad_mode == PAD_REFLECT else (KH - 1 - cfg.pad)      # zero margin around dy
            OHd, OWd = dy.shape[1], dy.shape[2]
            p = cfg.pad
            if (cfg.stride == 1 and KH == KW and q >= 0 and p > 0 and thin is None and USE_WINDOW
                    and _window_ok(x.dtype, Cout, Cin, KH, KW, H + 2 * p, W + 2 * p, True)):
                # ONE copy of dy serves both gradients: dy at the pitch of the padded input (what the wgrad kernel wants)
                # behind a zero prefix; the flipped-tap window reads wrap into the zero margins of the previous row / image
                dy_shared = dy_pre if dy_pre is not None else _dy_wrap(dy, H + 2 * p, W + 2 * p, KH, KW)
                buf = _dgrad_wrap(dy_shared, wq, x.dtype, N, H, W, p, KH, KW, Cout, Cin, cfg.pad_mode == PAD_REFLECT)
                passes = []
            elif (cfg.stride == 1 and KH == KW and q >= 0
                    and _window_ok(x.dtype, Cout, Cin, KH, KW, OHd + 2 * q, OWd + 2 * q, True)):
                dyp = pad_nhwc(dy, q, q, q, q, False) if q else dy
                buf = conv_window(dyp, wq, None, x.dtype, KH, KW, Cin, True, Cout * Cin, Cin, 1, "dgrad")
                passes = []
            else:
                passes, (DH, DW) = conv_dgrad_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.pad_mode)
                if passes is None:
                    raise NotImplementedError("input gradient of a convolution whose stride skips whole parity classes")
                buf = torch.empty((N, DH, DW, Cin), dtype=x.dtype, device=x.device)
                if _window_passes(passes, dy, wq, None, buf, KH * KW, "dgrad", dy_aux):
                    passes = []
            for P in passes:
                _run_pass(P, dy, wq, None, buf, "dgrad")
            if cfg.pad_mode == PAD_REFLECT and cfg.pad > 0:
                dx = torch.empty_like(x)
                fold = _lib.lib().p2p_reflect_pad_fold_f32 if buf.dtype == torch.float32 and x.dtype != torch.float32 \
                    else _lib.lib().p2p_reflect_pad_fold
                _lib.check(fold(_p(buf), _p(dx), N, H, W, Cin, cfg.pad, _DT[x.dtype], _stream()))
            else:
                dx = buf
        if Cin != Cin0:
            dx = dx[..., :Cin0].contiguous()
    if need_dw:
        direct = gphys is not None and thin is None and tuple(gphys.view.shape) == (KH, KW, Cout, Cin)
        dwp = gphys.view if direct else torch.zeros((KH, KW, Cout, Cin), dtype=torch.float32, device=x.device)
        fresh = gphys.first_write() if direct else True
        if cfg.transposed:
            wpasses = convT_fwd_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.out_pad)[0]
        else:
            wpasses = conv_fwd_passes(N, H, W, Cin, Cout, KH, KW, cfg.stride, cfg.pad, cfg.pad_mode)
        # 2: zero_grad left this slice alone (FusedAdam lazy zeroing), the kernel must overwrite it
        must_store = direct and fresh and gphys.unzeroed_epoch == gphys.owner.grad_epoch
        # direct gradients feed nothing until the optimiser runs: their kernels go to the side stream (see _SideWgrad)
        ring = dy_shared[3] if dy_shared is not None and len(dy_shared) > 3 else None
        with _wgrad_side(direct, (x, dy, xpad, dy_shared[:3] if dy_shared is not None else None, dy_aux), ring):
            if _window_wgrad(wpasses, x, dy, dwp, KH * KW, xpad if thin is None else None, 2 if must_store else fresh, dy_shared,
                             dy_aux[0] if dy_aux else None):
                if direct and fresh and _lib.lib().p2p_conv_wgrad_window_stored() == 1:
                    gphys.stored_epoch = gphys.owner.grad_epoch
            else:
                if must_store:
                    dwp.zero_()             # zero_grad skipped this slice expecting a store: the fallback kernels accumulate
                x, dy = x.contiguous(), dy.contiguous()
                for P in wpasses:
                    _run_wgrad(P, x, dy, dwp)
        if direct:
            gphys.written()
            return dx, None, db
        if thin is not None:
            dwp = dwp[:, :, :Cout0, :Cin0]
        dw = dwp.permute(3, 2, 0, 1) if cfg.transposed else _dw_logical(dwp, cfg)
    return dx, dw, db


def _alg_ratio(x, weight, cfg):
    """algorithmic / launched FLOPs for the thin-layer paths (1 for every other layer)"""
    _, Cout, Cin, KH, KW = _phys(weight, cfg)
    if x.dtype != torch.bfloat16 or (Cin % 16 == 0 and Cout % 16 == 0):
        return 1.0
    if _colable(x, cfg, Cin, Cout, KH, KW) == "in":
        return 1.0          # set per launch by the im2col path
    pk = _packable(x, cfg, Cin, Cout, KH, KW)
    if pk == "in":
        return KW * Cin / 16.0
    if pk == "out":
        W = x.shape[2] + 2 * cfg.pad
        return (KW * Cout / 16.0) * ((W - KW + 1) / W)
    return (Cin * Cout) / float(_ceil16(Cin) * _ceil16(Cout))


def conv_forward(x, weight, bias, w16, cfg, out_dtype, aux=None, xin=None, stats=None):
    """aux: optional list; the window paths append (key, padded planes) for conv_backward(xpad=...) to reuse.
    xin / stats: see _conv_forward; with `stats` the result is (y, stats_taken), else y alone."""
    global _FLOP_RATIO
    if _TIMING is None:
        y, took = _conv_forward(x, weight, bias, w16, cfg, out_dtype, aux, xin, stats)
    else:
        _FLOP_RATIO = _alg_ratio(x, weight, cfg)
        try:
            y, took = _conv_forward(x, weight, bias, w16, cfg, out_dtype, aux, xin, stats)
        finally:
            _FLOP_RATIO = 1.0
    return (y, took) if stats is not None else y


def conv_backward(x, weight, w16, dy, cfg, need_dx, need_dw, need_db, gphys=None, xpad=None, dy_pre=None):
    global _FLOP_RATIO
    if _TIMING is None:
        return _conv_backward(x, weight, w16, dy, cfg, need_dx, need_dw, need_db, gphys, xpad, dy_pre)
    _FLOP_RATIO = _alg_ratio(x, weight, cfg)
    try:
        return _conv_backward(x, weight, w16, dy, cfg, need_dx, need_dw, need_db, gphys, xpad, dy_pre)
    finally:
        _FLOP_RATIO = 1.0


_STATS_WS = {}


def _ensure_stats_workspace(device):
    """zeroed scratch that lets the InstanceNorm reductions run as one launch each (p2p_instnorm_workspace); one per process
    (one process per GPU), created on first use and kept alive here"""
    if not USE_STATS_WS:
        return
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    if _STATS_WS.get("device") == key:
        return
    ws = torch.zeros(1 << 20, dtype=torch.float32, device=device)
    cnt = torch.zeros(1 << 16, dtype=torch.int32, device=device)
    _STATS_WS.update(device=key, ws=ws, cnt=cnt)
    _lib.check(_lib.lib().p2p_instnorm_workspace(_p(ws), ws.numel(), _p(cnt), cnt.numel()))


USE_STATS_WS = os.environ.get("P2P_STATS_WS", "1") == "1"


def norm_act_forward(y, res, norm, act, slope, eps, out_dtype):
    """z = act(InstanceNorm2d(affine=False)(y)) [+ res]; returns (z, stats)"""
    N, H, W, C = y.shape
    _ensure_stats_workspace(y.device)
    z = torch.empty(y.shape, dtype=out_dtype, device=y.device)
    stats = torch.empty((N, C, 2), dtype=torch.float32, device=y.device) if norm else None
    if res is not None:
        assert res.shape == y.shape and res.dtype == out_dtype and res.is_contiguous()
    _lib.check(_lib.lib().p2p_instnorm_act_fwd(_p(y), _p(res), _p(z), _p(stats), N, H * W, C, int(norm), act, slope, eps,
                                               _DT[y.dtype], _DT[out_dtype], _stream()))
    return z, stats


def planes_apply_ok(y, out_dtype):
    """the fused normalise + activate + pad kernel (csrc/in_fused.cu) covers this tensor"""
    return (out_dtype == torch.bfloat16 and y.dtype in (torch.float32, torch.bfloat16) and y.shape[3] % 8 == 0
            and y.shape[3] <= 2048 and y.shape[0] < 65535)


def norm_act_forward_planes(y, bias, raw, res, norm, act, slope, eps, key, want_plain):
    """z = act(InstanceNorm2d(y)) [+ res] (bf16) written DIRECTLY as the padded planes `key` of the layer that consumes it
    (plan_input_layout; None: plain NHWC only).  raw: the conv epilogue's [N,C,2] sums of (y - bias), or None (a separate
    reduction runs).  -> (z, layout or None, stats): z is the plain tensor when want_plain or the consumer is strided, else a
    strided view of the planes' interior; layout = (key, planes) for conv_forward(xin=...); stats = (mean, rstd) [N,C,2]."""
    N, H, W, C = y.shape
    lib = _lib.lib()
    _ensure_stats_workspace(y.device)
    plain_only = key is None
    if plain_only:
        key = (0, 0, H, W, N * H * W, False, 1, C)
    pt, pl, Hh, Wh, rows, reflect, s, Ck = key
    assert Ck == C
    planes = torch.empty((s * s, rows, C), dtype=torch.bfloat16, device=y.device)
    dual = (not plain_only) and (want_plain or s != 1)
    z_plain = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=y.device) if dual else None
    stats = stats_in = stats_out = None
    raw_mode = 0
    if norm:
        stats = torch.empty((N, C, 2), dtype=torch.float32, device=y.device)
        if raw is None:
            _lib.check(lib.p2p_instnorm_stats(_p(y), _p(stats), N, H * W, C, eps, _DT[y.dtype], _stream()))
            stats_in = stats
        else:
            raw_mode, stats_in, stats_out = 1, raw, stats
    rs = (0, 0, 0)
    if res is not None:
        assert res.shape == y.shape and res.dtype == torch.bfloat16
        if res.stride(3) != 1 or any(st % 8 for st in res.stride()[:3]) or res.data_ptr() % 16:
            res = res.contiguous()
        rs = res.stride()[:3]
    _lib.check(lib.p2p_instnorm_act_planes(_p(y), _p(bias) if raw_mode else None, _p(stats_in), raw_mode, _p(stats_out), _p(res),
                                           rs[0], rs[1], rs[2], _p(planes), _p(z_plain), N, H, W, C, pt, pl, Hh, Wh, s, rows,
                                           int(reflect), int(norm), act, slope, eps, _DT[y.dtype], _stream()))
    if plain_only:
        return planes.view(N, H, W, C), None, stats
    if dual:
        z = z_plain
    else:
        z = planes[0, :N * Hh * Wh].view(N, Hh, Wh, C)[:, pt:pt + H, pl:pl + W, :]
    return z, (key, planes), stats


def norm_act_backward(y, stats, dz, norm, act, slope):
    N, H, W, C = y.shape
    _ensure_stats_workspace(y.device)
    dy = torch.empty(y.shape, dtype=dz.dtype, device=y.device)
    work = torch.empty((N, C, 2), dtype=torch.float32, device=y.device) if norm else None
    _lib.check(_lib.lib().p2p_instnorm_act_bwd(_p(y), _p(stats), _p(dz), _p(dy), _p(work), N, H * W, C, int(norm), act,
                                               slope, _DT[y.dtype], _DT[dz.dtype], _stream()))
    return dy


# The InstanceNorm backward writes dy directly as the zero-margined grid the stride-1 gradient kernels of the convolution in
# front of it read (p2p_instnorm_act_bwd_wrap): one pass less over dy per 3x3 / 4x4 stride-1 layer.  P2P_FUSE_BWD_WRAP=0: off.
FUSE_BWD_WRAP = os.environ.get("P2P_FUSE_BWD_WRAP", "1") == "1"


def norm_act_backward_wrap(y, stats, dz, norm, act, slope, plan):
    """norm_act_backward whose result lands in the shared gradient grid of `plan` (_dy_wrap_plan) -> (strided view of dy
    inside the grid, the dy_shared triple _conv_backward takes as dy_pre)"""
    N, H, W, C = y.shape
    Hh, Wh, KH, KW, _ = plan
    _ensure_stats_workspace(y.device)
    buf, P64, rows, slot = _wrap_buffer(y.device, dz.dtype, N, Hh, Wh, KH, KW, C)
    work = torch.empty((N, C, 2), dtype=torch.float32, device=y.device) if norm else None
    _lib.check(_lib.lib().p2p_instnorm_act_bwd_wrap(_p(y), _p(stats), _p(dz), _p(buf[P64:]), _p(work), N, H, W, C, Hh, Wh, rows,
                                                    int(norm), act, slope, _DT[y.dtype], _DT[dz.dtype], _stream()))
    view = buf[P64:P64 + N * Hh * Wh].view(N, Hh, Wh, C)[:, :H, :W, :]
    return view, ((Hh, Wh, rows, C), buf, P64, slot)


# ---------------------------------------------------------------------------------------------------
# autograd Functions (x, outputs: NHWC)
# ---------------------------------------------------------------------------------------------------
NORM_BIAS_GRAD = os.environ.get("P2P_NORM_BIAS_GRAD", "0") == "1"


class GradSink:
    """a conv weight's slice of FusedAdam's flat gradient buffer (physical layout).  first_write() is True for the first
    gradient written after the optimizer's zero_grad -- the kernels may then store instead of accumulate."""

    def __init__(self, view, owner, offset=0):
        self.view, self.owner, self.epoch, self.offset = view, owner, -1, offset
        self.stored_epoch = -1      # epoch whose first write was a STORE by the window wgrad kernel (FusedAdam.zero_grad)
        self.unzeroed_epoch = -1    # epoch for which zero_grad left this slice alone

    def first_write(self):
        fresh = self.epoch != self.owner.grad_epoch
        self.epoch = self.owner.grad_epoch
        return fresh

    def written(self):
        """the wgrad kernels of this weight are enqueued: data-parallel training may start reducing it (train.GradSync)"""
        hook = getattr(self.owner, "grad_written_hook", None)
        if hook is not None:
            hook(self)


# ---- backward-pass phases (train.train_step with a SHARED discriminator forward on the generated image) -------------------------
# The reference runs D three times per step (pix2pixHD_model.py:351-361): on (label, fake.detach()) for the discriminator's
# loss, on (label, real), and once more on (label, fake) for the generator's loss -- the first and the third forward compute
# the same numbers.  train_step runs that forward ONCE and walks its graph twice:
#   phase G (loss_G.backward): gradients flow through D's activations to G; D's weight gradients are not formed
#                              (BACKWARD_OWNERS = {optimizer_G}); the gate in front of D is open;
#   phase D (loss_D.backward): D's weight gradients are formed, the gate is closed (the reference detaches the fake image).
BACKWARD_OWNERS = None      # ids of the optimizers whose parameters receive gradients in the running backward pass (None: all)
GATE_OPEN = True
# Batched discriminator pass (train.BATCH_D): D runs ONCE on cat(real, fake) along the batch -- InstanceNorm is per image, so
# every image sees exactly the arithmetic of a separate pass.  Forward: BATCH_HALVES = B marks the layers of that pass.
# Phase G back-propagates through the FAKE half only (FAKE_HALF_ONLY): those layers slice their saved tensors to [B:], and
# the real half of every gradient tensor that travels between them is never written nor read.
BATCH_HALVES = 0
FAKE_HALF_ONLY = False


def set_backward_phase(owners=None, gate_open=True, fake_half_only=False):
    global BACKWARD_OWNERS, GATE_OPEN, FAKE_HALF_ONLY
    BACKWARD_OWNERS = None if owners is None else {id(o) for o in owners}
    GATE_OPEN = bool(gate_open)
    FAKE_HALF_ONLY = bool(fake_half_only)


def set_batch_halves(B):
    """forward-time marker: the layers run until the next call belong to a pass over cat(real[B], fake[B])"""
    global BATCH_HALVES
    old, BATCH_HALVES = BATCH_HALVES, int(B)
    return old


def _owner_active(gphys):
    return BACKWARD_OWNERS is None or gphys is None or id(gphys.owner) in BACKWARD_OWNERS


class BatchHalvesFn(torch.autograd.Function):
    """x [2B, ...] -> (x[:B], x[B:]) (views).  Backward: in phase G only the fake half carries a gradient and only that half
    of the result is written; otherwise missing halves are zeros."""

    @staticmethod
    def forward(ctx, x, B):
        ctx.set_materialize_grads(False)
        ctx.meta = (int(B), tuple(x.shape), tuple(x.stride()), x.dtype, x.device)
        return x[:B], x[B:]

    @staticmethod
    def backward(ctx, d_real, d_fake):
        if d_real is None and d_fake is None:
            return None, None
        B, shape, stride, dtype, device = ctx.meta
        out = torch.empty_strided(shape, stride, dtype=dtype, device=device)
        if d_real is None and FAKE_HALF_ONLY:
            out[B:].copy_(d_fake)               # the real half is never read in this phase
            return out, None
        for half, d in ((out[:B], d_real), (out[B:], d_fake)):
            if d is None:
                half.zero_()
            else:
                half.copy_(d)
        return out, None


class GradGateFn(torch.autograd.Function):
    """identity whose gradient passes only while the gate is open (set_backward_phase)"""

    @staticmethod
    def forward(ctx, x):
        ctx.set_materialize_grads(False)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g if GATE_OPEN else None


class ConvBlockFn(torch.autograd.Function):
    """[ReflectionPad2d] Conv2d | ConvTranspose2d  ->  [InstanceNorm2d]  ->  [ReLU | LeakyReLU]  [+ residual]
    as ONE differentiable op.  In bf16 mode the pre-normalisation conv output stays fp32 inside the op (see
    include/p2p_b200.h), so only O(1)-scaled tensors are ever rounded to bf16.

    Round 2: the InstanceNorm sums come out of the conv epilogue, and the normalise / activate / residual pass writes its
    result directly as the padded planes of the NEXT convolution (`out_plan(shape, dtype)` -> that layer's
    plan_input_layout; `xin` = the (key, planes) this layer's producer wrote).  Returns (z, layout or None)."""

    @staticmethod
    def forward(ctx, x, weight, bias, w16, res, cfg, norm, act, slope, eps, gphys=None, xin=None, out_plan=None,
                want_plain=False, gated=False):
        _need_cuda(x, weight)
        ctx.gated = bool(gated)         # x sits right behind a GradGateFn: no input gradient while the gate is closed
        ctx.halves = BATCH_HALVES       # > 0: this call is part of a batched (real, fake) discriminator pass
        ctx.set_materialize_grads(False)    # an undefined upstream gradient (closed gate) must not run this layer's backward on zeros
        plain = (not norm) and act == ACT_NONE and res is None
        aux = [] if ctx.needs_input_grad[1] else None
        Cout = weight.shape[1] if cfg.transposed else weight.shape[0]
        fused = (not plain) and FUSE_NORM_PAD and x.dtype == torch.bfloat16 and Cout % 8 == 0 and Cout <= 2048 and x.shape[0] < 65535
        # bf16 mode, fused path: the pre-normalisation output is stored in bf16 WITHOUT the conv bias (InstanceNorm2d(affine=False)
        # removes any per-channel constant, so the bias only cost mantissa bits before the mean subtraction); the statistics are
        # taken of the stored values, as torch.autocast does for the reference.  P2P_BF16_PRENORM=0 keeps fp32 (round-1 behaviour).
        y16 = fused and norm and BF16_PRENORM and not (EPI_STATS and not deterministic())
        y_dtype = (x.dtype if y16 else torch.float32) if norm else x.dtype
        cbias = None if y16 else bias
        raw = None
        if fused and norm and EPI_STATS and not y16:
            raw = torch.empty((x.shape[0], Cout, 2), dtype=torch.float32, device=x.device)
            y, took = conv_forward(x, weight, cbias, w16, cfg, y_dtype, aux, xin, raw)
            if not took:
                raw = None
        else:
            y = conv_forward(x, weight, cbias, w16, cfg, y_dtype, aux, xin)
        ctx.xpad = aux[0] if aux else None
        layout = None
        if plain:
            z, stats, ysave = y, None, None
        elif fused and planes_apply_ok(y, x.dtype):
            key = out_plan(tuple(y.shape), x.dtype) if out_plan is not None else None
            z, layout, stats = norm_act_forward_planes(y, bias, raw, res, norm, act, slope, eps, key, want_plain)
            ysave = y
        else:
            if norm and raw is not None:
                raise RuntimeError("epilogue statistics without the fused apply kernel")
            if res is not None:
                res = res.contiguous()
            z, stats = norm_act_forward(y, res, norm, act, slope, eps, x.dtype)
            ysave = y
        ctx.save_for_backward(x, weight, w16, ysave, stats)
        ctx.cfg = (cfg, norm, act, slope, bias is not None, res is not None, plain)
        ctx.gphys = gphys
        ctx.bias_n = bias.numel() if bias is not None else 0
        return z, layout

    @staticmethod
    def backward(ctx, dz, _unused=None):
        if dz is None:
            return (None,) * 15
        x, weight, w16, y, stats = ctx.saved_tensors
        cfg, norm, act, slope, has_bias, has_res, plain = ctx.cfg
        dz = dz.contiguous()
        xpad = ctx.xpad
        half = ctx.halves if (ctx.halves and FAKE_HALF_ONLY and not has_res) else 0
        if half:                        # phase G of a batched discriminator pass: the fake images only
            assert x.shape[0] == 2 * half and dz.shape[0] == 2 * half
            dz, x, xpad = dz[half:], x[half:], None
            y = y[half:] if y is not None else None
            stats = stats[half:] if stats is not None else None
        # a bias in front of InstanceNorm2d(affine=False) has an exactly zero gradient (the mean subtraction removes it; the
        # reference's autograd returns rounding noise of ~1e-7 relative size): not computed unless P2P_NORM_BIAS_GRAD=1
        active = _owner_active(ctx.gphys)       # this layer's parameters take gradients in the running backward phase
        need_db = has_bias and ctx.needs_input_grad[2] and (not norm or NORM_BIAS_GRAD) and active
        need_dx = ctx.needs_input_grad[0] and not (ctx.gated and not GATE_OPEN)
        need_dw = ctx.needs_input_grad[1] and active
        dy_pre = None
        plan = None if plain or y.shape[0] >= 65535 else _dy_wrap_plan(x, weight, cfg, need_dx, need_dw, need_db)
        if plain:
            dy = dz
        elif plan is not None:
            dy, dy_pre = norm_act_backward_wrap(y, stats, dz, norm, act, slope, plan)
        else:
            dy = norm_act_backward(y, stats, dz, norm, act, slope)
        dx, dw, db = conv_backward(x, weight, w16, dy, cfg, need_dx, need_dw, need_db, ctx.gphys, xpad, dy_pre)
        if half:
            if need_dw or need_db:
                raise RuntimeError("the fake-half backward phase must not form parameter gradients of the batched pass")
            if dx is not None:          # autograd wants the full batch's shape; the real half stays unwritten (never read)
                full = torch.empty((2 * half,) + tuple(dx.shape[1:]), dtype=dx.dtype, device=dx.device)
                full[half:].copy_(dx)
                dx = full
        if has_bias and ctx.needs_input_grad[2] and not need_db and ctx.gphys is None and active:
            db = torch.zeros(ctx.bias_n, dtype=torch.float32, device=x.device)    # no flat gradient buffer behind .grad: give autograd the zero
        return dx, dw, db, None, (dz if has_res else None), None, None, None, None, None, None, None, None, None, None


FUSE_NORM_PAD = os.environ.get("P2P_FUSE_NORM_PAD", "1") == "1"
# measured (round 2): storing the pre-normalisation output in bf16 moves the G3L2 step by 0.4 % (these passes are launch- and
# latency-bound, not byte-bound) and costs four per-layer 2e-2 tolerance checks: fp32 stays the default
BF16_PRENORM = os.environ.get("P2P_BF16_PRENORM", "0") == "1"
# InstanceNorm sums from the conv epilogue instead of a separate reduction pass.  Measured on B200 (round 2, G3L2 step): the
# window kernel's forward launches get 0.74 ms slower per step (column sums + red.add in the epilogue warps) while the
# reduction kernels they replace cost 0.4 ms -- a net loss, so the separate pass stays the default (P2P_EPI_STATS=1 opts in).
EPI_STATS = os.environ.get("P2P_EPI_STATS", "0") == "1"


class InstNormActFn(torch.autograd.Function):
    """stand-alone z = act(InstanceNorm2d(y)) [+ res] (norm=False: bare activation); same dtype in and out"""

    @staticmethod
    def forward(ctx, y, res, norm, act, slope, eps):
        _need_cuda(y)
        ctx.set_materialize_grads(False)
        y = y.contiguous()
        if res is not None:
            res = res.contiguous()
        z, stats = norm_act_forward(y, res, norm, act, slope, eps, y.dtype)
        ctx.save_for_backward(y, stats)
        ctx.cfg = (bool(norm), act, slope, res is not None)
        return z

    @staticmethod
    def backward(ctx, dz):
        if dz is None:
            return (None,) * 6
        y, stats = ctx.saved_tensors
        norm, act, slope, has_res = ctx.cfg
        dz = dz.contiguous()
        return norm_act_backward(y, stats, dz, norm, act, slope), (dz if has_res else None), None, None, None, None


class AvgPool3s2Fn(torch.autograd.Function):
    """nn.AvgPool2d(3, stride=2, padding=[1,1], count_include_pad=False), networks.py:166, 309."""

    @staticmethod
    def forward(ctx, x, gated=False):
        _need_cuda(x)
        ctx.gated = bool(gated)
        ctx.halves = BATCH_HALVES
        ctx.set_materialize_grads(False)
        x = x.contiguous()
        N, H, W, C = x.shape
        y = torch.empty((N, (H - 1) // 2 + 1, (W - 1) // 2 + 1, C), dtype=x.dtype, device=x.device)
        _lib.check(_lib.lib().p2p_avgpool3s2_fwd(_p(x), _p(y), N, H, W, C, _DT[x.dtype], _stream()))
        ctx.shape = (N, H, W, C)
        return y

    @staticmethod
    def backward(ctx, dy):
        if dy is None or (ctx.gated and not GATE_OPEN):
            return None, None
        N, H, W, C = ctx.shape
        dy = dy.contiguous()
        dx = torch.empty((N, H, W, C), dtype=dy.dtype, device=dy.device)
        if ctx.halves and FAKE_HALF_ONLY and N == 2 * ctx.halves:       # batched discriminator pass, phase G: the fake images only
            B = ctx.halves
            _lib.check(_lib.lib().p2p_avgpool3s2_bwd(_p(dy[B:]), _p(dx[B:]), B, H, W, C, _DT[dy.dtype], _stream()))
        else:
            _lib.check(_lib.lib().p2p_avgpool3s2_bwd(_p(dy), _p(dx), N, H, W, C, _DT[dy.dtype], _stream()))
        return dx, None


class TanhFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _need_cuda(x)
        ctx.set_materialize_grads(False)
        x = x.contiguous()
        y = torch.empty_like(x)
        _lib.check(_lib.lib().p2p_tanh_fwd(_p(x), _p(y), x.numel(), _DT[x.dtype], _stream()))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        if dy is None:
            return None
        (y,) = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(y)
        _lib.check(_lib.lib().p2p_tanh_bwd(_p(y), _p(dy), _p(dx), y.numel(), _DT[y.dtype], _stream()))
        return dx


class MseConstFn(torch.autograd.Function):
    """weight * mean((x - target)^2) as a 0-dim fp32 tensor (nn.MSELoss against a constant label map)."""

    @staticmethod
    def forward(ctx, x, target, weight):
        _need_cuda(x)
        x = x.contiguous()
        loss = torch.zeros((), dtype=torch.float32, device=x.device)
        _lib.check(_lib.lib().p2p_loss_mse_const(_p(x), target, weight, _p(loss), None, 0.0, None, x.numel(), _DT[x.dtype], _stream()))
        ctx.save_for_backward(x)
        ctx.cfg = (target, weight)
        return loss

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        target, weight = ctx.cfg
        dx = torch.empty_like(x)
        g = g.to(torch.float32).contiguous()          # upstream scalar stays on the device
        _lib.check(_lib.lib().p2p_loss_mse_const(_p(x), target, weight, None, _p(dx), 1.0, _p(g), x.numel(), _DT[x.dtype], _stream()))
        return dx, None, None


class L1Fn(torch.autograd.Function):
    """weight * mean(|a - b|), gradient to `a` only (b is the detached real feature, pix2pixHD_model.py:396-398)."""

    @staticmethod
    def forward(ctx, a, b, weight):
        _need_cuda(a, b)
        a, b = a.contiguous(), b.contiguous()
        assert a.shape == b.shape and a.dtype == b.dtype
        loss = torch.zeros((), dtype=torch.float32, device=a.device)
        _lib.check(_lib.lib().p2p_loss_l1(_p(a), _p(b), weight, _p(loss), None, 0.0, None, a.numel(), _DT[a.dtype], _stream()))
        ctx.save_for_backward(a, b)
        ctx.weight = weight
        return loss

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        da = torch.empty_like(a)
        g = g.to(torch.float32).contiguous()
        _lib.check(_lib.lib().p2p_loss_l1(_p(a), _p(b), ctx.weight, None, _p(da), 1.0, _p(g), a.numel(), _DT[a.dtype], _stream()))
        return da, None, None


class MseConstSumFn(torch.autograd.Function):
    """sum_i weight * mean((x_i - target)^2) over several tensors as ONE 0-dim fp32 tensor: the loss kernels accumulate into
    the same scalar (GANLoss over the discriminator's scales, networks.py:98-106) -- one zero-fill and no add kernels"""

    @staticmethod
    def forward(ctx, target, weight, *xs):
        _need_cuda(*xs)
        xs = tuple(x.contiguous() for x in xs)
        loss = torch.zeros((), dtype=torch.float32, device=xs[0].device)
        for x in xs:
            _lib.check(_lib.lib().p2p_loss_mse_const(_p(x), target, weight, _p(loss), None, 0.0, None, x.numel(), _DT[x.dtype], _stream()))
        ctx.save_for_backward(*xs)
        ctx.cfg = (target, weight)
        return loss

    @staticmethod
    def backward(ctx, g):
        target, weight = ctx.cfg
        g = g.to(torch.float32).contiguous()
        out = []
        for i, x in enumerate(ctx.saved_tensors):
            if not ctx.needs_input_grad[2 + i]:
                out.append(None)
                continue
            dx = torch.empty_like(x)
            _lib.check(_lib.lib().p2p_loss_mse_const(_p(x), target, weight, None, _p(dx), 1.0, _p(g), x.numel(), _DT[x.dtype], _stream()))
            out.append(dx)
        return (None, None) + tuple(out)


class L1SumFn(torch.autograd.Function):
    """sum_i weight * mean(|a_i - b_i|) as ONE 0-dim fp32 tensor, gradients to the a_i only (the feature-matching sum of
    pix2pixHD_model.py:390-398); arguments: weight, a_0 .. a_{n-1}, b_0 .. b_{n-1}"""

    @staticmethod
    def forward(ctx, weight, *ts):
        n = len(ts) // 2
        _need_cuda(*ts)
        ts = tuple(t.contiguous() for t in ts)
        loss = torch.zeros((), dtype=torch.float32, device=ts[0].device)
        for a, b in zip(ts[:n], ts[n:]):
            assert a.shape == b.shape and a.dtype == b.dtype
            _lib.check(_lib.lib().p2p_loss_l1(_p(a), _p(b), weight, _p(loss), None, 0.0, None, a.numel(), _DT[a.dtype], _stream()))
        ctx.save_for_backward(*ts)
        ctx.weight = weight
        return loss

    @staticmethod
    def backward(ctx, g):
        ts = ctx.saved_tensors
        n = len(ts) // 2
        g = g.to(torch.float32).contiguous()
        out = []
        for i, (a, b) in enumerate(zip(ts[:n], ts[n:])):
            if not ctx.needs_input_grad[1 + i]:
                out.append(None)
                continue
            da = torch.empty_like(a)
            _lib.check(_lib.lib().p2p_loss_l1(_p(a), _p(b), ctx.weight, None, _p(da), 1.0, _p(g), a.numel(), _DT[a.dtype], _stream()))
            out.append(da)
        return (None,) + tuple(out) + (None,) * n


class FramesDbFn(torch.autograd.Function):
    """dB(|gain * window * frames|) = amplitude_to_DB(|.|, 20, min_value, 1) of discriminate_time_D
    (pix2pixHD_model.py:317-318) fused with the sqrt(up_ratio-1) * window scaling of :375; frames [..., N] fp32."""

    @staticmethod
    def forward(ctx, frames, window, gain, min_value):
        _need_cuda(frames, window)
        frames = frames.to(torch.float32).contiguous()
        out = torch.empty_like(frames)
        N = frames.shape[-1]
        _lib.check(_lib.lib().p2p_frames_db_fwd(_p(frames), _p(window), gain, _p(out), frames.numel() // N, N, min_value, _stream()))
        ctx.save_for_backward(frames, window)
        ctx.cfg = (gain, min_value)
        return out

    @staticmethod
    def backward(ctx, dout):
        frames, window = ctx.saved_tensors
        gain, min_value = ctx.cfg
        dout = dout.to(torch.float32).contiguous()
        dframes = torch.empty_like(frames)
        N = frames.shape[-1]
        _lib.check(_lib.lib().p2p_frames_db_bwd(_p(frames), _p(window), gain, _p(dout), _p(dframes), frames.numel() // N, N,
                                                min_value, _stream()))
        return dframes, None, None, None


class FrameMatchFn(torch.autograd.Function):
    """weight * MSE(frames[:, :-1, half:] * w[:half], frames[:, 1:, :half] * w[half:]) (pix2pixHD_model.py:408-415)."""

    @staticmethod
    def forward(ctx, frames, window, weight):
        _need_cuda(frames, window)
        frames = frames.to(torch.float32).contiguous()
        B, F, N = frames.shape
        loss = torch.zeros((), dtype=torch.float32, device=frames.device)
        _lib.check(_lib.lib().p2p_frame_match(_p(frames), _p(window), weight, _p(loss), None, 0.0, None, B, F, N, _stream()))
        ctx.save_for_backward(frames, window)
        ctx.weight = weight
        return loss

    @staticmethod
    def backward(ctx, g):
        frames, window = ctx.saved_tensors
        B, F, N = frames.shape
        dframes = torch.empty_like(frames)
        g = g.to(torch.float32).contiguous()
        _lib.check(_lib.lib().p2p_frame_match(_p(frames), _p(window), ctx.weight, None, _p(dframes), 1.0, _p(g), B, F, N, _stream()))
        return dframes, None, None


class SpectroDecodeFn(torch.autograd.Function):
    """denormalize + explicit +/- decode + permute(0,2,1) (pix2pixHD_model.py:229-237, 246-248) with a gradient:
    x [B,2,M,F] image (any strides, fp32 or bf16) -> coefficients [B,F,M] fp32.  The gradient comes back as a
    channels-last buffer behind an NCHW shape, in x's dtype."""

    @staticmethod
    def forward(ctx, x, norm, alpha, min_value, gain):
        _need_cuda(x, norm)
        B, C, M, F = x.shape
        assert C == 2, "the differentiable decode is the explicit-encoding branch"
        coef = torch.empty((B, F, M), dtype=torch.float32, device=x.device)
        sb, sc, sh, sw = x.stride()
        _lib.check(_lib.lib().p2p_spectro_decode(_p(x), sb, sc, sh, sw, _p(norm), None, _p(coef), B, C, F, M, alpha, min_value,
                                                 gain, _DT[x.dtype], _stream()))
        ctx.save_for_backward(x, norm)
        ctx.cfg = (alpha, gain)
        return coef

    @staticmethod
    def backward(ctx, dcoef):
        x, norm = ctx.saved_tensors
        alpha, gain = ctx.cfg
        B, C, M, F = x.shape
        dcoef = dcoef.to(torch.float32).contiguous()
        dx = torch.empty((B, M, F, 2), dtype=torch.float32, device=x.device)
        sb, sc, sh, sw = x.stride()
        _lib.check(_lib.lib().p2p_spectro_decode_bwd(_p(x), sb, sc, sh, sw, _p(norm), _p(dcoef), _p(dx), B, F, M, alpha, gain,
                                                     _DT[x.dtype], _stream()))
        return dx.permute(0, 3, 1, 2).to(x.dtype), None, None, None, None


def frame_window(audio, window, n_frames, hop, offset):
    """windowed frames [B, F, win] of a zero-padded signal (models/mdct.py:497-503)"""
    _need_cuda(audio, window)
    audio = audio.to(torch.float32).contiguous()
    B, T = audio.shape
    win = window.numel()
    frames = torch.empty((B, n_frames, win), dtype=torch.float32, device=audio.device)
    _lib.check(_lib.lib().p2p_frame_window(_p(audio), _p(window), _p(frames), B, T, n_frames, hop, win, offset, _stream()))
    return frames


def adam_step(w, g, m, v, w16, lr, beta1, beta2, eps, step):
    _need_cuda(w, g, m, v)
    assert w.is_contiguous() and g.is_contiguous() and w.dtype == torch.float32 and g.dtype == torch.float32
    _lib.check(_lib.lib().p2p_adam_step(_p(w), _p(g), _p(m), _p(v), _p(w16), w.numel(), lr, beta1, beta2, eps, step, _stream()))


def adam_step_dev(w, g, m, v, w16, lr_dev, beta1, beta2, eps, grad_scale, step_dev, scratch):
    """Adam with the step counter / learning rate on the device (CUDA-graph friendly)"""
    _need_cuda(w, g, m, v, lr_dev, step_dev, scratch)
    assert w.is_contiguous() and g.is_contiguous() and w.dtype == torch.float32 and g.dtype == torch.float32
    _lib.check(_lib.lib().p2p_adam_step_dev(_p(w), _p(g), _p(m), _p(v), _p(w16), w.numel(), _p(lr_dev), beta1, beta2, eps,
                                            grad_scale, _p(step_dev), _p(scratch), _stream()))
