"""ctypes binding of libp2p_b200.so (include/p2p_b200.h).  No fallback: if the library is missing
or a call fails, this raises."""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("P2P_LIB_PATH") or os.path.join(_PKG, "libp2p_b200.so")      # override: instrumented builds (profiles/)

_c = ctypes
_SIGS = {
    "p2p_last_error": (_c.c_char_p, []),
    "p2p_abi_version": (_c.c_int, []),
    "p2p_kernel_launches": (_c.c_longlong, []),
    "p2p_set_deterministic": (_c.c_int, [_c.c_int]),
    "p2p_mdct4_force_general": (_c.c_int, [_c.c_int]),
    "p2p_mdct2_force_general": (_c.c_int, [_c.c_int]),
    "p2p_mdct4_table_words": (_c.c_int, [_c.c_int]),
    "p2p_mdct4_fill_tables": (_c.c_int, [_c.c_int, _c.c_void_p]),
    "p2p_mdct4_analysis": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int64,
                                      _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int64, _c.c_float, _c.c_void_p]),
    "p2p_mdct4_synthesis": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int,
                                       _c.c_int64, _c.c_int, _c.c_int, _c.c_int, _c.c_int64, _c.c_float, _c.c_void_p]),
}
_vp, _i, _f, _ll = _c.c_void_p, _c.c_int, _c.c_float, _c.c_longlong
MAX_TAPS = 49


class ConvPass(ctypes.Structure):
    """p2p_conv_pass of include/p2p_b200.h."""
    _fields_ = [("N", _i), ("H", _i), ("W", _i), ("K", _i), ("OH", _i), ("OW", _i), ("J", _i),
                ("gy", _i), ("gx", _i), ("sy", _i), ("sx", _i),
                ("oys", _i), ("oyo", _i), ("oxs", _i), ("oxo", _i),
                ("ntaps", _i), ("pad_mode", _i),
                ("w_tap", _ll), ("w_k", _ll), ("w_j", _ll),
                ("dh", _c.c_byte * MAX_TAPS), ("dw", _c.c_byte * MAX_TAPS), ("wtap", _c.c_ubyte * MAX_TAPS)]


class WindowDesc(ctypes.Structure):
    """p2p_conv_window_desc of include/p2p_b200.h."""
    _fields_ = [("nplanes", _i), ("N", _i), ("Hp", _i), ("Wp", _i), ("C", _i), ("OHf", _i), ("OWf", _i), ("J", _i),
                ("wtaps_total", _i), ("w_tap", _ll), ("w_k", _ll), ("w_j", _ll), ("nsub", _i), ("oys", _i), ("oxs", _i),
                ("gy", _i * 4), ("gx", _i * 4), ("oyo", _i * 4), ("oxo", _i * 4), ("ntaps", _i),
                ("tap_sub", _c.c_byte * MAX_TAPS), ("tap_plane", _c.c_byte * MAX_TAPS),
                ("tap_dh", _c.c_short * MAX_TAPS), ("tap_dw", _c.c_short * MAX_TAPS), ("tap_w", _c.c_ubyte * MAX_TAPS),
                ("wrap", _i), ("plane_rows", _ll)]


_SIGS.update({
    "p2p_conv_forward": (_i, [_c.POINTER(ConvPass), _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "p2p_conv_force_simple": (_i, [_i]),
    "p2p_conv_umma_launches": (_ll, []),
    "p2p_conv_wgrad": (_i, [_c.POINTER(ConvPass), _vp, _vp, _vp, _i, _vp]),
    "p2p_pad_nhwc": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_conv_window": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _ll, _ll, _ll, _i, _vp]),
    "p2p_pad_planes": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _ll, _i, _i, _vp]),
    "p2p_conv_wgrad_window": (_i, [_vp, _i, _vp, _i, _ll, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "p2p_conv_wgrad_window_stored": (_i, []),
    "p2p_conv_window_debug_buffer": (_i, [_vp]),
    "p2p_conv_window_run": (_i, [_c.POINTER(WindowDesc), _vp, _vp, _vp, _vp, _i, _vp]),
    "p2p_conv_window_stats": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _ll, _ll, _ll, _i, _vp, _vp]),
    "p2p_conv_window_run_stats": (_i, [_c.POINTER(WindowDesc), _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "p2p_instnorm_act_planes": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _ll, _ll, _ll, _vp, _vp] + [_i] * 9 + [_ll, _i, _i, _i, _f, _f, _i, _vp]),
    "p2p_bias_grad": (_i, [_vp, _vp, _ll, _i, _i, _vp]),
    "p2p_reflect_pad_fold": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_reflect_pad_fold_f32": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_pack_in": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_fold_out": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_fold_out_rows": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_pack_dy": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_im2col": (_i, [_vp, _vp] + [_i] * 12 + [_ll, _vp]),
    "p2p_col2im": (_i, [_vp, _vp] + [_i] * 11 + [_vp]),
    "p2p_pack_dy_full": (_i, [_vp, _vp] + [_i] * 7 + [_ll, _vp]),
    "p2p_instnorm_act_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _ll, _i, _i, _i, _f, _f, _i, _i, _vp]),
    "p2p_instnorm_act_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _ll, _i, _i, _i, _f, _i, _i, _vp]),
    "p2p_instnorm_act_bwd_wrap": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _ll, _i, _i, _f, _i, _i, _vp]),
    "p2p_instnorm_stats": (_i, [_vp, _vp, _i, _ll, _i, _f, _i, _vp]),
    "p2p_workspace_bytes": (_i, [_i, _i, _vp, _vp]),
    "p2p_instnorm_stats_static": (_i, [_vp, _vp, _i, _ll, _i, _f, _i, _vp]),
    "p2p_instnorm_workspace": (_i, [_vp, _ll, _vp, _ll]),
    "p2p_avgpool3s2_fwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "p2p_avgpool3s2_bwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "p2p_tanh_fwd": (_i, [_vp, _vp, _ll, _i, _vp]),
    "p2p_tanh_bwd": (_i, [_vp, _vp, _vp, _ll, _i, _vp]),
    "p2p_loss_mse_const": (_i, [_vp, _f, _f, _vp, _vp, _f, _vp, _ll, _i, _vp]),
    "p2p_loss_l1": (_i, [_vp, _vp, _f, _vp, _vp, _f, _vp, _ll, _i, _vp]),
    "p2p_spectro_encode": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _vp]),
    "p2p_spectro_normalize": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "p2p_spectro_decode": (_i, [_vp, _ll, _ll, _ll, _ll, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _f, _i, _vp]),
    "p2p_mdct2_table_words": (_i, [_i]),
    "p2p_mdct2_fill_tables": (_i, [_i, _vp]),
    "p2p_mdct2_analysis": (_i, [_vp, _vp, _vp, _vp, _i, _ll, _i, _i, _i, _i, _ll, _f, _f, _vp]),
    "p2p_mdct2_synthesis": (_i, [_vp, _vp, _vp, _vp, _i, _i, _ll, _i, _i, _i, _ll, _f, _f, _vp]),
    "p2p_resample": (_i, [_vp, _vp, _vp, _i, _ll, _ll, _i, _i, _i, _vp]),
    "p2p_frame_window": (_i, [_vp, _vp, _vp, _i, _ll, _i, _i, _i, _ll, _vp]),
    "p2p_frames_db_fwd": (_i, [_vp, _vp, _f, _vp, _ll, _i, _f, _vp]),
    "p2p_frames_db_bwd": (_i, [_vp, _vp, _f, _vp, _vp, _ll, _i, _f, _vp]),
    "p2p_frame_match": (_i, [_vp, _vp, _f, _vp, _vp, _f, _vp, _i, _i, _i, _vp]),
    "p2p_spectro_decode_bwd": (_i, [_vp, _ll, _ll, _ll, _ll, _vp, _vp, _vp, _i, _i, _i, _f, _f, _i, _vp]),
    "p2p_adam_step_dev": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _vp, _f, _f, _f, _f, _vp, _vp, _vp]),
    "p2p_adam_step": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _f, _f, _f, _f, _i, _vp]),
})
_lib = None


class P2PError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise P2PError(
                f"{LIB_PATH} is missing: build it with `python -m pix2pixhdaudiosr_b200.build` "
                "(there is no CPU or PyTorch fallback for this path)")
        h = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        _lib = h
    return _profile if _profile is not None else _lib


class _ProfiledLib:
    """proxy used by start_call_profile(): brackets every C-ABI call that takes a stream with a CUDA event pair"""

    def __init__(self, h, records, external):
        self._h, self._records, self._external = h, records, external

    def __getattr__(self, name):
        fn = getattr(self._h, name)
        sig = _SIGS.get(name)
        if sig is None or not sig[1] or sig[1][-1] is not _vp or name in ("p2p_mdct4_fill_tables", "p2p_mdct2_fill_tables", "p2p_conv_window_debug_buffer"):
            return fn
        import torch

        def call(*a):
            e0 = torch.cuda.Event(enable_timing=True, external=self._external)
            e1 = torch.cuda.Event(enable_timing=True, external=self._external)
            e0.record()
            rc = fn(*a)
            e1.record()
            self._records.append((name, e0, e1))
            return rc
        return call


_profile = None


def start_call_profile(external=False):
    """from now on lib() hands out a proxy that times every kernel-launching entry point (bench.py's per-entry-point
    breakdown of one step; external=True while the step is being captured into a CUDA graph)"""
    global _profile
    lib()
    _profile = _ProfiledLib(_lib, [], external)


def stop_call_profile():
    """-> {entry point: (calls, milliseconds)} since start_call_profile(); call after the work has completed"""
    global _profile
    import torch
    recs, _profile = (_profile._records if _profile is not None else []), None
    torch.cuda.synchronize()
    out = {}
    for name, e0, e1 in recs:
        c, ms = out.get(name, (0, 0.0))
        out[name] = (c + 1, ms + e0.elapsed_time(e1))
    return out


def pause_call_profile():
    """stop handing out the proxy but keep the records (graph capture finished; replay, then stop_call_profile())"""
    global _profile, _held
    _held, _profile = _profile, None


def resume_call_profile():
    global _profile, _held
    _profile, _held = _held, None


_held = None


def exported_symbols():
    return list(_SIGS)


def check(rc):
    if rc != 0:
        msg = lib().p2p_last_error().decode(errors="replace")
        exc = {-1: ValueError, -3: NotImplementedError}.get(rc, P2PError)
        raise exc(f"libp2p_b200 error {rc}: {msg}")
