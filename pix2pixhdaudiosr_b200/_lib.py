"""ctypes binding of libp2p_b200.so (include/p2p_b200.h).  No fallback: if the library is missing
or a call fails, this raises."""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libp2p_b200.so")

_c = ctypes
_SIGS = {
    "p2p_last_error": (_c.c_char_p, []),
    "p2p_abi_version": (_c.c_int, []),
    "p2p_mdct4_force_general": (_c.c_int, [_c.c_int]),
    "p2p_mdct4_table_words": (_c.c_int, [_c.c_int]),
    "p2p_mdct4_fill_tables": (_c.c_int, [_c.c_int, _c.c_void_p]),
    "p2p_mdct4_analysis": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int64,
                                      _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int64, _c.c_float, _c.c_void_p]),
    "p2p_mdct4_synthesis": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int,
                                       _c.c_int64, _c.c_int, _c.c_int, _c.c_int, _c.c_int64, _c.c_float, _c.c_void_p]),
}
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
    return _lib


def exported_symbols():
    return list(_SIGS)


def check(rc):
    if rc != 0:
        msg = lib().p2p_last_error().decode(errors="replace")
        exc = {-1: ValueError, -3: NotImplementedError}.get(rc, P2PError)
        raise exc(f"libp2p_b200 error {rc}: {msg}")
