"""CPU: the C-ABI library loads, exports every symbol include/p2p_b200.h declares, its host-side
functions are right, and the Python mirror's host logic (pads, frame counts, asserts) matches the
reference.  No device compute here."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import fast_model as FM
from cases import CASES
from pix2pixhdaudiosr_b200 import _lib
from pix2pixhdaudiosr_b200.models.mdct import IMDCT4, MDCT4
from pix2pixhdaudiosr_b200.util.util import kbdwin

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "p2p_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(p2p_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    h = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert "p2p_mdct4_analysis" in syms and "p2p_mdct4_synthesis" in syms
    for s in syms:
        assert hasattr(h, s), s
    assert set(_lib.exported_symbols()) == set(syms)
    assert _lib.lib().p2p_abi_version() >= 1


@pytest.mark.parametrize("n_fft", [16, 64, 256, 512, 1024, 2048, 4096])
def test_tables(n_fft):
    lib = _lib.lib()
    L = n_fft // 4
    words = lib.p2p_mdct4_table_words(n_fft)
    assert words == 9 * L
    buf = np.zeros(words, np.float32)
    assert lib.p2p_mdct4_fill_tables(n_fft, buf.ctypes.data) == 0
    pre, post = FM.tables(n_fft)
    got_pre = buf[0:2 * L].view(np.complex64)
    got_post = buf[2 * L:4 * L].view(np.complex64)
    roots = buf[4 * L:6 * L].view(np.complex64)
    perm = buf[6 * L:7 * L].view(np.int32)
    xtw = buf[7 * L:9 * L].view(np.complex64)
    VP = {64: (8, 8), 128: (16, 8), 256: (16, 16), 512: (32, 16), 1024: (32, 32)}.get(L)
    if VP:
        V, P = VP
        want = np.exp(-2j * np.pi * np.outer(np.arange(V), np.arange(P)) / L).reshape(-1)
        assert np.abs(xtw - want).max() < 1e-7
    assert np.abs(got_pre - pre).max() < 1e-7 and np.abs(got_post - post).max() < 1e-7
    assert np.abs(roots - np.exp(-2j * np.pi * np.arange(L) / L)).max() < 1e-7
    assert sorted(perm.tolist()) == list(range(L))
    # the permutation is the one an in-place DIF (radix 4.., then 2) produces: emulate it
    z = (np.random.default_rng(1).standard_normal(L) + 1j * np.random.default_rng(2).standard_normal(L))
    a, S = z.copy(), L
    while S >= 4:
        q = S // 4
        b = a.reshape(L // S, 4, q)
        j = np.arange(q)
        out = np.empty_like(b)
        for k1 in range(4):
            acc = sum(b[:, n1, :] * (-1j) ** (n1 * k1) for n1 in range(4))
            out[:, k1, :] = acc * np.exp(-2j * np.pi * j * k1 / S)
        a, S = out.reshape(L), q
    if S == 2:
        b = a.reshape(L // 2, 2)
        a = np.stack([b[:, 0] + b[:, 1], b[:, 0] - b[:, 1]], 1).reshape(L)
    ref = np.fft.fft(z)
    assert np.abs(a - ref[perm]).max() < 1e-9 * L


def test_bad_arguments_are_reported_not_crashed():
    lib = _lib.lib()
    assert lib.p2p_mdct4_table_words(1000) == -3
    assert b"power of two" in lib.p2p_last_error()
    rc = lib.p2p_mdct4_analysis(None, None, None, None, 1, 10, 1, 1024, 2048, 1024, 0, 1.0, None)
    assert rc == -1 and b"hop_length" in lib.p2p_last_error()
    with pytest.raises(NotImplementedError):
        _lib.check(lib.p2p_mdct4_table_words(48))


def test_module_host_logic_matches_reference_shapes(golden):
    for name, (n_fft, hop, win, center, kbd, one_d) in CASES.items():
        m = MDCT4(n_fft, hop, win, kbdwin if kbd else None, center, device="cpu")
        x = golden[name + "/x"]
        sp, ep = m.pads(x.shape[0])
        F = (x.shape[-1] + sp + ep - win) // hop + 1
        assert F == golden[name + "/spec"].shape[-2], name
        im = IMDCT4(n_fft, hop, win, kbdwin if kbd else None, center, device="cpu")
        crop, t_out = im.out_geometry(F)
        assert t_out == golden[name + "/y"].shape[-1], name
        assert crop == (win // 2 if center else 0)
        assert torch.allclose(m.window, torch.from_numpy(golden[name + "/window"]), atol=5e-7)
    im = IMDCT4(1024, 512, 1024, kbdwin, True, out_length=130816, device="cpu")
    assert im.out_geometry(257) == (512, 130816)


def test_constructor_contract():
    with pytest.raises(AssertionError):
        MDCT4(512, 256, 1024, kbdwin, device="cpu")       # win > n_fft  (mdct.py:480)
    with pytest.raises(AssertionError):
        IMDCT4(512, 512, 256, kbdwin, device="cpu")       # hop > win    (mdct.py:481)
    w = kbdwin(256)
    m = MDCT4(256, 128, None, w, device="cpu")            # tensor window: win_length = len(window)
    assert m.win_length == 256
    assert MDCT4(256, 128, 200, None, device="cpu").window.sum() == 200    # None -> ones
    im = IMDCT4(256, 128, 256, kbdwin, device="cpu")
    with pytest.raises(AssertionError):
        im(torch.zeros(4, 128))                           # exactly 3-D (mdct.py:543)
    with pytest.raises(AssertionError):
        im(torch.zeros(1, 4, 256))                        # last dim = n_fft/2 (mdct.py:544)


def test_no_cpu_fallback():
    m = MDCT4(256, 128, 256, kbdwin, device="cpu")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(2, 1024))


def test_workspace_query():
    """p2p_workspace_bytes: sizes of the InstanceNorm workspace for N images x C channels; no GPU needed"""
    import ctypes
    from pix2pixhdaudiosr_b200 import _lib
    ws, cnt = ctypes.c_longlong(), ctypes.c_longlong()
    assert _lib.lib().p2p_workspace_bytes(16, 1536, ctypes.byref(ws), ctypes.byref(cnt)) == 0
    assert ws.value == 2 * 16 * 1536 * 4 and cnt.value == 16 * 12 * 4
    assert _lib.lib().p2p_workspace_bytes(-1, 8, ctypes.byref(ws), ctypes.byref(cnt)) == -1
