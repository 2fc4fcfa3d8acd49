"""Generate golden vectors by running the REFERENCE itself (imported from /root/reference, CPU).

Run once in the build container:  python tests/golden/make_golden.py
/root/reference does not exist on the GPU box, so the outputs are committed as small .npz files.
Stub set needed to import the reference: pysepm, matplotlib(.pyplot), turtle.forward.
"""
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("P2P_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def import_reference():
    for name in ("pysepm", "matplotlib", "matplotlib.pyplot", "turtle"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["turtle"].forward = lambda *a, **k: None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from models.mdct import MDCT4, IMDCT4  # noqa
    from util.util import kbdwin            # noqa
    return MDCT4, IMDCT4, kbdwin


def mdct_cases():
    # (name, B, T, n_fft, hop, win, center, window kind, 1-D input?)
    return [
        ("c1_small", 2, 3000, 1024, 512, 1024, True, "kbd", False),     # config-1 geometry, short
        ("n256", 2, 3001, 256, 128, 256, True, "kbd", False),           # ragged T
        ("n512_nocenter", 2, 2048, 512, 256, 512, False, "kbd", False),
        ("n2048", 1, 6000, 2048, 1024, 2048, True, "kbd", True),        # 1-D input path (len = T)
        ("n4096", 1, 12000, 4096, 2048, 4096, True, "kbd", False),
        ("hop_quarter", 2, 3000, 512, 128, 512, True, "kbd", False),    # hop < win/2
        ("win_lt_nfft", 2, 2500, 512, 192, 384, True, "ones", False),   # win < n_fft, window=None -> ones
        ("batch_quirk", 8, 100, 16, 8, 16, True, "kbd", False),         # H8(a): B % hop == 0, T % hop != 0
        ("n64", 5, 777, 64, 32, 64, True, "kbd", False),
    ]


def main():
    MDCT4, IMDCT4, kbdwin = import_reference()
    torch.manual_seed(1234)
    out = {}
    for name, B, T, n_fft, hop, win, center, wk, one_d in mdct_cases():
        window = kbdwin if wk == "kbd" else None
        m = MDCT4(n_fft=n_fft, hop_length=hop, win_length=win, window=window, center=center, device="cpu")
        im = IMDCT4(n_fft=n_fft, hop_length=hop, win_length=win, window=window, center=center, device="cpu")
        x = (0.1 * torch.randn(T if one_d else (B, T))).float().requires_grad_(True)
        spec = m(x)                                   # float64
        gs = torch.randn_like(spec)
        (gx,) = torch.autograd.grad(spec, x, gs)      # MDCT4 backward
        spec_in = spec.detach().clone()
        if spec_in.dim() == 2:
            spec_in = spec_in[None]
        spec_in.requires_grad_(True)
        y = im(spec_in.clone())                       # clone: reference transposes in place
        gy = torch.randn_like(y)
        (gspec,) = torch.autograd.grad(y, spec_in, gy)
        out[name + "/x"] = x.detach().numpy()
        out[name + "/window"] = m.window.numpy()
        out[name + "/spec"] = spec.detach().numpy()
        out[name + "/gs"] = gs.numpy()
        out[name + "/gx"] = gx.numpy()
        out[name + "/y"] = y.detach().numpy()
        out[name + "/gy"] = gy.numpy()
        out[name + "/gspec"] = gspec.numpy()
        print(name, tuple(x.shape), "->", tuple(spec.shape), "->", tuple(y.shape), spec.dtype, gx.dtype)
    np.savez_compressed(os.path.join(HERE, "mdct4_ref.npz"), **out)

    # README.md:93-118 / DCT_test.ipynb cell 32: 130816 samples, n_fft 1024, hop 512 -> [257,512]
    torch.manual_seed(1234)
    x = 0.1 * torch.randn(16, 130816)
    m = MDCT4(n_fft=1024, hop_length=512, win_length=1024, window=kbdwin, center=True, device="cpu")
    im = IMDCT4(n_fft=1024, hop_length=512, win_length=1024, window=kbdwin, center=True, device="cpu")
    s = m(x)
    y = im(s)
    mse = float(((x.double() - y[:, 0, 0, :130816]) ** 2).mean())
    one = m(x[0])
    # keep a strided sample of coefficients so the fixture stays small
    np.savez_compressed(
        os.path.join(HERE, "mdct4_config1.npz"),
        spec_shape=np.array(s.shape), one_d_shape=np.array(one.shape), y_shape=np.array(y.shape),
        mse=np.array(mse), spec_rows=s[::5, ::37, :].numpy(), y_sample=y[:, 0, 0, ::101].numpy(),
        kbd1024=kbdwin(1024).numpy(), kbd16=kbdwin(16).numpy(),
    )
    print("config1", tuple(s.shape), tuple(one.shape), tuple(y.shape), "mse", mse)


if __name__ == "__main__":
    main()
