"""Golden vectors for MDCT2 / IMDCT2 / DCT_2N_native / IDCT_2N_native from the REFERENCE itself, run on
CPU in the build container.   python tests/golden/make_golden_mdct2.py -> tests/golden/mdct2_ref.npz
Outputs and the autograd gradients (torch autograd through the reference's own lines)."""
import os

import numpy as np
import torch

from make_golden import import_reference

HERE = os.path.dirname(os.path.abspath(__file__))
# name: (B, T, n_fft, hop, win, center, kbd window?)
CASES = {
    "model": (2, 63 * 64, 128, 64, 128, True, True),          # the miniature model's geometry
    "recipe": (2, 4000, 512, 256, 512, True, True),           # generated/GEN_VCTK_G3L2_48ngf/opt.txt: n_fft 512 hop 256
    "nocenter": (3, 1500, 256, 128, 256, False, True),
    "hop_quarter": (2, 2000, 256, 64, 256, True, True),
    "win_lt_nfft": (2, 1800, 256, 96, 192, True, False),
    "n1024": (1, 6000, 1024, 512, 1024, True, True),
    "n16": (8, 100, 16, 8, 16, True, True),
}


def main():
    _, _, kbdwin = import_reference()
    from dct.dct_native import DCT_2N_native, IDCT_2N_native
    from models.mdct import IMDCT2, MDCT2
    torch.manual_seed(4321)
    out = {}
    x16 = torch.arange(1, 17, dtype=torch.float64)
    out["known/dct_arange"] = DCT_2N_native()(x16.clone()).numpy()            # test/DCT_test.ipynb cell 34
    out["known/idct_dct_arange"] = IDCT_2N_native()(DCT_2N_native()(x16.clone())).numpy()
    for N in (8, 64, 512):
        x = torch.randn(3, 5, N, dtype=torch.float64, requires_grad=True)
        y = DCT_2N_native()(x)
        g = torch.randn_like(y)
        (gx,) = torch.autograd.grad(y, x, g)
        X = torch.randn(3, 5, N, dtype=torch.float64, requires_grad=True)
        yi = IDCT_2N_native()(X)
        (gX,) = torch.autograd.grad(yi, X, g)
        for k, v in (("x", x), ("dct", y), ("g", g), ("dct_gx", gx), ("X", X), ("idct", yi), ("idct_gX", gX)):
            out[f"dct{N}/{k}"] = v.detach().numpy()
    for name, (B, T, n_fft, hop, win, center, kbd) in CASES.items():
        window = kbdwin if kbd else None
        m = MDCT2(n_fft=n_fft, hop_length=hop, win_length=win, window=window, center=center, device="cpu", dct_op=DCT_2N_native())
        im = IMDCT2(n_fft=n_fft, hop_length=hop, win_length=win, window=window, center=center, device="cpu", idct_op=IDCT_2N_native())
        x = (0.1 * torch.randn(B, T)).float().requires_grad_(True)
        spec, frames = m(x, True)
        gs = torch.randn_like(spec)
        (gx,) = torch.autograd.grad(spec, x, gs)
        spec_in = spec.detach().clone().requires_grad_(True)
        y = im(spec_in)
        gy = torch.randn_like(y)
        (gspec,) = torch.autograd.grad(y, spec_in, gy)
        for k, v in (("x", x), ("spec", spec), ("frames", frames), ("gs", gs), ("gx", gx), ("y", y), ("gy", gy), ("gspec", gspec)):
            out[f"{name}/{k}"] = v.detach().numpy()
        print(name, tuple(spec.shape), tuple(y.shape), float(((y[:, 0, 0, :T] - x[:, :y.shape[-1]]) ** 2).mean()) if hop * 2 == win and kbd else "")
    np.savez_compressed(os.path.join(HERE, "mdct2_ref.npz"), **out)


if __name__ == "__main__":
    main()
