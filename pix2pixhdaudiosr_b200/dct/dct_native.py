"""Drop-in `DCT_2N_native` / `IDCT_2N_native` (reference: dct/dct_native.py:7-67) on the MDCT2 kernels of
csrc/mdct2.cu: a transform over the last dimension is the lapped transform at hop = win = n_fft with a window of ones.

    DCT_2N_native(x)[k]  = (2/N) sum_n x[n] cos(pi (2n+1) k / (2N))
    IDCT_2N_native(X)[n] = X[0] + 2 sum_{k>=1} X[k] cos(pi (2n+1) k / (2N))            (IDCT(DCT(x)) = 2x)

As in the reference the length N is fixed by the first call (dct_native.py:18-19).  float32 on a CUDA device; autograd."""
import torch

from ..models.mdct import _Analysis2Fn, _Plan2, _Synthesis2Fn


class _LastDim(torch.nn.Module):
    def __init__(self, expk=None):
        super().__init__()
        self.expk = expk            # accepted and ignored: the twiddles live in the kernel's tables
        self.N = None
        self._plan = None

    def _rows(self, x):
        if self.N is None:
            self.N = x.size(-1)
            self._plan = _Plan2(self.N, self.N, self.N, torch.ones(self.N))
        assert x.size(-1) == self.N, 'the transform length was fixed to %d by the first call' % self.N
        return x.reshape(1, -1)     # rows side by side: frame f of a hop = N framing is row f


class DCT_2N_native(_LastDim):
    def forward(self, x):
        rows = self._rows(x)
        y = _Analysis2Fn.apply(rows, self._plan, rows.shape[1] // self.N, 0, 1.0, 1.0)
        return y.reshape(x.shape)


class IDCT_2N_native(_LastDim):
    def forward(self, x):
        rows = self._rows(x)
        F = rows.shape[1] // self.N
        y = _Synthesis2Fn.apply(rows.reshape(1, F, self.N), self._plan, F * self.N, 0, 1.0, 0.5)
        return y.reshape(x.shape)
