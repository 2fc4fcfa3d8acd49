"""Mirror of the reference's util/util.py for the symbols on the hot path."""
import torch


def kbdwin(N: int, beta: float = 12.0, device='cpu') -> torch.Tensor:
    """MATLAB-style Kaiser-Bessel-derived window; same arithmetic (fp32, beta*pi, non-periodic
    kaiser of N/2+1 points, sqrt of the normalised cumulative sum, mirrored) as the reference's
    util/util.py:186-193, so w[n]^2 + w[n+N/2]^2 = 1.  Host-side, computed once per module."""
    assert N % 2 == 0, "N must be even"
    kaiser = torch.kaiser_window(window_length=N // 2 + 1, beta=beta * torch.pi, periodic=False, device=device)
    rising = torch.sqrt(torch.cumsum(kaiser, dim=0) / kaiser.sum())[:-1]
    return torch.cat((rising, rising.flip(dims=(0,))), dim=0)
