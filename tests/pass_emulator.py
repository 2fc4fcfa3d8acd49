"""numpy emulation of one p2p_conv_pass (include/p2p_b200.h) -- checks the HOST-side pass builders on CPU."""
import numpy as np


def _pad(i, n, mode):
    if mode == 1:
        i = -i if i < 0 else (2 * (n - 1) - i if i >= n else i)
    return i if 0 <= i < n else -1


def run_pass(P, x, wflat, bias, out):
    """x [N,H,W,K], wflat 1-D physical weight buffer, out [N,OH,OW,J] written in place"""
    for y in range(P.gy):
        for xx in range(P.gx):
            acc = np.zeros((P.N, P.J))
            for t in range(P.ntaps):
                ih = _pad(y * P.sy + P.dh[t], P.H, P.pad_mode)
                iw = _pad(xx * P.sx + P.dw[t], P.W, P.pad_mode)
                if ih < 0 or iw < 0:
                    continue
                k = np.arange(P.K)[:, None]
                j = np.arange(P.J)[None, :]
                wt = wflat[P.wtap[t] * P.w_tap + k * P.w_k + j * P.w_j]        # [K, J]
                acc += x[:, ih, iw, :] @ wt
            if bias is not None:
                acc += bias
            out[:, y * P.oys + P.oyo, xx * P.oxs + P.oxo, :] = acc


def run_wgrad(P, x, dout, dwflat):
    for y in range(P.gy):
        for xx in range(P.gx):
            d = dout[:, y * P.oys + P.oyo, xx * P.oxs + P.oxo, :]             # [N, J]
            for t in range(P.ntaps):
                ih = _pad(y * P.sy + P.dh[t], P.H, P.pad_mode)
                iw = _pad(xx * P.sx + P.dw[t], P.W, P.pad_mode)
                if ih < 0 or iw < 0:
                    continue
                k = np.arange(P.K)[:, None]
                j = np.arange(P.J)[None, :]
                np.add.at(dwflat, P.wtap[t] * P.w_tap + k * P.w_k + j * P.w_j, x[:, ih, iw, :].T @ d)


def fold_reflect(dxp, p):
    N, HP, WP, C = dxp.shape
    H, W = HP - 2 * p, WP - 2 * p
    dx = np.zeros((N, H, W, C))
    for q in range(HP):
        h = _pad(q - p, H, 1)
        for r in range(WP):
            w = _pad(r - p, W, 1)
            dx[:, h, w, :] += dxp[:, q, r, :]
    return dx
