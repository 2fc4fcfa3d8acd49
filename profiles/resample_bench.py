"""Timing of the device-side data path: p2p_resample on batches of 48 kHz segments, CUDA events, inputs
larger than L2.  python profiles/resample_bench.py"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pix2pixhdaudiosr_b200.data import resample as R      # noqa: E402


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    torch.cuda.synchronize()
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[reps // 2]


def main():
    B, T = 1024, 65024
    x = 0.1 * torch.randn(B, T, device="cuda")
    rows = []
    for orig, new in ((48000, 8000), (8000, 48000), (44100, 48000), (48000, 16000)):
        src = x if orig != 8000 else R.resample(x, 48000, 8000)
        ms = timed(lambda: R.resample(src, orig, new))
        out = R.resample(src, orig, new)
        nbytes = 4 * (src.numel() + out.numel())
        rows.append({"orig": orig, "new": new, "segments": B, "T_in": src.shape[1], "T_out": out.shape[1], "ms": round(ms, 4),
                     "GBps": round(nbytes / ms / 1e6, 1), "Msamples_out_per_s": round(out.numel() / ms / 1e3, 1)})
    ms = timed(lambda: R.make_lr_hr(x, 48000, 8000, 48000, T))
    item = {"make_lr_hr_ms": round(ms, 4), "segments_per_s": round(B / ms * 1e3, 1)}
    # the reference's CPU path for the same item (torchaudio on the host threads), bounded sample of 32 segments
    import torchaudio.functional as aF
    xc = x[:32].cpu()
    t0 = time.perf_counter()
    lr = aF.resample(aF.resample(xc, 48000, 8000), 8000, 48000)
    cpu_s = time.perf_counter() - t0
    item["cpu_torchaudio_segments_per_s"] = round(32 / cpu_s, 1)
    item["cpu_threads"] = torch.get_num_threads()
    print(json.dumps({"rows": rows, "item": item}))


if __name__ == "__main__":
    main()
