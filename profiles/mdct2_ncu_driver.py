"""Driver for ncu captures of the MDCT2 / IMDCT2 kernels: one n_fft (argv[1], default 512), 1024 segments of 130 816 samples.
    ncu --set full --clock-control none --import-source on -k regex:mdct2 -c 4 -o gpurun_out/r02_mdct2 python profiles/mdct2_ncu_driver.py 512
Without ncu it prints CUDA-event times (GB/s = algorithmic bytes / time; inputs exceed L2)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pix2pixhdaudiosr_b200.models import mdct as pm      # noqa: E402
from pix2pixhdaudiosr_b200.util.util import kbdwin        # noqa: E402

n_fft = int(sys.argv[1]) if len(sys.argv) > 1 else 512
B, T = 1024, 130816
dev = torch.device("cuda")
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6550.0
m = pm.MDCT2(n_fft, n_fft // 2, n_fft, kbdwin, True, device=dev)
im = pm.IMDCT2(n_fft, n_fft // 2, n_fft, kbdwin, True, device=dev)
x = torch.randn(B, T, device=dev)
spec = m(x)
y = im(spec)
torch.cuda.synchronize()


def tm(f, n=20):
    for _ in range(3):
        f()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    s.record()
    for _ in range(n):
        f()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


if os.environ.get("NCU_SHORT") != "1":
    ta, ts = tm(lambda: m(x)), tm(lambda: im(spec))
    ba, bs = 4 * (x.numel() + spec.numel()), 4 * (spec.numel() + y.numel())
    print(json.dumps({"n_fft": n_fft, "segments": B, "analysis_ms": round(ta, 4), "analysis_GBps": round(ba / ta / 1e6, 1),
                      "analysis_frac": round(ba / ta / 1e6 / peak, 3), "synthesis_ms": round(ts, 4),
                      "synthesis_GBps": round(bs / ts / 1e6, 1), "synthesis_frac": round(bs / ts / 1e6 / peak, 3)}))
