"""BASELINE.json configs[4]: MDCT4 / IMDCT4 throughput sweep, n_fft 256-4096 (hop = n_fft/2, KBD, center), batch 1-1024
segments of 130 816 samples, fp32.   python profiles/mdct_sweep.py            (1 GPU)
                                     torchrun --nproc-per-node 8 profiles/mdct_sweep.py   (batch split over the ranks)
Each launch is CUDA-event timed; between timed launches a 256 MB buffer is rewritten (flushes the 126 MB L2) unless
the launch's own data exceeds L2.  GB/s = algorithmic bytes / time; the multi-GPU figure sums the ranks'
bytes over the slowest rank's time (segments are independent: no collective on the data path)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pix2pixhdaudiosr_b200.models import mdct as pm      # noqa: E402
from pix2pixhdaudiosr_b200.util.util import kbdwin        # noqa: E402

T = 130816
L2_BYTES = 126e6


def main():
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6550.0
    flush = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    rows = []
    mdct2 = "--mdct2" in sys.argv               # the DCT-II lapped transform (MDCT2 / IMDCT2): n_fft bins per frame
    bins = (lambda n: n) if mdct2 else (lambda n: n // 2)
    ana = (lambda x, plan, F, sp: pm._analysis2(x, plan, F, sp, 1.0, 1.0)) if mdct2 else (lambda x, plan, F, sp: pm._analysis(x, plan, F, sp, 1.0))
    syn = ((lambda s, plan, T_out, crop, n: pm._synthesis2(s, plan, T_out, crop, 0.5, 0.5)) if mdct2
           else (lambda s, plan, T_out, crop, n: pm._synthesis(s, plan, T_out, crop, 4.0 / n)))
    for n_fft in (256, 512, 1024, 2048, 4096):
        hop = n_fft // 2
        m = (pm.MDCT2 if mdct2 else pm.MDCT4)(n_fft, hop, n_fft, kbdwin, True, device=dev)
        im = (pm.IMDCT2 if mdct2 else pm.IMDCT4)(n_fft, hop, n_fft, kbdwin, True, device=dev)
        sizes = (1, 16, 256, 1024) if mdct2 else (1, 4, 16, 64, 256, 1024)
        if world > 1:
            sizes = sizes + (1024 * world,)          # the per-GPU plateau size on every rank
        for B_total in sizes:
            B = max(1, B_total // world)            # per rank (B_total < world: every rank runs one segment)
            x = 0.1 * torch.randn(B, T, device=dev)
            sp, ep = m.pads(B)
            F = (T + sp + ep - n_fft) // hop + 1
            crop, T_out = im.out_geometry(F)
            ba = B * (4 * T + 4 * F * bins(n_fft))
            bs = B * (4 * F * bins(n_fft) + 4 * T_out)
            reps = 20 if ba < 64e6 else 8
            need_flush = ba < 2 * L2_BYTES
            for _ in range(3):
                spec = ana(x, m._plan, F, sp)
                y = syn(spec, m._plan, T_out, crop, n_fft)
            ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(reps)]
            torch.cuda.synchronize()
            for e in ev:
                if need_flush:
                    flush.fill_(1.0)
                e[0].record()
                spec = ana(x, m._plan, F, sp)
                e[1].record()
                if need_flush:
                    flush.fill_(2.0)
                e[2].record()
                y = syn(spec, m._plan, T_out, crop, n_fft)
                e[3].record()
            torch.cuda.synchronize()
            ms = torch.tensor([sorted(e[0].elapsed_time(e[1]) for e in ev)[reps // 2],
                               sorted(e[2].elapsed_time(e[3]) for e in ev)[reps // 2]], dtype=torch.float64, device=dev)
            nb = torch.tensor([ba, bs], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                dist.all_reduce(nb, op=dist.ReduceOp.SUM)
            err = float((y[:, : (F - 2) * hop] - x[:, : (F - 2) * hop]).abs().max())
            rows.append({"n_fft": n_fft, "segments": B * world, "ms_analysis": round(ms[0].item(), 4), "ms_synthesis": round(ms[1].item(), 4),
                         "GBps_analysis": round(nb[0].item() / ms[0].item() / 1e6, 1), "GBps_synthesis": round(nb[1].item() / ms[1].item() / 1e6, 1),
                         "l2_flushed": need_flush, "roundtrip_max_abs_err": err})
            del x, spec, y
    if rank == 0:
        print(json.dumps({"transform": "mdct2" if mdct2 else "mdct4", "n_gpus": world, "T": T, "hbm_peak_GBps_per_gpu": peak, "rows": rows}))
        print("| n_fft | segments | MDCT4 ms | GB/s (frac of %d-GPU peak) | IMDCT4 ms | GB/s (frac) |" % world, file=sys.stderr)
        print("|---|---|---|---|---|---|", file=sys.stderr)
        for r in rows:
            print("| %d | %d | %.4f | %.0f (%.2f) | %.4f | %.0f (%.2f) |" % (
                r["n_fft"], r["segments"], r["ms_analysis"], r["GBps_analysis"], r["GBps_analysis"] / (peak * world),
                r["ms_synthesis"], r["GBps_synthesis"], r["GBps_synthesis"] / (peak * world)), file=sys.stderr)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
