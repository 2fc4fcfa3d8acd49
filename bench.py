#!/usr/bin/env python
"""bench.py -- the driver's measurement contract for the hot path (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload train|train_time|mdct|global_g]

One JSON line on stdout (rank 0).
Workload "train" (default; BASELINE.json configs[2]): one step = one full G3L2-ngf48 optimisation step
(train.py:148-184) on 8 synthetic 48 kHz segments per GPU: MDCT4 -> dB/+- encoding -> LocalEnhancer ->
3 x MultiscaleDiscriminator -> LSGAN + feature matching -> two backward passes -> two Adam steps, bf16 storage with
fp32 accumulation; ranks are data-parallel replicas exchanging only the gradient all-reduce.  `value` = samples/s with
the audio resident in HBM; `e2e` = the same step fed from pinned HOST audio with the losses read back every step.
The line also carries "mdct4": the MDCT4 / IMDCT4 kernels' GB/s against the HBM peak (second half of the metric).
Workload "global_g" (BASELINE.json configs[1]): one step = one forward of define_G(1,1,64,'global',4,9) on [8,1,512,256]
(bench_global.py).
Workload "mdct": one step = MDCT4.forward + IMDCT4.forward over a batch of B synthetic 48 kHz segments (n_fft 1024,
hop 512, KBD window, center), inputs larger than L2 so no flush is needed.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "G3L2 train samples/s @1/2/4/8 B200; MDCT4/IMDCT4 GB/s vs HBM peak"
N_FFT, HOP, WIN, T_SEG = 1024, 512, 1024, 130816


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="train", choices=["train", "train_time", "mdct", "global_g"])
    ap.add_argument("--transform", default="mdct4", choices=["mdct4", "mdct2"],
                    help="train workloads: the lapped transform in front of the networks (mdct2 = the shipped wiring)")
    ap.add_argument("--batch", type=int, default=None, help="segments per GPU per step (train: 8, mdct: 512)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="train workload: launch the step eagerly instead of replaying a CUDA graph")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def tensor_peak():
    """bf16 dense TFLOP/s: the SUSTAINED figure (the conv kernels are timed inside a long step)"""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured (sustained)"
    return 1400.0, "fallback"


def tensor_peak_burst():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["bf16_tflops"])
    return 1590.0


def kept_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the ncu capture kept under profiles/
    (profiles/r02_dram_traffic.json, written by profiles/dram_traffic.py from an ncu pass over one step), or None"""
    p = os.path.join(ROOT, "profiles", "r02_dram_traffic.json")
    try:
        return json.load(open(p)).get(kernel)
    except (OSError, ValueError):
        return None


def mdct_bytes(B, T, F, M, T_out):
    """Algorithmic bytes: read signal + write coefficients; read coefficients + write audio."""
    analysis = B * (4 * T + 4 * F * M)
    synthesis = B * (4 * F * M + 4 * T_out)
    return analysis, synthesis


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [c.strip() for c in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- CPU arm: the oracle port of the reference algorithm on the host cores ---------------------------
def cpu_mdct_roundtrip(n_rows, threads, repeats=1):
    """Times oracle MDCT4.forward + IMDCT4.forward (reference algorithm, numpy/pocketfft) on n_rows
    segments split over `threads` host threads (numpy's FFT releases the GIL).  Returns seconds/pass."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from oracle import mdct4_oracle as O
    rng = np.random.default_rng(1234)
    x = (0.1 * rng.standard_normal((n_rows, T_SEG))).astype(np.float32)
    w = O.kbdwin(WIN)
    parts = [p for p in np.array_split(np.arange(n_rows), threads) if len(p)]

    def work(idx):
        s = O.mdct4_forward(x[idx], w, N_FFT, HOP, True, signal_len=512)   # same frame count as the GPU arm
        return O.imdct4_forward(s, w, N_FFT, HOP, True).shape

    best = None
    with ThreadPoolExecutor(max_workers=len(parts)) as ex:
        list(ex.map(work, parts[:1]))           # warm-up
        for _ in range(repeats):
            t0 = time.perf_counter()
            list(ex.map(work, parts))
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
    return best


def cpu_baseline_entry(n_rows=None):
    cores = os.cpu_count() or 1
    n_rows = n_rows or max(cores * 4, 32)
    sec = cpu_mdct_roundtrip(n_rows, cores, repeats=2)
    F = (T_SEG + 2 * HOP - WIN) // HOP + 1
    a, s = mdct_bytes(n_rows, T_SEG, F, N_FFT // 2, (F - 1) * HOP)
    return {"value": round((a + s) / sec / 1e9, 4), "unit": "GB/s", "cores": cores, "kind": "port",
            "sample": f"{n_rows} segments x {T_SEG} samples, MDCT4+IMDCT4 n_fft {N_FFT}, oracle/mdct4_oracle.py "
                      f"(numpy float64 restatement of models/mdct.py:486-566) on {cores} threads, best of 2",
            "seconds": round(sec, 4)}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for this path (the oracle port; the reference is
    Python and /root/reference does not exist on the GPU box), all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_rows = max(cores * 2, 16)
    F = (T_SEG + 2 * HOP - WIN) // HOP + 1
    a, s = mdct_bytes(n_rows, T_SEG, F, N_FFT // 2, (F - 1) * HOP)
    for _ in range(min(args.warmup, 1)):
        cpu_mdct_roundtrip(n_rows, cores)
    t = [cpu_mdct_roundtrip(n_rows, cores) for _ in range(args.steps)]
    sec = sum(t) / len(t)
    val = round((a + s) / sec / 1e9, 4)
    sample = (f"{n_rows} segments x {T_SEG} samples per step (bounded sample of the B={args.batch} workload), "
              f"oracle port on {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(sec * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"mdct4+imdct4 round trip n_fft={N_FFT} hop={HOP} KBD center, T={T_SEG}",
                   "segments_per_step": n_rows},
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ---- GPU arm ------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from pix2pixhdaudiosr_b200.models import mdct as pm
    from pix2pixhdaudiosr_b200.util.util import kbdwin

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    B, K, W = args.batch, args.steps, args.warmup
    m = pm.MDCT4(N_FFT, HOP, WIN, kbdwin, True, device=dev)
    im = pm.IMDCT4(N_FFT, HOP, WIN, kbdwin, True, device=dev)
    torch.manual_seed(1234 + rank)
    x_host = (0.1 * torch.randn(B, T_SEG)).pin_memory()
    x = x_host.to(dev)
    sp, ep = m.pads(B)
    F = (T_SEG + sp + ep - WIN) // HOP + 1
    M = N_FFT // 2
    crop, T_out = im.out_geometry(F)
    bytes_a, bytes_s = mdct_bytes(B, T_SEG, F, M, T_out)
    plan = m._plan

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        spec = pm._analysis(x, plan, F, sp, 1.0)
        return pm._synthesis(spec, plan, T_out, crop, 4.0 / N_FFT)

    # ---- device-resident timing: K steps, per-kernel CUDA events on the launching (current) stream
    for _ in range(W):
        y = step()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    t_begin = torch.cuda.Event(enable_timing=True); t_end = torch.cuda.Event(enable_timing=True)
    t_begin.record()
    for i in range(K):
        ev[i][0].record()
        spec = pm._analysis(x, plan, F, sp, 1.0)
        ev[i][1].record()
        y = pm._synthesis(spec, plan, T_out, crop, 4.0 / N_FFT)
        ev[i][2].record()
    t_end.record()
    barrier()
    clk = clocks.stop()
    total_ms = t_begin.elapsed_time(t_end)
    ms_a = sum(e[0].elapsed_time(e[1]) for e in ev) / K
    ms_s = sum(e[1].elapsed_time(e[2]) for e in ev) / K

    # ---- end to end through the public modules: pinned host in, host result out, every step
    y_host = torch.empty((B, 1, 1, T_out), dtype=torch.float32).pin_memory()
    def e2e_step():
        xd = x_host.to(dev, non_blocking=True)
        yd = im(m(xd))
        y_host.copy_(yd, non_blocking=True)
    for _ in range(min(W, 3)):
        e2e_step()
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    err = float((y_host[:, 0, 0, : 255 * HOP] - x_host[:, : 255 * HOP]).abs().max())

    times = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = times.tolist()
    if rank == 0:
        peak, peak_src = peaks()
        step_bytes = (bytes_a + bytes_s) * world
        dom = ("mdct4_analysis_kernel", bytes_a, ms_a) if ms_a >= ms_s else ("mdct4_synthesis_kernel", bytes_s, ms_s)
        ach = dom[1] / (dom[2] * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": round(step_bytes * K / (total_ms * 1e-3) / 1e9, 2), "unit": "GB/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(total_ms / K, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"mdct4+imdct4 round trip n_fft={N_FFT} hop={HOP} win={WIN} KBD center, "
                                   f"{B} segments x {T_SEG} samples per GPU per step",
                       "l2": "inputs larger than L2 (%.0f MB signal, %.0f MB coefficients per GPU); no flush" %
                             (B * T_SEG * 4 / 1e6, B * F * M * 4 / 1e6),
                       "segments_per_s": round(B * world * K / (total_ms * 1e-3), 1),
                       "roundtrip_max_abs_err": err},
            "roofline": {"bound": "hbm", "kernel": dom[0], "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(ach / peak, 4), "peak_source": peak_src, "traffic": None,
                         "bytes_per_launch": dom[1], "ms_per_launch": round(dom[2], 4)},
            "kernels": {"mdct4_analysis_kernel": {"ms": round(ms_a, 4), "GBps": round(bytes_a / ms_a / 1e6, 1)},
                        "mdct4_synthesis_kernel": {"ms": round(ms_s, 4), "GBps": round(bytes_s / ms_s / 1e6, 1)}},
            "e2e": {"value": round(step_bytes * K / (e2e_ms * 1e-3) / 1e9, 2), "unit": "GB/s",
                    "h2d_bytes_per_step": B * T_SEG * 4 * world, "d2h_bytes_per_step": B * T_out * 4 * world,
                    "ms_per_step": round(e2e_ms / K, 3)},
            "gpu_launches": 2 * K, "clocks": clk,
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline_entry()
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    # Native libraries write to file descriptor 1 (NCCL prints "NCCL version ..." there): keep stdout = ONE JSON line by
    # pointing fd 1 at stderr and giving Python's own stdout a private copy of the original descriptor
    sys.stdout.flush()
    sys.stdout = os.fdopen(os.dup(1), "w", buffering=1)
    os.dup2(2, 1)
    if args.batch is None:
        args.batch = 8 if args.workload.startswith("train") or args.workload == "global_g" else 512
    if args.workload == "global_g":
        from bench_global import run_global, run_global_reference
        (run_global_reference if args.impl == "reference" else run_global)(args)
    elif args.workload.startswith("train"):
        from bench_train import run_train, run_train_reference
        (run_train_reference if args.impl == "reference" else run_train)(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
