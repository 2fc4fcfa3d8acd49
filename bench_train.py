"""The "train" workload of bench.py: one G3L2-ngf48 optimisation step per bench step (see bench.py's docstring)."""
import json
import os
import time

METRIC = "G3L2 train samples/s @1/2/4/8 B200; MDCT4/IMDCT4 GB/s vs HBM peak"
T_SEG = 127 * 512          # 65 024 samples -> 128 frames of 512 bins (S128 of SURVEY 8: [B, 2, 512, 128])
WORKLOAD = ("G3L2 ngf48 LocalEnhancer + 2-scale D (ndf 64), GAN + feature matching, full optimisation step "
            "(train.py:148-184), batch %d x 65024-sample 48 kHz segments per GPU -> [%d,2,512,128] MDCT4 spectrograms")


def workload_name(args):
    """the `config.workload` string, identical in both arms"""
    B = args.batch
    name = WORKLOAD % (B, B)
    if args.workload == "train_time":
        name += ("; + time-domain D on [%d,2,128,%d] frames and frame-match loss "
                 "(--use_time_D --lambda_time 10 --use_match_loss)" % (B, 512 if getattr(args, "transform", "mdct4") == "mdct2" else 1024))
    return name


def synth_audio(B, seed, T=None):
    import torch
    g = torch.Generator().manual_seed(seed)
    hr = 0.1 * torch.randn(B, T or T_SEG, generator=g)
    lr = 0.5 * (hr + torch.roll(hr, 1, dims=1))             # a crude low-pass stands in for the 8 kHz resample
    return lr, hr


def _mdct_roofline(dev, steps=10, B=512, T=130816):
    """MDCT4 / IMDCT4 kernels alone (second half of the metric): algorithmic bytes / CUDA-event time"""
    import torch
    from bench import HOP, N_FFT, WIN, mdct_bytes, peaks
    from pix2pixhdaudiosr_b200.models import mdct as pm
    from pix2pixhdaudiosr_b200.util.util import kbdwin
    m = pm.MDCT4(N_FFT, HOP, WIN, kbdwin, True, device=dev)
    im = pm.IMDCT4(N_FFT, HOP, WIN, kbdwin, True, device=dev)
    x = 0.1 * torch.randn(B, T, device=dev)
    sp, ep = m.pads(B)
    F = (T + sp + ep - WIN) // HOP + 1
    crop, T_out = im.out_geometry(F)
    ba, bs = mdct_bytes(B, T, F, N_FFT // 2, T_out)
    for _ in range(3):
        y = pm._synthesis(pm._analysis(x, m._plan, F, sp, 1.0), m._plan, T_out, crop, 4.0 / N_FFT)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
    torch.cuda.synchronize()
    for e in ev:
        e[0].record()
        spec = pm._analysis(x, m._plan, F, sp, 1.0)
        e[1].record()
        y = pm._synthesis(spec, m._plan, T_out, crop, 4.0 / N_FFT)
        e[2].record()
    torch.cuda.synchronize()
    ms_a = sum(e[0].elapsed_time(e[1]) for e in ev) / steps
    ms_s = sum(e[1].elapsed_time(e[2]) for e in ev) / steps
    peak, src = peaks()
    del x, spec, y
    return {"workload": f"{B} segments x {T} samples, n_fft {N_FFT} hop {HOP} KBD center; inputs larger than L2",
            "analysis": {"ms": round(ms_a, 4), "GBps": round(ba / ms_a / 1e6, 1), "frac_hbm_peak": round(ba / ms_a / 1e6 / peak, 4)},
            "synthesis": {"ms": round(ms_s, 4), "GBps": round(bs / ms_s / 1e6, 1), "frac_hbm_peak": round(bs / ms_s / 1e6 / peak, 4)},
            "hbm_peak_GBps": peak, "peak_source": src}


def cpu_train_step_seconds(batch=1, steps=1, workload="train", transform="mdct4"):
    """the oracle port of the reference's optimisation step (oracle/pix2pixhd_oracle.py: plain torch CPU ops on
    reference-format state_dicts) on all host threads; returns (seconds per step, threads)"""
    import torch
    from oracle import pix2pixhd_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = {'n_fft': 1024, 'hop_length': 512, 'win_length': 1024, 'alpha': 0.6, 'min_value': 1e-7, 'explicit_encoding': True,
           'up_ratio': 6.0, 'lambda_feat': 10.0,
           'G': {'netG': 'local', 'n_downsample_global': 4, 'n_blocks_global': 3, 'n_local_enhancers': 1, 'n_blocks_local': 2},
           'D': {'num_D': 2, 'n_layers_D': 3}}
    sdT, t_seg = None, None
    if transform == "mdct2":                # the shipped wiring: n_fft 512 / hop 256 on 32 512-sample segments
        cfg.update(transform="mdct2", n_fft=512, hop_length=256, win_length=512)
        t_seg = 127 * 256
    if workload == "train_time":            # config 4: + time-domain D on frames + frame-match loss
        cfg.update(use_time_D=True, use_match_loss=True, lambda_time=10.0, lambda_mat=10.0)
        sdT = O.random_time_D_state_dict()
    sdG, sdD = O.random_state_dicts()
    st = O.TrainState(sdG, sdD, cfg, sdT=sdT)
    lr, hr = synth_audio(batch, 7, t_seg)
    t = []
    for _ in range(steps):
        t0 = time.perf_counter()
        st.step(lr, hr)
        t.append(time.perf_counter() - t0)
    return sum(t) / len(t), threads


def reference_available(workload, transform):
    """the reference's own modules (oracle/_ref, vendored by oracle/build_ref.py) can run this workload on the host cores"""
    from oracle import ref_runner
    # the reference wires its time-domain path to MDCT2 only: config 4 on MDCT4 frames exists only in the port
    return ref_runner.available() and not (workload == "train_time" and transform != "mdct2")


def cpu_reference_steps(batch, steps, warmup, workload="train", transform="mdct4", budget_s=200.0):
    """the REFERENCE's own Pix2PixHDModel (oracle/_ref) on all host threads at the bench's batch size:
    -> (seconds per step, timed steps, warm-up steps, threads).  K and W are honoured when the run fits `budget_s`
    (estimated from the first step), else cut to fit."""
    from oracle import ref_runner
    t_seg = 127 * 256 if transform == "mdct2" else T_SEG
    tr = ref_runner.ReferenceTrainer(batch=batch, transform=transform, time_d=(workload == "train_time"))
    lr, hr = synth_audio(batch, 7, t_seg)
    first = tr.time_steps(lr, hr, 1)[0]                       # warm-up step 1 (allocator, thread pool), also the estimate
    w_done = 1
    if warmup > 1 and first * (warmup - 1 + steps) <= budget_s:
        tr.time_steps(lr, hr, warmup - 1)
        w_done = warmup
    k = steps if first * steps <= budget_s else max(2, int(budget_s / first))
    t = tr.time_steps(lr, hr, k)
    return sum(t) / len(t), len(t), w_done, tr.threads


def run_train_reference(args):
    """--impl reference: the reference's CPU implementation of the step on the box's host cores, all threads.  With
    oracle/_ref in place (oracle/build_ref.py) it is the reference's own, unmodified Pix2PixHDModel at the bench's batch size
    (cpu_baseline.kind "reference"); otherwise the oracle port at batch 1 (kind "port")."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    transform = getattr(args, "transform", "mdct4")
    if reference_available(args.workload, transform):
        sec, steps, warm, threads = cpu_reference_steps(args.batch, args.steps, args.warmup, args.workload, transform)
        batch, kind = args.batch, "reference"
        sample = (f"the reference's own Pix2PixHDModel (oracle/_ref, unmodified; fp32, --gpu_ids -1) at batch {batch}, "
                  f"{steps} timed step(s) of the requested {args.steps} after {warm} warm-up, {threads} threads")
    else:
        steps = max(1, min(args.steps, 6))
        warm = min(args.warmup, 1)
        sec, threads = cpu_train_step_seconds(batch=1, steps=steps + warm, workload=args.workload, transform=transform)
        batch, kind = 1, "port"
        sample = (f"batch 1 (bounded sample of the batch-{args.batch} step), {steps} timed step(s) of the requested "
                  f"{args.steps}, oracle/pix2pixhd_oracle.py on {threads} threads (oracle/_ref absent or workload not wired in the reference)")
    val = round(batch / sec, 5)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "samples/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": round(sec * 1e3, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "transform": transform, "global_batch": batch, "sample": sample},
        "cpu_baseline": {"value": val, "unit": "samples/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def run_train(args):
    import torch
    import torch.distributed as dist
    from bench import ClockSampler, kept_traffic, tensor_peak, tensor_peak_burst
    from pix2pixhdaudiosr_b200 import _lib, ops, options
    from pix2pixhdaudiosr_b200.models.models import create_model
    from pix2pixhdaudiosr_b200.train import GradSync, GraphedTrainStep, broadcast_parameters, replica_drift, train_step

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
    lib = _lib.lib()

    import contextlib
    import io
    torch.manual_seed(1234)
    with contextlib.redirect_stdout(io.StringIO()):
        # --workload train_time: BASELINE.json configs[3] (train_script.sh:93, 96: --use_time_D --lambda_time 10 --use_match_loss)
        extra = dict(use_time_D=True, use_match_loss=True, lambda_time=10.0) if args.workload == "train_time" else {}
        t_seg = T_SEG
        if getattr(args, "transform", "mdct4") == "mdct2":
            # the recipe exactly as generated/GEN_VCTK_G3L2_48ngf/opt.txt wires it: MDCT2 n_fft 512 / hop 256 on 32 512-sample
            # segments -> the same [B,2,512,128] image
            t_seg = 127 * 256
            extra.update(transform="mdct2", n_fft=512, hop_length=256, win_length=512, segment_length=t_seg)
        model = create_model(options.g3l2_opt(gpu_ids=[local], batchSize=B, precision=args.precision, **extra))
    broadcast_parameters(model)
    # bf16: sharded data parallelism (reduce-scatter, Adam on 1/world of the parameters per rank, all-gather of the bf16 shadow);
    # P2P_DP_SHARD=0 keeps one all-reduce of the fp32 gradients per network (round 1)
    sync = GradSync(shard=args.precision == "bf16" and os.environ.get("P2P_DP_SHARD", "1") == "1")
    if os.environ.get("P2P_EARLY_SYNC", "0") == "1":      # measured slower at 8 GPUs (12.7 vs 12.1 ms/step): off by default
        sync.attach_early(model.optimizer_G)
    lr_host, hr_host = synth_audio(B, 1000 + rank, t_seg)
    lr_host, hr_host = lr_host.pin_memory(), hr_host.pin_memory()
    lr_d, hr_d = lr_host.to(dev), hr_host.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.no_graph:
        def step(l, h):
            return train_step(model, l, h, sync=sync)
    else:
        step = GraphedTrainStep(model, lr_d, hr_d, sync=sync)
        lr_d, hr_d = step.lr_static, step.hr_static

    # ---- device-resident: K full optimisation steps
    for _ in range(W):
        loss_dict, _ = step(lr_d, hr_d)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(K):
        loss_dict, _ = step(lr_d, hr_d)
    t1.record()
    barrier()
    clk = clocks.stop()
    total_ms = t0.elapsed_time(t1)
    final_losses = {k: float(v) for k, v in loss_dict.items()}

    # ---- end to end: pinned host audio in, losses out to the host, every step
    names = list(loss_dict)
    loss_host = torch.empty(len(names), dtype=torch.float32).pin_memory()

    def e2e_step():
        if args.no_graph:
            l, h = lr_host.to(dev, non_blocking=True), hr_host.to(dev, non_blocking=True)
        else:
            l, h = lr_d.copy_(lr_host, non_blocking=True), hr_d.copy_(hr_host, non_blocking=True)
        ld, _ = step(l, h)
        loss_host.copy_(torch.stack([ld[n].float() for n in names]), non_blocking=True)

    for _ in range(min(W, 2)):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)

    # ---- one extra, un-timed step with CUDA events around every convolution launch: the roofline leg
    n0 = lib.p2p_kernel_launches()
    torch.cuda.synchronize(dev)
    if args.no_graph:
        torch.cuda._sleep(int(0.035 * 1.9e9))     # ~35 ms of device spin: the host enqueues the eager step behind it
        ops.start_conv_timing()
        train_step(model, lr_d, hr_d, sync=sync)
    else:
        # the profiled step is itself captured in a CUDA graph with the event pairs as record nodes and replayed once: every
        # kernel runs back to back exactly as in the timed region (an eager step outruns the host's launch rate, and the launch
        # queue holds only ~1000 of its ~1300 launches, so event pairs late in an eager step include host gaps)
        # (with the weight gradients IN the main stream: in the timed step they run on a side stream next to the memory-bound
        # kernels, which speeds the step up but makes an event pair measure two kernels sharing the SMs -- the roofline leg
        # wants each kernel's own duration)
        pg = torch.cuda.CUDAGraph()
        from pix2pixhdaudiosr_b200 import train as _train
        overlap0, ops.WGRAD_OVERLAP = ops.WGRAD_OVERLAP, False
        gstep0, _train.OVERLAP_G_STEP = _train.OVERLAP_G_STEP, False      # (likewise the generator's Adam next to D's backward)
        ops.start_conv_timing(external=True)
        try:
            with torch.cuda.graph(pg):
                train_step(model, lr_d, hr_d, sync=sync)
        finally:
            ops.WGRAD_OVERLAP = overlap0
            _train.OVERLAP_G_STEP = gstep0
        ops.pause_conv_timing()
        pg.replay()
        torch.cuda.synchronize(dev)
        # a second profiled replay with an event pair around EVERY C-ABI entry point: the step's breakdown
        pg2 = torch.cuda.CUDAGraph()
        _lib.start_call_profile(external=True)
        with torch.cuda.graph(pg2):
            train_step(model, lr_d, hr_d, sync=sync)
        _lib.pause_call_profile()
        pg2.replay()
        _lib.resume_call_profile()
        calls = _lib.stop_call_profile()
        del pg2
    rec = ops.stop_conv_timing()
    launches = (lib.p2p_kernel_launches() - n0) * K       # kernels of ours per step (counted eagerly) x timed steps
    if args.no_graph:
        calls = {}

    drift = replica_drift(model)          # 0 = every rank holds bit-identical bf16 weights after all the steps above
    times = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = times.tolist()
    if rank == 0:
        dump = os.environ.get("P2P_BENCH_DUMP_LAUNCHES")
        if dump:        # per-launch table of the profiled step: kind, GEMM shape, algorithmic GFLOP, ms, TFLOP/s
            with open(dump, "w") as f:
                for kind, umma, flops, ms, shape in rec:
                    f.write("%-14s %s M=%-7d J=%-5d K=%-5d taps=%-3d HxW=%dx%d  %8.2f GF %8.4f ms %7.1f TF\n" % (
                        kind, "tc" if umma else "cc", shape["M"], shape["J"], shape["K"], shape["taps"], shape["H"], shape["W"],
                        flops / 1e9, ms, flops / max(ms, 1e-9) / 1e9))
        peak, peak_src = tensor_peak()
        step_ms = total_ms / K
        classes = {}
        for kind, umma, flops, ms, _ in rec:
            c = classes.setdefault(("tcgen05 " if umma else "cuda-core ") + kind, [0, 0.0, 0.0])
            c[0] += 1; c[1] += flops; c[2] += ms
        um_flops = sum(r[2] for r in rec if r[1])
        um_ms = sum(r[3] for r in rec if r[1])
        um_n = sum(1 for r in rec if r[1])
        all_flops = sum(r[2] for r in rec)
        ach = um_flops / (um_ms * 1e-3) / 1e12 if um_ms > 0 else 0.0
        # the dominant kernel: conv_window_kernel (shifted-window implicit GEMM, forward and input-gradient passes)
        dom = [r for r in rec if r[1] and r[0] in ("fwd-window", "dgrad-window")]
        dom_flops, dom_ms, dom_n = sum(r[2] for r in dom), sum(r[3] for r in dom), len(dom)
        dom_ach = dom_flops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else 0.0
        out = {
            "metric": METRIC, "value": round(B * world * K / (total_ms * 1e-3), 2), "unit": "samples/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": workload_name(args),
                       "transform": getattr(args, "transform", "mdct4") + (" (n_fft 512, hop 256, 32512-sample segments)" if t_seg != T_SEG else ""),
                       "global_batch": B * world, "parallelism": f"dp{world}" + (" (reduce-scatter + sharded Adam + all-gather)" if sync.shard else ""),
                       "replica_drift": drift,
                       "launch": "eager" if args.no_graph else "one CUDA graph replay per step",
                       "l2": "per-step working set (0.31 GB bf16 weights + 0.65 GB fp32 master + activations) exceeds the 126 MB L2; no flush",
                       "weights": "random init N(0,0.02), seed 1234; mask_mode mode2; Adam lr 2e-4 beta1 0.5",
                       "final_losses": final_losses},
            "roofline": {"bound": "tensor", "kernel": "conv_window_kernel (TMA-fed tcgen05 shifted-window implicit GEMM: forward + input-gradient passes)",
                         "achieved": round(dom_ach, 1), "peak": peak, "unit": "TFLOP/s", "frac": round(dom_ach / peak, 4),
                         "peak_source": peak_src + ": bf16_tflops_sustained of MEASURED_PEAKS.json (cuBLAS, seconds-long loop); the kernel was "
                                        "timed inside the step at the SM clock in `clocks`",
                         "frac_of_burst_peak": round(dom_ach / tensor_peak_burst(), 4),
                         "traffic": kept_traffic("conv_window_kernel"), "launches_per_step": dom_n,
                         "flops_per_launch": round(dom_flops / max(dom_n, 1)), "ms_per_launch": round(dom_ms / max(dom_n, 1), 5),
                         "ms_per_step_in_kernel": round(dom_ms, 3), "share_of_step": round(dom_ms / step_ms, 3),
                         "all_tcgen05": {"launches_per_step": um_n, "achieved": round(ach, 1), "frac": round(ach / peak, 4),
                                         "ms_per_step_in_kernels": round(um_ms, 3), "share_of_step": round(um_ms / step_ms, 3)},
                         "note": "algorithmic conv FLOPs (2*MACs) of the kernel's launches in one step / their summed "
                                 "CUDA-event durations, measured in one extra profiled step after the timed region "
                                 "(a CUDA graph replay with the event pairs as graph nodes and the weight-gradient kernels and the "
                                 "generator's optimiser update in the main stream -- the timed step runs them on side streams next to "
                                 "other kernels, where an event pair would time two kernels sharing the SMs; eager with --no-graph); share_of_step therefore "
                                 "compares un-overlapped kernel time with the overlapped step"},
            "step_tensor_utilisation": round(all_flops / (step_ms * 1e-3) / 1e12 / peak, 4),
            "conv_flops_per_step": round(all_flops),
            "kernels": {k: {"launches": v[0], "GFLOP": round(v[1] / 1e9, 1), "ms": round(v[2], 3),
                            "TFLOPs": round(v[1] / (v[2] * 1e-3) / 1e12, 1) if v[2] > 0 else None}
                        for k, v in sorted(classes.items())},
            "entry_points": {"note": "one profiled graph replay: device ms per step inside each C-ABI entry point (event pairs as graph "
                                     "nodes); `other` = step time not inside any of them (ATen kernels, NCCL, graph launch gaps)",
                             "ms": {k: round(v[1], 3) for k, v in sorted(calls.items(), key=lambda kv: -kv[1][1])},
                             "calls": {k: v[0] for k, v in sorted(calls.items(), key=lambda kv: -kv[1][1])},
                             "other_ms": round(step_ms - sum(v[1] for v in calls.values()), 3)},
            "e2e": {"value": round(B * world * K / (e2e_ms * 1e-3), 2), "unit": "samples/s",
                    "h2d_bytes_per_step": 2 * B * t_seg * 4 * world, "d2h_bytes_per_step": 4 * len(names) * world,
                    "ms_per_step": round(e2e_ms / K, 3)},
            "gpu_launches": int(launches), "clocks": clk,
        }
        del model
        torch.cuda.empty_cache()
        out["mdct4"] = _mdct_roofline(dev)
        if world == 1 and not args.no_cpu_baseline:
            transform = getattr(args, "transform", "mdct4")
            if reference_available(args.workload, transform):
                sec, n, warm, threads = cpu_reference_steps(B, 3, 1, args.workload, transform, budget_s=30.0)
                out["cpu_baseline"] = {"value": round(B / sec, 5), "unit": "samples/s", "cores": threads, "kind": "reference",
                                       "sample": f"{n} optimisation step(s) at batch {B} after {warm} warm-up: the reference's own "
                                                 f"Pix2PixHDModel (oracle/_ref, unmodified, fp32) on {threads} threads",
                                       "seconds_per_step": round(sec, 2)}
            else:
                sec, threads = cpu_train_step_seconds(batch=1, steps=1, workload=args.workload, transform=transform)
                out["cpu_baseline"] = {"value": round(1.0 / sec, 5), "unit": "samples/s", "cores": threads, "kind": "port",
                                       "sample": "one optimisation step at batch 1 (bounded sample of the batch-8 step), "
                                                 "oracle/pix2pixhd_oracle.py (torch CPU restatement of networks.py / "
                                                 f"pix2pixHD_model.py / train.py) on {threads} threads",
                                       "seconds_per_step": round(sec, 2)}
        print(json.dumps(out), flush=True)
    if world > 1:
        # CUDA graphs that captured NCCL kernels keep the communicator busy: drop every graph, drain the device, line the ranks
        # up, then tear the group down.  A watchdog ends the process if the teardown still does not return (seen in round 1).
        import gc
        import sys
        import threading
        step = pg = pg2 = None
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        threading.Timer(20.0, lambda: os._exit(0)).start()
        dist.destroy_process_group()
        os._exit(0)
