"""The "global_g" workload of bench.py: BASELINE.json configs[1] -- one forward of the pix2pixHD GlobalGenerator
define_G(1, 1, 64, 'global', n_downsample_global=4, n_blocks_global=9) (options/base_options.py:48-51,
models/networks.py:183-211; 182 430 721 parameters, 1713.96 GFLOP per batch of 8) on a synthetic [8,1,512,256] MDCT
spectrogram, bf16 storage / fp32 accumulation.  `value` = samples/s with the batch resident in HBM, `e2e` = the same forward
fed from pinned host memory with the output image read back, `roofline` = the conv window kernel's launches of one forward
(CUDA events as graph nodes) against the measured bf16 peak.  fp32-vs-bf16 parity of this configuration against the
reference's float64 run: tests/test_nets_g3l2_gpu.py::test_config2_global_generator_forward."""
import json
import os
import time

METRIC = "G3L2 train samples/s @1/2/4/8 B200; MDCT4/IMDCT4 GB/s vs HBM peak"
GFLOP_PER_SAMPLE = 1713.96 / 8
WORKLOAD = "GlobalGenerator define_G(1,1,64,'global',4,9) forward on [%d,1,512,256] (BASELINE config 2)"


def _build(dev, precision):
    import contextlib
    import io
    import torch
    from pix2pixhdaudiosr_b200.models import networks as N
    N.set_default_precision(precision)
    try:
        torch.manual_seed(1234)
        with contextlib.redirect_stdout(io.StringIO()):
            G = N.define_G(1, 1, 64, 'global', 4, 9, 1, 3, 'instance', [])
    finally:
        N.set_default_precision("fp32")
    G.to(dev)
    N.sync_shadows(G)
    return G


def run_global_reference(args):
    """--impl reference: the reference's own GlobalGenerator (oracle/_ref networks.py, fp32, all host threads) at the same batch;
    falls back to the oracle port's generator when oracle/_ref is absent"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    from oracle import ref_runner
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    B = args.batch
    if ref_runner.available():
        import contextlib
        import importlib
        import io
        ref_runner._import_reference()
        R = importlib.import_module("models.networks")
        torch.manual_seed(1234)
        with contextlib.redirect_stdout(io.StringIO()):
            G = R.define_G(1, 1, 64, 'global', 4, 9, 1, 3, 'instance', [])
        kind, what = "reference", "the reference's own networks.define_G (oracle/_ref, unmodified, fp32)"
    else:
        from oracle import pix2pixhd_oracle as O
        G = O.build_global_generator(1, 1, 64, 4, 9) if hasattr(O, "build_global_generator") else None
        if G is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref absent and the port has no GlobalGenerator builder"}))
            return
        kind, what = "port", "oracle/pix2pixhd_oracle.py"
    x = torch.rand(B, 1, 512, 256)
    with torch.no_grad():
        t0 = time.perf_counter()
        G(x)
        first = time.perf_counter() - t0
        warm = 1
        if args.warmup > 1 and first * (args.warmup - 1 + args.steps) <= 200.0:
            for _ in range(args.warmup - 1):
                G(x)
            warm = args.warmup
        k = args.steps if first * args.steps <= 200.0 else max(2, int(200.0 / first))
        t = []
        for _ in range(k):
            t0 = time.perf_counter()
            G(x)
            t.append(time.perf_counter() - t0)
    sec = sum(t) / len(t)
    val = round(B / sec, 4)
    sample = f"{what} at batch {B}, {k} timed forward(s) of the requested {args.steps} after {warm} warm-up, {threads} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "samples/s", "n_gpus": args.gpus, "steps": k, "warmup": warm,
        "ms_per_step": round(sec * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": WORKLOAD % B, "global_batch": B, "sample": sample},
        "cpu_baseline": {"value": val, "unit": "samples/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


def run_global(args):
    import torch
    import torch.distributed as dist
    from bench import ClockSampler, kept_traffic, tensor_peak, tensor_peak_burst
    from pix2pixhdaudiosr_b200 import _lib, ops

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
    G = _build(dev, args.precision)
    g = torch.Generator().manual_seed(77 + rank)
    x_host = torch.rand(B, 1, 512, 256, generator=g).pin_memory()
    x_static = x_host.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                y = G(x_static)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        n0 = lib.p2p_kernel_launches()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            y_static = G(x_static)
        launches_per_step = lib.p2p_kernel_launches() - n0
        for _ in range(W):
            graph.replay()
        clocks = ClockSampler(local)
        barrier()
        clocks.start()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(K):
            graph.replay()
        t1.record()
        barrier()
        clk = clocks.stop()
        total_ms = t0.elapsed_time(t1)

        y_host = torch.empty(tuple(y_static.shape), dtype=y_static.dtype).pin_memory()

        def e2e_step():
            x_static.copy_(x_host, non_blocking=True)
            graph.replay()
            y_host.copy_(y_static, non_blocking=True)
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

        pg = torch.cuda.CUDAGraph()
        ops.start_conv_timing(external=True)
        with torch.cuda.graph(pg):
            G(x_static)
        ops.pause_conv_timing()
        pg.replay()
        rec = ops.stop_conv_timing()

    times = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = times.tolist()
    if rank == 0:
        peak, peak_src = tensor_peak()
        step_ms = total_ms / K
        dom = [r for r in rec if r[1] and r[0] == "fwd-window"]
        dom_flops, dom_ms, dom_n = sum(r[2] for r in dom), sum(r[3] for r in dom), len(dom)
        dom_ach = dom_flops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else 0.0
        all_flops = sum(r[2] for r in rec)
        out = {
            "metric": METRIC, "value": round(B * world * K / (total_ms * 1e-3), 2), "unit": "samples/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.precision, "data": "synthetic",
            "config": {"workload": WORKLOAD % B, "global_batch": B * world, "parallelism": f"dp{world} (independent replicas: inference has no exchange)",
                       "launch": "one CUDA graph replay per step",
                       "l2": "0.36 GB of bf16 weights per forward exceed the 126 MB L2; no flush", "weights": "random init N(0,0.02), seed 1234"},
            "roofline": {"bound": "tensor", "kernel": "conv_window_kernel (forward passes of one GlobalGenerator forward)",
                         "achieved": round(dom_ach, 1), "peak": peak, "unit": "TFLOP/s", "frac": round(dom_ach / peak, 4),
                         "peak_source": peak_src, "frac_of_burst_peak": round(dom_ach / tensor_peak_burst(), 4),
                         "traffic": kept_traffic("conv_window_kernel(global_g)"), "launches_per_step": dom_n,
                         "flops_per_launch": round(dom_flops / max(dom_n, 1)), "ms_per_launch": round(dom_ms / max(dom_n, 1), 5),
                         "ms_per_step_in_kernel": round(dom_ms, 3), "share_of_step": round(dom_ms / step_ms, 3)},
            "step_tensor_utilisation": round(all_flops / (step_ms * 1e-3) / 1e12 / peak, 4),
            "conv_flops_per_step": round(all_flops), "algorithmic_gflop_per_step": round(GFLOP_PER_SAMPLE * B, 2),
            "e2e": {"value": round(B * world * K / (e2e_ms * 1e-3), 2), "unit": "samples/s", "h2d_bytes_per_step": B * 512 * 256 * 4 * world,
                    "d2h_bytes_per_step": y_host.numel() * y_host.element_size() * world, "ms_per_step": round(e2e_ms / K, 3)},
            "gpu_launches": int(launches_per_step * K), "clocks": clk,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import ref_runner
            if ref_runner.available():
                import contextlib
                import importlib
                import io
                threads = os.cpu_count() or 1
                torch.set_num_threads(threads)
                ref_runner._import_reference()
                R = importlib.import_module("models.networks")
                torch.manual_seed(1234)
                with contextlib.redirect_stdout(io.StringIO()):
                    Gr = R.define_G(1, 1, 64, 'global', 4, 9, 1, 3, 'instance', [])
                xr = torch.rand(B, 1, 512, 256)
                with torch.no_grad():
                    Gr(xr)
                    ts = []
                    for _ in range(3):
                        t0_ = time.perf_counter()
                        Gr(xr)
                        ts.append(time.perf_counter() - t0_)
                        if sum(ts) > 25:
                            break
                sec = sum(ts) / len(ts)
                out["cpu_baseline"] = {"value": round(B / sec, 4), "unit": "samples/s", "cores": threads, "kind": "reference",
                                       "sample": f"{len(ts)} forward(s) at batch {B} after 1 warm-up: the reference's own define_G "
                                                 f"(oracle/_ref, fp32) on {threads} threads", "seconds_per_step": round(sec, 2)}
        print(json.dumps(out), flush=True)
    if world > 1:
        import gc
        import sys
        import threading
        graph = pg = None
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        threading.Timer(20.0, lambda: os._exit(0)).start()
        dist.destroy_process_group()
        os._exit(0)
