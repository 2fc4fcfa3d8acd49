"""Realistic-extent fixture for the bf16 tolerance: G3L2 / 2-scale D of generated/GEN_VCTK_G3L2_48ngf/opt.txt at the
true spatial size [1,*,512,128] but ngf=8 / ndf=16 so the CPU reference runs in seconds.  Weights are NOT stored:
the same seed reproduces them bit-exactly (tests/test_nets_gpu.py checks that property on the small fixture).
Stored: outputs, input gradient, per-parameter gradient norms and a strided gradient sample."""
import contextlib
import importlib
import io
import os

import numpy as np
import torch

from make_golden import import_reference

HERE = os.path.dirname(os.path.abspath(__file__))
SEED = 4321


def main():
    import_reference()
    R = importlib.import_module("models.networks")
    out = {}

    def run_once(net, x, flat):
        for p in net.parameters():
            p.grad = None
        x = x.clone().requires_grad_(True)
        outs = flat(net(x))
        g = torch.Generator().manual_seed(77)
        loss = 0
        for o in outs:
            loss = loss + (o * torch.randn(o.shape, generator=g).to(o.dtype)).mean()
        loss.backward()
        return [o.detach().double().numpy() for o in outs], x.grad.double().numpy(), \
            {k: p.grad.double().reshape(-1).numpy() for k, p in net.named_parameters()}

    def run(tag, net, x, flat):
        """the golden values come from the reference run in float64; the reference's own fp32 run is compared
        to it and its error recorded (`calib/...`) so tolerances for our fp32 path are calibrated, not guessed"""
        o32, gx32, g32 = run_once(net, x, flat)
        net.double()
        o64, gx64, g64 = run_once(net, x.double(), flat)
        net.float()
        for i, o in enumerate(o64):
            out[f"{tag}/out{i}"] = o.astype(np.float32)
        out[f"{tag}/gx"] = gx64.astype(np.float32)
        rel = lambda a, b: float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))
        calib = {"out": max(float(np.abs(a - b).max() / np.abs(b).max()) for a, b in zip(o32, o64)), "gx": rel(gx32, gx64)}
        gmax = max(np.linalg.norm(v) for v in g64.values())
        worst = 0.0
        for k, v in g64.items():
            out[f"{tag}/gnorm/{k}"] = np.array(float(np.linalg.norm(v)))
            out[f"{tag}/gsample/{k}"] = v[:: max(1, v.size // 97)].astype(np.float32)
            if np.linalg.norm(v) > 1e-4 * gmax:
                worst = max(worst, rel(g32[k], v))
        calib["grad"] = worst
        # the reference under its own reduced-precision mode (torch.autocast, CPU) against its float64 run: what
        # "bf16 / fp16 parity" can mean for this randomly initialised, perturbation-amplifying net
        for name, dt in (("bf16", torch.bfloat16), ("fp16", torch.float16)):
            xx = x.clone().requires_grad_(True)
            for p in net.parameters():
                p.grad = None
            with torch.autocast("cpu", dtype=dt):
                oo = flat(net(xx))
            g = torch.Generator().manual_seed(77)
            loss = 0
            for o in oo:
                loss = loss + (o.float() * torch.randn(o.shape, generator=g)).mean()
            loss.backward()
            calib[f"autocast_{name}_out"] = max(float(np.abs(a.detach().double().numpy() - b).max() / np.abs(b).max())
                                                for a, b in zip(oo, o64))
            calib[f"autocast_{name}_gx"] = rel(xx.grad.double().numpy(), gx64)
            calib[f"autocast_{name}_grad"] = max(rel(p.grad.double().reshape(-1).numpy(), g64[k])
                                                 for k, p in net.named_parameters() if np.linalg.norm(g64[k]) > 1e-4 * gmax)
        for k, v in calib.items():
            out[f"{tag}/calib/{k}"] = np.array(v)
        print(tag, [o.shape for o in o64], sum(p.numel() for p in net.parameters()), "reference fp32 vs fp64:", calib)

    torch.manual_seed(SEED)
    with contextlib.redirect_stdout(io.StringIO()):
        G = R.define_G(2, 2, 8, 'local', 4, 3, 1, 2, 'instance', [])
        D = R.define_D(4, 16, 3, 'instance', False, 2, True, [])
    xg = torch.rand(1, 2, 512, 128)
    xd = torch.rand(1, 4, 512, 128)
    out["G/x"], out["D/x"] = xg.numpy(), xd.numpy()
    run("G", G, xg, lambda o: [o])
    run("D", D, xd, lambda o: [t for s in o for t in s][4::5])     # the two logit maps only (features are big)
    np.savez_compressed(os.path.join(HERE, "nets_big_ref.npz"), **out)
    print("saved", os.path.getsize(os.path.join(HERE, "nets_big_ref.npz")))


if __name__ == "__main__":
    main()
