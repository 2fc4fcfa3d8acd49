"""Fixtures at the BENCHMARKED sizes, produced by the reference's own networks.py on CPU.

(a) `G3L2`: define_G(2, 2, 48, 'local', 4, 3, 1, 2, 'instance') and define_D(4, 64, 3, 'instance', False, 2, True)
    (generated/GEN_VCTK_G3L2_48ngf/opt.txt, models/networks.py:27-54) at [1,*,512,128]: float64 reference run ->
    outputs, D logits, input gradient, per-parameter gradient norms + a strided sample of every gradient.
(b) `GG`: BASELINE.json config 2, define_G(1, 1, 64, 'global', 4, 9) (options/base_options.py:48-51,
    models/networks.py:183-211) forward on [2,1,512,256] in float64.
Weights are NOT stored (156 M / 182 M parameters): the seed reproduces them bit-exactly
(tests/test_nets_gpu.py::test_state_dict_keys_and_seeded_init_match_reference checks that property).  The reference's
own fp32 run and its torch.autocast(bf16) run are compared to the float64 run and recorded as calib/*: what "fp32 /
bf16 parity" means for these seeded, perturbation-amplifying nets.

Run once in the build container:  python tests/golden/make_golden_nets_g3l2.py   (~10 min on 8 cores)."""
import contextlib
import importlib
import io
import os
import sys
import time

import numpy as np
import torch

from make_golden import import_reference

HERE = os.path.dirname(os.path.abspath(__file__))
SEED = 4848


def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def maxrel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def run_once(net, x, flat, dtype=None, autocast=None):
    for p in net.parameters():
        p.grad = None
    x = x.clone().requires_grad_(True)
    ctx = torch.autocast("cpu", dtype=autocast) if autocast is not None else contextlib.nullcontext()
    with ctx:
        outs = flat(net(x))
    g = torch.Generator().manual_seed(77)
    loss = 0
    for o in outs:
        loss = loss + (o.to(x.dtype) * torch.randn(o.shape, generator=g).to(x.dtype)).mean()
    loss.backward()
    return ([o.detach().double().numpy() for o in outs], x.grad.double().numpy(),
            {k: p.grad.double().reshape(-1).numpy().copy() for k, p in net.named_parameters()})


def fixture_train_nets(R, out, do_autocast):
    torch.manual_seed(SEED)
    with contextlib.redirect_stdout(io.StringIO()):
        G = R.define_G(2, 2, 48, 'local', 4, 3, 1, 2, 'instance', [])
        D = R.define_D(4, 64, 3, 'instance', False, 2, True, [])
    assert sum(p.numel() for p in G.parameters()) == 156050690 and sum(p.numel() for p in D.parameters()) == 5531522
    xg = torch.rand(1, 2, 512, 128)
    xd = torch.rand(1, 4, 512, 128)
    out["G/x"], out["D/x"] = xg.numpy(), xd.numpy()
    for tag, net, x, flat in (("G", G, xg, lambda o: [o]),
                              ("D", D, xd, lambda o: [t for s in o for t in s])):
        t0 = time.time()
        o32, gx32, g32 = run_once(net, x, flat)
        t1 = time.time()
        net.double()
        o64, gx64, g64 = run_once(net, x.double(), flat)
        net.float()
        t2 = time.time()
        print(tag, "fp32 run %.1f s, fp64 run %.1f s" % (t1 - t0, t2 - t1), flush=True)
        for i, o in enumerate(o64):
            if tag == "D" and i % 5 != 4:
                # intermediate feature maps are large: keep the norm and a strided sample
                out[f"{tag}/onorm{i}"] = np.array(float(np.linalg.norm(o)))
                out[f"{tag}/osample{i}"] = o.reshape(-1)[:: max(1, o.size // 4099)].astype(np.float32)
            else:
                out[f"{tag}/out{i}"] = o.astype(np.float32)
        out[f"{tag}/gx"] = gx64.astype(np.float32)
        calib = {"out": max(maxrel(a, b) for a, b in zip(o32, o64)), "gx": rel(gx32, gx64)}
        gmax = max(np.linalg.norm(v) for v in g64.values())
        worst = 0.0
        for k, v in g64.items():
            out[f"{tag}/gnorm/{k}"] = np.array(float(np.linalg.norm(v)))
            out[f"{tag}/gsample/{k}"] = v[:: max(1, v.size // 997)].astype(np.float32)
            if np.linalg.norm(v) > 1e-4 * gmax:
                worst = max(worst, rel(g32[k], v))
        calib["grad"] = worst
        if do_autocast:
            t3 = time.time()
            oa, gxa, ga = run_once(net, x, flat, autocast=torch.bfloat16)
            print(tag, "autocast bf16 run %.1f s" % (time.time() - t3), flush=True)
            calib["autocast_bf16_out"] = max(maxrel(a, b) for a, b in zip(oa, o64))
            calib["autocast_bf16_gx"] = rel(gxa, gx64)
            calib["autocast_bf16_grad"] = max(rel(ga[k], g64[k]) for k in g64 if np.linalg.norm(g64[k]) > 1e-4 * gmax)
        for k, v in calib.items():
            out[f"{tag}/calib/{k}"] = np.array(v)
        print(tag, [o.shape for o in o64], "reference vs its float64 run:", calib, flush=True)


def fixture_global_generator(R, out, do_autocast):
    torch.manual_seed(SEED + 1)
    with contextlib.redirect_stdout(io.StringIO()):
        GG = R.define_G(1, 1, 64, 'global', 4, 9, 1, 3, 'instance', [])
    assert sum(p.numel() for p in GG.parameters()) == 182430721
    x = torch.rand(2, 1, 512, 256)
    out["GG/x"] = x.numpy()
    with torch.no_grad():
        t0 = time.time()
        y32 = GG(x).double().numpy()
        t1 = time.time()
        GG.double()
        y64 = GG(x.double()).numpy()
        GG.float()
        print("GG fp32 forward %.1f s, fp64 %.1f s" % (t1 - t0, time.time() - t1), flush=True)
        out["GG/out0"] = y64.astype(np.float32)
        out["GG/calib/out"] = np.array(maxrel(y32, y64))
        if do_autocast:
            t2 = time.time()
            with torch.autocast("cpu", dtype=torch.bfloat16):
                ya = GG(x).double().numpy()
            print("GG autocast forward %.1f s" % (time.time() - t2), flush=True)
            out["GG/calib/autocast_bf16_out"] = np.array(maxrel(ya, y64))
    print("GG", y64.shape, "fp32 vs fp64", float(out["GG/calib/out"]), "autocast bf16 vs fp64",
          float(out.get("GG/calib/autocast_bf16_out", np.nan)), flush=True)


def main():
    import_reference()
    R = importlib.import_module("models.networks")
    do_autocast = "--no-autocast" not in sys.argv
    torch.set_num_threads(os.cpu_count())
    out = {}
    fixture_train_nets(R, out, do_autocast)
    fixture_global_generator(R, out, do_autocast)
    path = os.path.join(HERE, "nets_g3l2_ref.npz")
    np.savez_compressed(path, **out)
    print("saved", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
