"""Parity of the BENCHMARKED arithmetic: the real G3L2 ngf-48 LocalEnhancer + ndf-64 2-scale D
(generated/GEN_VCTK_G3L2_48ngf/opt.txt) and BASELINE config 2 (define_G(1,1,64,'global',4,9)) against fixtures made by the
reference's own networks.py in float64 (tests/golden/make_golden_nets_g3l2.py), plus the tcgen05 kernels directly against
torch.nn.functional.conv2d / conv_transpose2d (fp32, TF32 off) at the dominant layer shapes.

Tolerances (written next to each assert):
  fp32 path     outputs 1e-4 (max-abs / max-abs), gradients max(1e-3, 3 x the reference's own fp32-vs-fp64 error)
  bf16 path     per layer, fp32 activations injected: 2e-2 relative (BASELINE.json north_star)
                end to end: max(2e-2, the reference's own torch.autocast(bf16) error vs its float64 run) -- the measured
                errors are written to gpurun_out/r02_parity_ngf48.json (kept under profiles/)."""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from cases import relerr
from pix2pixhdaudiosr_b200 import ops
from pix2pixhdaudiosr_b200.models import networks as N

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 4848
_REPORT = {}


@pytest.fixture(scope="module")
def fix():
    return np.load(os.path.join(ROOT, "tests", "golden", "nets_g3l2_ref.npz"))


@pytest.fixture(scope="module", autouse=True)
def _write_report():
    yield
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "r02_parity_ngf48.json"), "w") as f:
        json.dump(_REPORT, f, indent=1, sort_keys=True)


def _normerr(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def _build(precision):
    N.set_default_precision(precision)
    try:
        torch.manual_seed(SEED)
        G = N.define_G(2, 2, 48, 'local', 4, 3, 1, 2, 'instance', [])
        D = N.define_D(4, 64, 3, 'instance', False, 2, True, [])
    finally:
        N.set_default_precision("fp32")
    assert sum(p.numel() for p in G.parameters()) == 156050690          # train_script.sh:59
    assert sum(p.numel() for p in D.parameters()) == 5531522
    for net in (G, D):
        net.cuda()
        N.sync_shadows(net)
    return G, D


def _flat_D(o):
    return [t for s in o for t in s]


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_g3l2_ngf48_against_reference(fix, precision):
    G, D = _build(precision)
    for tag, net, flat in (("G", G, lambda o: [o]), ("D", D, _flat_D)):
        calib = {k: float(fix[f"{tag}/calib/{k}"]) for k in ("out", "gx", "grad", "autocast_bf16_out", "autocast_bf16_gx",
                                                             "autocast_bf16_grad")}
        if precision == "fp32":
            tol_o, tol_g = 1e-4, max(1e-3, 3 * calib["grad"], 3 * calib["gx"])
        else:
            tol_o = max(2e-2, calib["autocast_bf16_out"])
            tol_g = max(2e-2, calib["autocast_bf16_grad"], calib["autocast_bf16_gx"])
        x = torch.from_numpy(fix[tag + "/x"]).cuda().requires_grad_(True)
        outs = flat(net(x))
        g = torch.Generator().manual_seed(77)
        loss, worst_o = 0, 0.0
        for i, o in enumerate(outs):
            r = torch.randn(o.shape, generator=g).cuda()
            oc = o.detach().float().cpu().numpy()
            if f"{tag}/out{i}" in fix.files:
                eo = relerr(oc, fix[f"{tag}/out{i}"])
            else:           # D's feature maps: strided sample + norm
                samp = oc.reshape(-1)[:: max(1, oc.size // 4099)]
                ref = fix[f"{tag}/osample{i}"]
                eo = max(float(np.abs(samp - ref).max() / np.abs(ref).max()),
                         abs(float(np.linalg.norm(oc)) - float(fix[f"{tag}/onorm{i}"])) / float(fix[f"{tag}/onorm{i}"]))
            worst_o = max(worst_o, eo)
            assert eo < tol_o, (tag, precision, i, eo, tol_o)
            loss = loss + (o.float() * r).mean()
        loss.backward()
        egx = _normerr(x.grad.cpu().numpy(), fix[tag + "/gx"])
        assert egx < tol_g, (tag, precision, egx, tol_g)
        gmax = max(float(fix[f"{tag}/gnorm/{k}"]) for k, _ in net.named_parameters())
        worst = 0.0
        for k, p in net.named_parameters():
            ref_norm = float(fix[f"{tag}/gnorm/{k}"])
            if ref_norm < 1e-4 * gmax:
                continue                                               # zero-gradient bias in front of InstanceNorm
            gflat = p.grad.reshape(-1)
            samp = gflat[:: max(1, gflat.numel() // 997)].cpu().numpy()
            ref = fix[f"{tag}/gsample/{k}"]
            err_n = abs(float(gflat.norm()) - ref_norm) / ref_norm
            err_s = float(np.linalg.norm(samp - ref) / max(np.linalg.norm(ref), 1e-30))
            worst = max(worst, err_n, err_s)
            assert err_n < tol_g and err_s < 2.5 * tol_g, (tag, precision, k, err_n, err_s, tol_g)
        _REPORT[f"{tag}/{precision}"] = dict(output_err=worst_o, gx_err=egx, worst_param_grad_err=worst, tol_out=tol_o, tol_grad=tol_g,
                                             reference_fp32_vs_fp64=dict(out=calib["out"], gx=calib["gx"], grad=calib["grad"]),
                                             reference_autocast_bf16_vs_fp64=dict(out=calib["autocast_bf16_out"], gx=calib["autocast_bf16_gx"],
                                                                                  grad=calib["autocast_bf16_grad"]))
        print(tag, precision, _REPORT[f"{tag}/{precision}"])
        del outs, loss
    del G, D
    torch.cuda.empty_cache()


def test_bf16_layers_at_ngf48_with_fp32_activations_injected(fix):
    """every conv block of the real ngf-48 G and ndf-64 D: the fp32 (exact CUDA-core) path runs the net and records each
    block's input; the bf16 tensor-core path then runs the SAME block on the SAME input (rounded to bf16) -- 2e-2 relative
    (north_star).  Together with the fp32 path's 1e-4 agreement with the reference fixture this bounds the benchmarked
    arithmetic layer by layer at its real shapes."""
    G, D = _build("fp32")
    records = []
    orig = N._conv_block

    def recording(conv, x, cfg, needs_wgrad, norm, act_mod, res, *rest):
        z = orig(conv, x, cfg, needs_wgrad, norm, act_mod, res, *rest)
        records.append((conv, x.detach(), cfg, norm, act_mod, None if res is None else res.detach(), z.detach()))
        return z

    N._conv_block = recording
    try:
        with torch.no_grad():
            G(torch.from_numpy(fix["G/x"]).cuda())
            D(torch.from_numpy(fix["D/x"]).cuda())
    finally:
        N._conv_block = orig
    assert len(records) == 23 + 10, len(records)        # G: 1 + 4 + 6 + 4 global, 2 + 4 + 1 + 1 local; D: 5 layers x 2 scales
    worst = {}
    for i, (conv, x, cfg, norm, act_mod, res, z32) in enumerate(records):
        conv.sync_shadow()
        xb = x.to(torch.bfloat16)
        rb = None if res is None else res.to(torch.bfloat16)
        with torch.no_grad():
            zb = orig(conv, xb, cfg, False, norm, act_mod, rb).float()
        ref = z32 - (res if res is not None else 0)
        got = zb - (rb.float() if rb is not None else 0)
        err = float((got - ref).norm() / ref.norm())
        name = f"{i:02d}_{type(conv).__name__}_{conv.in_channels}to{conv.out_channels}_k{conv.kernel_size}s{conv.stride}_{tuple(x.shape[1:3])}"
        worst[name] = err
        assert err < 2e-2, (name, err)
    _REPORT["per_layer_bf16_vs_fp32_injected"] = worst
    print(worst)


def test_config2_global_generator_forward(fix):
    """BASELINE.json config 2: define_G(1, 1, 64, 'global', 4, 9) forward on [2,1,512,256], fp32 and bf16, against the
    reference's float64 output.  fp32: 1e-4; bf16: max(2e-2, the reference's own autocast(bf16) error), measured value kept."""
    ref = fix["GG/out0"]
    for precision in ("fp32", "bf16"):
        N.set_default_precision(precision)
        try:
            torch.manual_seed(SEED + 1)
            GG = N.define_G(1, 1, 64, 'global', 4, 9, 1, 3, 'instance', [])
        finally:
            N.set_default_precision("fp32")
        assert sum(p.numel() for p in GG.parameters()) == 182430721
        GG.cuda()
        N.sync_shadows(GG)
        with torch.no_grad():
            y = GG(torch.from_numpy(fix["GG/x"]).cuda()).float().cpu().numpy()
        err = relerr(y, ref)
        tol = 1e-4 if precision == "fp32" else max(2e-2, float(fix["GG/calib/autocast_bf16_out"]))
        _REPORT[f"GG/{precision}"] = dict(output_err=err, tol=tol, reference_fp32_vs_fp64=float(fix["GG/calib/out"]),
                                          reference_autocast_bf16_vs_fp64=float(fix["GG/calib/autocast_bf16_out"]))
        print("config 2", precision, _REPORT[f"GG/{precision}"])
        assert err < tol, (precision, err, tol)
        del GG
        torch.cuda.empty_cache()


# (name, transposed, Cin, Cout, k, stride, pad, reflect, (N, H, W)): the layers that carry the step's FLOPs
DOMINANT = [
    ("g_res96_3x3_256x64", False, 96, 96, 3, 1, 1, True, (8, 256, 64)),
    ("g_bottleneck1536_3x3_16x4", False, 1536, 1536, 3, 1, 1, True, (8, 16, 4)),
    ("g_down_192to384_s2", False, 192, 384, 3, 2, 1, False, (8, 128, 32)),
    ("g_up_convT_384to192", True, 384, 192, 3, 2, 1, False, (8, 64, 16)),
    ("d_256to512_4x4s1_65x17", False, 256, 512, 4, 1, 2, False, (8, 65, 17)),
    ("d_128to256_4x4s2_129x33", False, 128, 256, 4, 2, 2, False, (8, 129, 33)),
]


@pytest.mark.parametrize("layer", DOMINANT, ids=[l[0] for l in DOMINANT])
def test_tcgen05_layers_against_torch_conv(layer):
    """tcgen05 forward / input-gradient / weight-gradient kernels against torch's fp32 convolution (TF32 off) on the same
    bf16-representable operands, at the real batch-8 shapes.  fp32 results: 2e-4 of the norm (accumulation order);
    bf16-rounded results: 4e-3 (one bf16 rounding)."""
    from pix2pixhdaudiosr_b200 import _lib
    name, transposed, Cin, Cout, k, stride, pad, reflect, (N_, H, W) = layer
    old_tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        torch.manual_seed(len(name))
        x = torch.randn(N_, H, W, Cin, device="cuda").bfloat16()
        shape = (Cin, Cout, k, k) if transposed else (Cout, Cin, k, k)
        perm = (2, 3, 1, 0) if transposed else (2, 3, 0, 1)
        inv = [perm.index(i) for i in range(4)]
        wl = (torch.randn(shape, device="cuda") * 0.05).bfloat16().float()
        b = torch.randn(Cout, device="cuda") * 0.1
        cfg = ops.ConvCfg(transposed, stride, pad, ops.PAD_REFLECT if reflect else ops.PAD_ZERO, 1 if transposed else 0)
        w = wl.permute(*perm).contiguous().permute(*inv)
        w16 = w.permute(*perm).contiguous().bfloat16()
        # torch reference (NCHW fp32)
        xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
        wr = wl.clone().requires_grad_(True)
        if transposed:
            yr = F.conv_transpose2d(xr, wr, b, stride=stride, padding=pad, output_padding=1)
        else:
            xp = F.pad(xr, (pad, pad, pad, pad), mode="reflect") if reflect else xr
            yr = F.conv2d(xp, wr, b, stride=stride, padding=0 if reflect else pad)
        g = torch.randn(yr.shape, device="cuda", generator=torch.Generator("cuda").manual_seed(3)).bfloat16()
        yr.backward(g.float())
        n0 = _lib.lib().p2p_conv_umma_launches()
        for out_dtype, tol_o in ((torch.float32, 2e-4), (torch.bfloat16, 4e-3)):
            y = ops.conv_forward(x, w, b, w16, cfg, out_dtype)
            err = float((y.float().permute(0, 3, 1, 2) - yr.detach()).norm() / yr.detach().norm())
            assert err < tol_o, (name, "out", out_dtype, err)
        dx, dw, db = ops.conv_backward(x, w, w16, g.permute(0, 2, 3, 1).contiguous(), cfg, True, True, True)
        torch.cuda.synchronize()
        assert _lib.lib().p2p_conv_umma_launches() - n0 >= 4, "the tensor-core kernels must be the ones that ran"
        e_dx = float((dx.float().permute(0, 3, 1, 2) - xr.grad).norm() / xr.grad.norm())
        e_dw = float((dw.float() - wr.grad).norm() / wr.grad.norm())
        e_db = float((db.float() - g.float().sum((0, 2, 3))).norm() / g.float().sum((0, 2, 3)).norm())
        assert e_dx < 4e-3 and e_dw < 2e-4 and e_db < 1e-4, (name, e_dx, e_dw, e_db)
        _REPORT[f"tcgen05_vs_torch/{name}"] = dict(dx=e_dx, dw=e_dw, db=e_db)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old_tf32
