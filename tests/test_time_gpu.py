"""GPU parity of SURVEY 8a rows a16-a18 (config 4: time-domain discriminator and frame-match loss, the per-frame IMDCT4
inverse in the loss path with backward) against the reference run of tests/golden/model_time_ref.npz
(make_golden_model_time.py) and against the CPU oracle on seeded inputs.  All calls go through the C ABI."""
import math
import os

import numpy as np
import pytest
import torch

from cases import relerr
from oracle import mdct4_oracle as M4
from oracle import pix2pixhd_oracle as O
from pix2pixhdaudiosr_b200 import ops
from pix2pixhdaudiosr_b200.models.mdct import IMDCT4, MDCT4, _Plan, _SynthesisFn
from pix2pixhdaudiosr_b200.models.models import create_model
from pix2pixhdaudiosr_b200.train import assemble_losses, train_step
from pix2pixhdaudiosr_b200.util.util import kbdwin
from test_model_gpu import batch, tiny_opt

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = {'n_fft': 128, 'hop_length': 64, 'win_length': 128, 'alpha': 0.6, 'min_value': 1e-7, 'explicit_encoding': True,
       'up_ratio': 3.0, 'lambda_feat': 10.0, 'use_time_D': True, 'use_match_loss': True, 'lambda_time': 10.0, 'lambda_mat': 10.0,
       'G': {'netG': 'local', 'n_downsample_global': 3, 'n_blocks_global': 2, 'n_local_enhancers': 1, 'n_blocks_local': 2},
       'D': {'num_D': 2, 'n_layers_D': 3}}
NAMES = ['G_GAN', 'G_GAN_Feat', 'G_mat', 'G_GAN_t', 'D_real_t', 'D_fake_t', 'D_real', 'D_fake']


# 'mdct4': the reference model with its two transform hooks switched to MDCT4 (make_golden_model_time.py);
# 'mdct2': the reference model exactly as shipped, nothing replaced (make_golden_model_mdct2.py)
GOLDEN = {"mdct4": "model_time_ref.npz", "mdct2": "model_mdct2_ref.npz"}


@pytest.fixture(scope="module", params=["mdct4", "mdct2"])
def ref(request):
    r = dict(np.load(os.path.join(ROOT, "tests", "golden", GOLDEN[request.param])))
    r["transform"] = request.param
    return r


def time_opt(ref, **kw):
    return tiny_opt(use_time_D=True, use_match_loss=True, lambda_time=10.0, lambda_mat=10.0, transform=ref["transform"], **kw)


# ---- kernels ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n_fft,hop,T,B", [(128, 64, 63 * 64, 2), (1024, 512, 5000, 3), (64, 16, 1000, 2)])
def test_frame_window_matches_oracle(n_fft, hop, T, B):
    torch.manual_seed(3)
    x = 0.1 * torch.randn(B, T)
    cfg = dict(hop_length=hop, win_length=n_fft)
    want = O.frames_of(x, cfg)
    m = MDCT4(n_fft, hop, n_fft, kbdwin, True, device="cuda")
    start_pad, _ = m.pads(B)
    got = ops.frame_window(x.cuda(), m.window.float(), want.shape[1], hop, start_pad)
    assert tuple(got.shape) == tuple(want.shape)
    assert relerr(got.cpu().numpy(), want.numpy()) < 1e-6


@pytest.mark.parametrize("n_fft", [128, 1024, 2048])
def test_per_frame_inverse_and_its_adjoint(n_fft):
    """synthesis kernel at hop = win = n_fft (no overlap) = the per-frame inverse of models/mdct.py:547-549; its
    autograd adjoint is the analysis kernel at hop = n_fft"""
    torch.manual_seed(4)
    B, F, M = 2, 9, n_fft // 2
    X = torch.randn(B, F, M)
    g = torch.randn(B, F, n_fft)
    exp1 = torch.exp(-1j * (math.pi / (2 * n_fft) + math.pi / 4) * torch.arange(1, n_fft, 2, dtype=torch.float64))
    exp2 = torch.exp(-1j * math.pi / (2 * n_fft) * torch.arange(0, 2 * n_fft, 2, dtype=torch.float64))
    Xd = X.double().requires_grad_(True)
    want = 8.0 / n_fft * torch.real(torch.fft.fft(exp1 * Xd, n=n_fft) * exp2)
    for window in (torch.ones(n_fft), 1.7 * kbdwin(n_fft).float()):
        plan = _Plan(n_fft, n_fft, n_fft, window)
        Xc = X.cuda().requires_grad_(True)
        got = _SynthesisFn.apply(Xc, plan, F * n_fft, 0, 8.0 / n_fft).view(B, F, n_fft)
        (ggot,) = torch.autograd.grad(got, Xc, g.cuda())
        w = window.double()
        assert relerr(got.detach().cpu().numpy(), (w * want.detach()).numpy()) < 1e-5
        (gw,) = torch.autograd.grad(w * want, Xd, g.double(), retain_graph=True)
        assert relerr(ggot.cpu().numpy(), gw.numpy()) < 1e-5


def test_frames_db_forward_backward():
    torch.manual_seed(5)
    y = torch.randn(3, 7, 128) * torch.logspace(-9, 0, 128)      # spans the clamp at 1e-7
    w = kbdwin(128).float()
    gain, amin = math.sqrt(2.0), 1e-7
    g = torch.randn_like(y)
    for window in (None, w):
        yd = y.double().requires_grad_(True)
        v = (gain * (window.double() if window is not None else 1.0) * yd).abs()
        want = 20.0 * torch.log10(torch.clamp(v, min=amin)) - 20.0
        (gwant,) = torch.autograd.grad(want, yd, g.double())
        yc = y.cuda().requires_grad_(True)
        got = ops.FramesDbFn.apply(yc, window.cuda() if window is not None else None, gain, amin)
        (ggot,) = torch.autograd.grad(got, yc, g.cuda())
        assert np.abs(got.detach().cpu().numpy() - want.detach().numpy()).max() < 2e-4        # dB, fp32 log10
        clear = (v.detach() < 0.5 * amin) | (v.detach() > 2 * amin)                               # away from the clamp's kink
        err = (ggot.cpu().double() - gwant)[clear].abs().max() / gwant.abs().max()
        assert err < 1e-5


@pytest.mark.parametrize("B,F,N", [(2, 64, 128), (3, 5, 1024)])
def test_frame_match_loss_and_gradient(B, F, N):
    torch.manual_seed(6)
    y = torch.randn(B, F, N)
    cfg = dict(win_length=N, lambda_mat=10.0)
    yd = y.clone().requires_grad_(True)
    want = O.match_loss(yd, cfg)
    (gwant,) = torch.autograd.grad(want * 0.7, yd)
    yc = y.cuda().requires_grad_(True)
    got = ops.FrameMatchFn.apply(yc, kbdwin(N).float().cuda(), 10.0)
    (ggot,) = torch.autograd.grad(got * 0.7, yc)
    assert abs(float(got) - float(want)) / float(want) < 1e-5
    assert relerr(ggot.cpu().numpy(), gwant.numpy()) < 1e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_spectro_decode_gradient(dtype):
    torch.manual_seed(7)
    B, M, F = 2, 64, 33
    x = (torch.rand(B, M, F, 2) * 2 - 1).to(dtype)                 # channels-last memory behind an NCHW shape
    norm = {'max': torch.tensor(-3.0), 'min': torch.tensor(-120.0)}
    cfg = dict(alpha=0.6, min_value=1e-7)
    g = torch.randn(B, F, M)
    xd = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    want = O.decode_coefficients(xd, norm, cfg)
    (gwant,) = torch.autograd.grad(want, xd, g)
    xc = x.cuda().permute(0, 3, 1, 2).requires_grad_(True)
    nc = torch.stack((norm['max'], norm['min'])).cuda()
    got = ops.SpectroDecodeFn.apply(xc, nc, 0.6, 1e-7, 1.0)
    (ggot,) = torch.autograd.grad(got, xc, g.cuda())
    assert ggot.dtype == dtype and tuple(ggot.shape) == (B, 2, M, F)
    assert relerr(got.detach().cpu().numpy(), want.detach().numpy()) < 1e-5
    assert relerr(ggot.float().cpu().numpy(), gwant.numpy()) < (1e-5 if dtype == torch.float32 else 1e-2)


# ---- model -----------------------------------------------------------------------------------------------------
def _load(model, ref):
    for tag, net in (("G0/", model.netG), ("D0/", model.netD), ("T0/", model.time_D)):
        net.load_state_dict({k[len(tag):]: torch.from_numpy(ref[k]) for k in ref if k.startswith(tag)})
    for opt in (model.optimizer_G, model.optimizer_D):
        opt.flat_bf16.copy_(opt.flat)


def test_to_frames_matches_reference(ref):
    model = create_model(time_opt(ref))
    assert model.loss_names == NAMES == list(ref["loss_names"])
    lr, _ = batch(0)
    _, _, npar = model.to_spectro(lr, mask=True)
    assert relerr(npar['frames'].cpu().numpy(), ref["lr_frames0"]) < 1e-6
    fr = model.to_frames(torch.from_numpy(ref["sr0"]).cuda(), npar)
    assert tuple(fr.shape) == ref["sr_frames0"].shape
    assert relerr(fr.cpu().numpy(), ref["sr_frames0"]) < 5e-3       # batch-minimum conditioning of to_spectro (test_model_gpu)


def test_first_step_with_time_D_matches_reference(ref):
    """all eight losses, every gradient of G, D and time_D, and the weights after the first Adam steps (exact fp32 path)"""
    model = create_model(time_opt(ref))
    _load(model, ref)
    lr, hr = batch(0)
    losses, sr = model(lr, None, hr, None, infer=True)
    got = np.array([float(x.detach()) for x in losses])
    want = ref["curve"][0]
    print("losses", got, want)
    assert np.allclose(got, want, rtol=2e-2), (got, want)
    _, loss_G, loss_D = assemble_losses(model, losses)
    model.optimizer_G.zero_grad()
    loss_G.backward()
    tol = 5e-2
    grads = {("G", k): p.grad.detach().cpu().numpy().copy() for k, p in model.netG.named_parameters()}
    model.optimizer_G.step()
    model.optimizer_D.zero_grad()
    loss_D.backward()
    grads.update({("D", k): p.grad.detach().cpu().numpy().copy() for k, p in model.netD.named_parameters()})
    grads.update({("T", k): p.grad.detach().cpu().numpy().copy() for k, p in model.time_D.named_parameters()})
    model.optimizer_D.step()
    for tag in "GDT":
        keys = [k for (t, k) in grads if t == tag]
        gmax = max(np.abs(ref[f"g{tag}0/{k}"]).max() for k in keys)
        for k in keys:
            r = ref[f"g{tag}0/{k}"]
            if np.abs(r).max() < 1e-4 * gmax:
                continue                    # bias in front of an InstanceNorm: true gradient 0
            err = np.linalg.norm(grads[(tag, k)] - r) / np.linalg.norm(r)
            assert err < tol, (tag, k, err)
    for tag, net in (("G", model.netG), ("D", model.netD), ("T", model.time_D)):
        for k, v in net.state_dict().items():
            assert np.abs(v.cpu().numpy() - ref[f"{tag}1/{k}"]).max() <= 2.0 * 2e-4 + 1e-7, (tag, k)


def test_first_step_is_bit_reproducible_in_deterministic_mode(ref):
    """P2P_DETERMINISTIC (p2p_set_deterministic): no split-K / split-pixel red.add between CTAs, single-chunk statistics and
    loss sums -- two runs of the config-4 first step from the same state must give BIT-IDENTICAL losses and gradients (fp32 and
    bf16 paths), and the fp32 losses must sit within 2e-3 of the reference's (2e-2 is what atomics reordering needed)"""
    from pix2pixhdaudiosr_b200 import _lib
    old = _lib.lib().p2p_set_deterministic(1)
    try:
        for precision in ("fp32", "bf16"):
            runs = []
            for _ in range(2):
                model = create_model(time_opt(ref, precision=precision))
                _load(model, ref)
                lr, hr = batch(0)
                losses, _ = model(lr, None, hr, None, infer=False)
                vals = torch.stack([x.detach().float().reshape(()) for x in losses]).cpu()
                _, loss_G, loss_D = assemble_losses(model, losses)
                model.optimizer_G.zero_grad()
                loss_G.backward()
                model.optimizer_D.zero_grad()
                loss_D.backward()
                runs.append((vals, model.optimizer_G.flat_grad.clone(), model.optimizer_D.flat_grad.clone()))
                del model
            assert torch.equal(runs[0][0], runs[1][0]), (precision, runs[0][0], runs[1][0])
            assert torch.equal(runs[0][1], runs[1][1]) and torch.equal(runs[0][2], runs[1][2]), precision
            if precision == "fp32":
                got, want = runs[0][0].numpy().astype(np.float64), ref["curve"][0]
                mat = NAMES.index('G_mat')
                err = np.abs(got - want) / np.abs(want)
                print("deterministic fp32 first-step loss errors", err)
                assert np.delete(err, mat).max() < 2e-3, err
    finally:
        _lib.lib().p2p_set_deterministic(old)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_time_D_loss_curve(ref, precision):
    """40 optimisation steps with --use_time_D --use_match_loss from the reference's seeded weights"""
    model = create_model(time_opt(ref, precision=precision))
    _load(model, ref)
    curve = []
    for i in range(40):
        loss_dict, _ = train_step(model, *batch(i))
        curve.append(torch.stack([loss_dict[n].detach().float() for n in NAMES]))
    curve = torch.stack(curve).cpu().numpy().astype(np.float64)
    want = ref["curve"]
    assert np.isfinite(curve).all()
    mat = NAMES.index('G_mat')
    others = [j for j in range(len(NAMES)) if j != mat]
    scale = np.abs(want).max(axis=0)
    early = (np.abs(curve[:8] - want[:8]) / scale)[:, others].max()
    k = np.ones(10) / 10
    sm = lambda c: np.stack([np.convolve(c[:, j], k, mode="valid") for j in range(len(NAMES))], 1)
    band = (np.abs(sm(curve) - sm(want)) / np.abs(sm(want)))[:, others]
    print(precision, "first-8-step error", early, "moving-average deviation max", band.max(), "mean", band.mean(),
          "G_mat last", curve[-1, mat], want[-1, mat])
    assert early < (0.05 if precision == "fp32" else 0.15)
    assert band.max() < (0.25 if precision == "fp32" else 0.5)
    # G_mat starts at 6e-6 (a difference of nearly equal numbers, four orders of magnitude below the other losses) and
    # grows 10^4-fold over the run: early on it follows the reference within a factor (bf16: 5x seen), and it lands on
    # the reference's level once it is large enough to be driven by the optimisation rather than by rounding
    assert np.abs(np.log(curve[:8, mat] / want[:8, mat])).max() < np.log(10.0)      # 2.6x seen in fp32 (atomics reorder sums), 5x in bf16
    assert abs(np.log(curve[-10:, mat].mean() / want[-10:, mat].mean())) < np.log(2.0)       # 1.00-1.07 seen


def test_inference_and_to_audio_match_shipped_reference(ref):
    if ref["transform"] != "mdct2":
        pytest.skip("the MDCT4 variant's inference is covered by test_model_gpu.py")
    model = create_model(time_opt(ref))
    model.netG.load_state_dict({k[3:]: torch.from_numpy(ref[k]) for k in ref if k.startswith("GN/")})
    lr, _ = batch(40)
    sr, pha, npar, ls = model.inference(lr, None)
    assert relerr(sr.float().cpu().numpy(), ref["inf_sr"]) < 5e-3
    audio = model.to_audio(torch.from_numpy(ref["inf_sr"]).cuda(), npar, pha)
    assert tuple(audio.shape) == ref["inf_audio"].shape
    assert relerr(audio.cpu().numpy(), ref["inf_audio"]) < 2e-3


def test_time_D_step_in_a_cuda_graph(ref):
    """the config-4 step is capturable (no host sync inside) and replays to finite losses"""
    from pix2pixhdaudiosr_b200.train import GraphedTrainStep
    model = create_model(time_opt(ref, precision="bf16"))
    lr, hr = batch(0)
    step = GraphedTrainStep(model, lr.cuda(), hr.cuda())
    for i in range(3):
        lr, hr = batch(i)
        losses, _ = step(lr.cuda(), hr.cuda())
    torch.cuda.synchronize()
    assert set(losses) == set(NAMES)
    assert all(math.isfinite(float(v)) for v in losses.values())
