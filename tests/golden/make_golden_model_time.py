"""Golden vectors for SURVEY 8a rows a16-a18 (config 4): the REFERENCE's own Pix2PixHDModel + train.py step with
--use_time_D --use_match_loss, run on CPU in the build container.
    python tests/golden/make_golden_model_time.py -> tests/golden/model_time_ref.npz

As in make_golden_model.py the transform is the reference's MDCT4 (the shipped model wires MDCT2).  The two MDCT2-side
hooks the time-domain path reads are given their MDCT4 counterparts, both made of the reference's own lines:
  * `_mdct(audio, True)` also returns the windowed frames (the pad / unfold / window lines models/mdct.py:488-503 that
    MDCT2.forward clones for return_ola, :377-394);
  * `_idct` = IMDCT4.forward's per-frame inverse (models/mdct.py:547-549 with IMDCT4's own exp1 / exp2), times
    8/n_fft: IDCT_2N is twice the frame IMDCT2 windows and overlap-adds (:440) and IMDCT4 overlap-adds 4/n_fft of it.
Every other line of forward() (discriminate_time_D, the :386 generator pass on raw frames, criterionMatch) and of
train.py:155-184 runs unmodified.  Config: make_golden_model.py's miniature G3L2 + lambda_time 10 (train_script.sh:93)."""
import contextlib
import importlib
import io
import os
import sys
import tempfile

import numpy as np
import torch
from torch.nn.functional import pad

from make_golden import import_reference
from make_golden_model import ARGS, B, T, batch

HERE = os.path.dirname(os.path.abspath(__file__))
STEPS = 40


def main(shipped_mdct2=False, out_name="model_time_ref.npz"):
    """shipped_mdct2=True (make_golden_model_mdct2.py): the reference model exactly as shipped -- MDCT2 / IMDCT2 /
    IDCT_2N_native, nothing replaced"""
    MDCT4, IMDCT4, kbdwin = import_reference()
    tmp = tempfile.mkdtemp()
    sys.argv = ['train.py', '--checkpoints_dir', tmp] + ARGS + ['--use_time_D', '--use_match_loss', '--lambda_time', '10']
    TrainOptions = importlib.import_module("options.train_options").TrainOptions
    create_model = importlib.import_module("models.models").create_model
    with contextlib.redirect_stdout(io.StringIO()):
        opt = TrainOptions().parse()
        torch.manual_seed(opt.seed)
        model = create_model(opt)
    ref_mdct4 = MDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=model.window, device='cpu')
    ref_imdct4 = IMDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=model.window, device='cpu')

    class Mdct4WithFrames(torch.nn.Module):
        def forward(self, signal, return_ola=False):
            spec = ref_mdct4(signal).float()
            if not return_ola:
                return spec
            m = ref_mdct4                                     # models/mdct.py:488-503, verbatim semantics
            start_pad = m.hop_length if m.center else 0
            additional_len = int(len(signal)) % m.hop_length
            end_pad = start_pad + (m.hop_length - additional_len if additional_len else 0)
            fr = pad(signal, (start_pad, end_pad), mode=m.pad_mode).unfold(dimension=-1, size=m.win_length, step=m.hop_length)
            return spec, torch.mul(fr, m.window)

    class FrameInverse(torch.nn.Module):
        def forward(self, x):
            im = ref_imdct4                                   # models/mdct.py:547-549
            y = torch.real(torch.fft.fft(im.exp1 * x.double(), n=im.n_fft) * im.exp2)
            return (8.0 / im.n_fft * y).float()

    if not shipped_mdct2:
        model._mdct = Mdct4WithFrames()
        model._idct = FrameInverse()
    out = {}
    for tag, net in (("G0/", model.netG), ("D0/", model.netD), ("T0/", model.time_D)):
        for k, v in net.state_dict().items():
            out[tag + k] = v.numpy().copy()
    names = None
    curve = []
    for i in range(STEPS):
        lr, hr = batch(i)
        losses, generated = model(lr, None, hr, None, infer=(i == 0))
        losses = [torch.mean(x) if not isinstance(x, int) else x for x in losses]
        loss_dict = dict(zip(model.loss_names, losses))
        # train.py:159-160
        loss_D = (loss_dict['D_fake'] + loss_dict['D_real']) * 0.5 + (loss_dict['D_fake_t'] + loss_dict['D_real_t']) * 0.5
        loss_G = loss_dict['G_GAN'] + loss_dict.get('G_mat', 0) + loss_dict.get('G_GAN_Feat', 0) + loss_dict.get('G_VGG', 0) + loss_dict['G_GAN_t']
        model.optimizer_G.zero_grad()
        loss_G.backward()
        if i == 0:
            for k, p in model.netG.named_parameters():
                out["gG0/" + k] = p.grad.numpy().copy()
        model.optimizer_G.step()
        model.optimizer_D.zero_grad()
        loss_D.backward()
        if i == 0:
            for k, p in model.netD.named_parameters():
                out["gD0/" + k] = p.grad.numpy().copy()
            for k, p in model.time_D.named_parameters():
                out["gT0/" + k] = p.grad.numpy().copy()
        model.optimizer_D.step()
        names = list(model.loss_names)
        curve.append([float(loss_dict[n]) for n in names])
        if i == 0:
            out["sr0"] = generated.detach().numpy()
            with torch.no_grad():
                ls, pha, npar = model.to_spectro(lr, mask=True)
                hs, _, _ = model.to_spectro(hr, mask=False)
                out["lr_spectro0"], out["hr_spectro0"], out["lr_pha0"] = ls.numpy(), hs.numpy(), pha.numpy()
                out["lr_norm0"] = np.array([float(npar[k]) for k in ('max', 'min', 'mean', 'std')])
                out["lr_frames0"] = npar['frames'].numpy()
                out["sr_frames0"] = model.to_frames(generated.detach(), npar).numpy()
            for tag, net in (("G1/", model.netG), ("D1/", model.netD), ("T1/", model.time_D)):
                for k, v in net.state_dict().items():
                    out[tag + k] = v.numpy().copy()
        if i % 10 == 0:
            print(i, dict(zip(names, curve[-1])))
    out["loss_names"] = np.array(names)
    out["curve"] = np.array(curve, np.float64)
    lr, hr = batch(STEPS)
    sr, pha, npar, ls = model.inference(lr, None)
    out["inf_sr"] = sr.numpy()
    if shipped_mdct2:
        out["inf_audio"] = model.to_audio(sr, npar, pha).numpy()
    for k, v in model.netG.state_dict().items():
        out["GN/" + k] = v.numpy().copy()
    np.savez_compressed(os.path.join(HERE, out_name), **out)
    print("params T", sum(p.numel() for p in model.time_D.parameters()), "frames", out["sr_frames0"].shape)


if __name__ == "__main__":
    main()
