"""Golden vectors for the model rows: the REFERENCE's own Pix2PixHDModel + train.py step,
run on CPU in the build container.   python tests/golden/make_golden_model.py -> tests/golden/model_ref.npz

The shipped model wires MDCT2 (pix2pixHD_model.py:38-40); README.md:133 invites MDCT4.  Per SURVEY H1 the build's
model uses MDCT4, so the oracle here is the reference model with `_mdct` replaced by the reference's own MDCT4
(cast to float32, as MDCT2's output is) -- every other line of forward() / train.py:148-184 runs unmodified.
Config: G3L2-shaped LocalEnhancer in miniature (ngf 4, 3 global downsamplings, 2+2 blocks), 2-scale D with
feature matching, explicit encoding, mask_mode None, pool_size 0, Adam lr 2e-4 beta1 0.5, batch 2."""
import contextlib
import importlib
import io
import os
import sys
import tempfile

import numpy as np
import torch

from make_golden import import_reference

HERE = os.path.dirname(os.path.abspath(__file__))
ARGS = ['--gpu_ids', '-1', '--name', 't', '--netG', 'local', '--ngf', '4', '--n_downsample_global', '3',
        '--n_blocks_global', '2', '--n_blocks_local', '2', '--n_local_enhancers', '1', '--ndf', '8', '--num_D', '2',
        '--n_layers_D', '3', '--input_nc', '2', '--output_nc', '2', '--label_nc', '0', '--no_instance', '--no_vgg_loss',
        '--explicit_encoding', '--n_fft', '128', '--hop_length', '64', '--win_length', '128',
        '--lr_sampling_rate', '16000', '--hr_sampling_rate', '48000', '--batchSize', '2', '--center']
T, B, STEPS = 63 * 64, 2, 200


def batch(i):
    """deterministic synthetic (lr, hr) pair of step i -- reproduced by tests/test_model_gpu.py"""
    g = torch.Generator().manual_seed(1000 + i)
    hr = 0.1 * torch.randn(B, T, generator=g)
    lr = 0.5 * (hr + torch.roll(hr, 1, dims=1))
    return lr, hr


def main():
    MDCT4, IMDCT4, kbdwin = import_reference()
    tmp = tempfile.mkdtemp()
    sys.argv = ['train.py', '--checkpoints_dir', tmp] + ARGS
    TrainOptions = importlib.import_module("options.train_options").TrainOptions
    create_model = importlib.import_module("models.models").create_model
    with contextlib.redirect_stdout(io.StringIO()):
        opt = TrainOptions().parse()
        torch.manual_seed(opt.seed)
        model = create_model(opt)
    ref_mdct4 = MDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=model.window,
                      device='cpu')

    class F32(torch.nn.Module):
        def __init__(self, m):
            super().__init__()
            self.m = m

        def forward(self, x):
            return self.m(x.double() if x.dtype != torch.float32 or x.dim() == 3 else x).float()

    model._mdct = F32(ref_mdct4)
    out = {}
    for k, v in model.netG.state_dict().items():
        out["G0/" + k] = v.numpy().copy()
    for k, v in model.netD.state_dict().items():
        out["D0/" + k] = v.numpy().copy()
    curve = np.zeros((STEPS, 4), np.float64)
    for i in range(STEPS):
        lr, hr = batch(i)
        losses, generated = model(lr, None, hr, None, infer=(i == 0))
        losses = [torch.mean(x) if not isinstance(x, int) else x for x in losses]
        loss_dict = dict(zip(model.loss_names, losses))
        loss_D = (loss_dict['D_fake'] + loss_dict['D_real']) * 0.5
        loss_G = loss_dict['G_GAN'] + loss_dict.get('G_GAN_Feat', 0)
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
        model.optimizer_D.step()
        curve[i] = [float(loss_dict[n]) for n in ('G_GAN', 'G_GAN_Feat', 'D_real', 'D_fake')]
        if i == 0:
            out["loss_names"] = np.array(model.loss_names)
            out["sr0"] = generated.detach().numpy()
            with torch.no_grad():
                ls, pha, npar = model.to_spectro(lr, mask=True)
                hs, hpha, hpar = model.to_spectro(hr, mask=False)
            out["lr_spectro0"], out["lr_pha0"] = ls.numpy(), pha.numpy()
            out["hr_spectro0"] = hs.numpy()
            out["lr_norm0"] = np.array([float(npar[k]) for k in ('max', 'min', 'mean', 'std')])
            for k, v in model.netG.state_dict().items():
                out["G1/" + k] = v.numpy().copy()
            for k, v in model.netD.state_dict().items():
                out["D1/" + k] = v.numpy().copy()
        if i % 20 == 0:
            print(i, curve[i])
    out["curve"] = curve
    # inference() on the trained weights (pix2pixHD_model.py:437-455) and the audio it reconstructs through the
    # reference's IMDCT4 (to_audio with explicit decoding, :234-249)
    lr, hr = batch(STEPS)
    sr, pha, npar, ls = model.inference(lr, None)
    out["inf_sr"] = sr.numpy()
    ref_imdct4 = IMDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=model.window,
                        device='cpu')
    model._imdct = F32(ref_imdct4)
    out["inf_audio"] = model.to_audio(sr, npar, pha).numpy()
    for k, v in model.netG.state_dict().items():
        out["GN/" + k] = v.numpy().copy()
    np.savez_compressed(os.path.join(HERE, "model_ref.npz"), **out)
    print("params G", sum(p.numel() for p in model.netG.parameters()), "D", sum(p.numel() for p in model.netD.parameters()))
    print("sr", out["sr0"].shape, "audio", out["inf_audio"].shape)


if __name__ == "__main__":
    main()
