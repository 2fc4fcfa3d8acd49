"""CPU ORACLE (test infrastructure, never the product path) for the conv-GAN rows of the hot path:
a functional restatement, on plain torch CPU ops, of

    models/networks.py:129-253   LocalEnhancer / GlobalGenerator / ResnetBlock      -> generator()
    models/networks.py:292-383   MultiscaleDiscriminator / NLayerDiscriminator      -> discriminator()
    models/networks.py:68-110    GANLoss (LSGAN branch)                             -> gan_loss()
    models/pix2pixHD_model.py:142-227, 229-249   to_spectro / denormalize / to_audio -> to_spectro(), to_audio()
    models/pix2pixHD_model.py:331-435            forward (GAN + feature matching)   -> model_losses()
    models/pix2pixHD_model.py:251-258, 314-320, 375-387, 408-415   to_frames, time-domain D, frame match
                                                                   -> to_frames(), frames_of(), time_losses(), match_loss()
    train.py:155-184             loss assembly, two backward passes, two Adam steps -> TrainState.step()

working directly on reference-format state_dicts (same keys, same shapes).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this file.

Parity pinned: tests/test_oracle_nets.py checks it against the reference's own outputs committed under
tests/golden/ (nets_ref.npz: network outputs and gradients; model_ref.npz: encoded images, first-step losses and
gradients, weights after the first Adam step, 200-step loss curve, inference + reconstructed audio; model_time_ref.npz:
the same step with --use_time_D --use_match_loss), all produced by importing /root/reference in the build container
(tests/golden/make_golden_nets.py, make_golden_model.py, make_golden_model_time.py).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import mdct2_oracle as M2
from . import mdct4_oracle as M4


# ---- layers ------------------------------------------------------------------------------------------
def _conv(x, sd, key, stride=1, pad=0, reflect=0):
    if reflect:
        x = F.pad(x, (reflect, reflect, reflect, reflect), mode='reflect')      # nn.ReflectionPad2d
    return F.conv2d(x, sd[key + '.weight'], sd[key + '.bias'], stride=stride, padding=pad)


def _convT(x, sd, key):
    # nn.ConvTranspose2d(k=3, stride=2, padding=1, output_padding=1)   networks.py:151, 203
    return F.conv_transpose2d(x, sd[key + '.weight'], sd[key + '.bias'], stride=2, padding=1, output_padding=1)


def _in(x):
    return F.instance_norm(x, eps=1e-5)         # InstanceNorm2d(affine=False), networks.py:22


def _resblock(x, sd, key):
    # networks.py:214-253
    h = F.relu(_in(_conv(x, sd, key + '.conv_block.1', reflect=1)))
    return x + _in(_conv(h, sd, key + '.conv_block.5', reflect=1))


def _pool(x):
    return F.avg_pool2d(x, 3, stride=2, padding=[1, 1], count_include_pad=False)   # networks.py:166, 309


def _global_trunk(x, sd, prefix, n_down, n_blocks, head):
    """GlobalGenerator.model (networks.py:183-211); head=False drops the last three modules (networks.py:138)"""
    i = 1
    x = F.relu(_in(_conv(x, sd, f'{prefix}.{i}', reflect=3)))
    i = 4
    for _ in range(n_down):
        x = F.relu(_in(_conv(x, sd, f'{prefix}.{i}', stride=2, pad=1)))
        i += 3
    for _ in range(n_blocks):
        x = _resblock(x, sd, f'{prefix}.{i}')
        i += 1
    for _ in range(n_down):
        x = F.relu(_in(_convT(x, sd, f'{prefix}.{i}')))
        i += 3
    if head:
        x = torch.tanh(_conv(x, sd, f'{prefix}.{i + 1}', reflect=3))
    return x


def generator(x, sd, netG, n_downsample_global, n_blocks_global, n_local_enhancers=1, n_blocks_local=3):
    """define_G(...)(x) for netG in {'global', 'local'} (networks.py:27-44)"""
    if netG == 'global':
        return _global_trunk(x, sd, 'model', n_downsample_global, n_blocks_global, True)
    assert netG == 'local'
    pyramid = [x]
    for _ in range(n_local_enhancers):
        pyramid.append(_pool(pyramid[-1]))
    out = _global_trunk(pyramid[-1], sd, 'model', n_downsample_global, n_blocks_global, False)
    for n in range(1, n_local_enhancers + 1):          # networks.py:174-181
        xi = pyramid[n_local_enhancers - n]
        d = f'model{n}_1'
        h = F.relu(_in(_conv(xi, sd, d + '.1', reflect=3)))
        h = F.relu(_in(_conv(h, sd, d + '.4', stride=2, pad=1)))
        h = h + out
        u = f'model{n}_2'
        for b in range(n_blocks_local):
            h = _resblock(h, sd, f'{u}.{b}')
        h = F.relu(_in(_convT(h, sd, f'{u}.{n_blocks_local}')))
        if n == n_local_enhancers:
            h = torch.tanh(_conv(h, sd, f'{u}.{n_blocks_local + 4}', reflect=3))
        out = h
    return out


def discriminator(x, sd, num_D, n_layers, get_interm_feat):
    """MultiscaleDiscriminator.forward: list (finest scale first) of feature lists (networks.py:319-331)"""
    result = []
    for i in range(num_D):
        s = num_D - 1 - i
        if get_interm_feat:
            keys = [f'scale{s}_layer{j}.0' for j in range(n_layers + 2)]
        else:
            keys = [f'layer{s}.{j}' for j in [0] + [2 + 3 * k for k in range(n_layers + 1)]]
        feats, h = [], x
        for j, key in enumerate(keys):
            stride = 2 if j < n_layers else 1
            h = _conv(h, sd, key, stride=stride, pad=2)
            if 0 < j <= n_layers:
                h = _in(h)
            if j <= n_layers:
                h = F.leaky_relu(h, 0.2)
            feats.append(h)
        result.append(feats if get_interm_feat else [feats[-1]])
        if i != num_D - 1:
            x = _pool(x)
    return result


def gan_loss(preds, target_is_real):
    """sum over scales of MSE(pred[-1], 1 or 0) (networks.py:100-110)"""
    t = 1.0 if target_is_real else 0.0
    return sum(((p[-1] - t) ** 2).mean() for p in preds)


def feature_matching(pred_fake, pred_real, num_D, n_layers_D, lambda_feat):
    """pix2pixHD_model.py:390-398"""
    w = (4.0 / (n_layers_D + 1)) * (1.0 / num_D) * lambda_feat
    loss = 0
    for i in range(num_D):
        for j in range(len(pred_fake[i]) - 1):
            loss = loss + w * (pred_fake[i][j] - pred_real[i][j].detach()).abs().mean()
    return loss


# ---- spectrogram encode / decode -------------------------------------------------------------------------
def _db(x, amin):
    return 20.0 * torch.log10(torch.clamp(x, min=amin)) - 20.0     # torchaudio amplitude_to_DB(x, 20, amin, 1)


def to_spectro(audio, cfg, mask):
    """pix2pixHD_model.py:142-227 with the transform = MDCT4 (float64 numpy oracle, cast to float32 like MDCT2's
    output); mask_mode None (zeros) is the deterministic variant the fixtures use"""
    w = M4.kbdwin(cfg['win_length'])
    if cfg.get('transform', 'mdct4') == 'mdct2':        # the shipped wiring (pix2pixHD_model.py:37-38)
        spec = M2.mdct2_forward(audio.numpy(), w, cfg['n_fft'], cfg['hop_length'], True, signal_len=len(audio))
    else:
        spec = M4.mdct4_forward(audio.numpy(), w, cfg['n_fft'], cfg['hop_length'], True, signal_len=len(audio))
    spectro = torch.from_numpy(spec.astype(np.float32)).unsqueeze(1).permute(0, 1, 3, 2)
    a, amin = cfg['alpha'], cfg['min_value']
    if cfg['explicit_encoding']:
        neg = 0.5 * (spectro.abs() - spectro)
        pos = spectro + neg
        logs = torch.cat((_db(a * pos + (1 - a) * neg, amin), _db((1 - a) * pos + a * neg, amin)), dim=1)
    else:
        logs = _db(spectro.abs() + amin, amin)
    pha = torch.sign(spectro)
    hi, lo = logs.max(), logs.min()
    norm = {'max': hi, 'min': lo, 'mean': logs.mean(), 'std': logs.var().sqrt()}
    if cfg.get('use_time_D'):
        norm['frames'] = frames_of(audio, cfg)
    logs = (logs - lo) / (hi - lo)
    if mask:
        size = logs.size()
        mask_size = int(size[2] * (1 - 1 / cfg['up_ratio']))
        if mask_size > 0:
            logs = torch.cat((logs[:, :, :-mask_size, :], torch.zeros(size[0], size[1], mask_size, size[3])), dim=2)
    return logs, pha, norm


def to_audio(log_spectro, norm, cfg):
    """explicit-encoding branch of pix2pixHD_model.py:229-249 with IMDCT4"""
    s = log_spectro.abs() * (norm['max'] - norm['min']) + norm['min']
    amp = 10.0 * torch.pow(10.0, 0.05 * s) - cfg['min_value']          # DB_to_amplitude(., 10, 0.5) - min_value
    spec = (amp[..., 0, :, :] - amp[..., 1, :, :]) / (2 * cfg['alpha'] - 1)
    w = M4.kbdwin(cfg['win_length'])
    inverse = M2.imdct2_forward if cfg.get('transform', 'mdct4') == 'mdct2' else M4.imdct4_forward
    y = inverse(spec.permute(0, 2, 1).double().numpy(), w, cfg['n_fft'], cfg['hop_length'], True)
    return math.sqrt(cfg['up_ratio'] - 1) * torch.from_numpy(y.astype(np.float32)).reshape(spec.shape[0], 1, 1, -1)


# ---- time-domain side (frames) ---------------------------------------------------------------------------
def frames_of(audio, cfg):
    """the windowed frames [B, F, win] MDCT2.forward(signal, return_ola=True) hands back (models/mdct.py:377-394; the
    same pad / unfold / window lines open MDCT4.forward, :488-503)"""
    hop, win = cfg['hop_length'], cfg['win_length']
    start_pad = hop                                               # center=True
    r = len(audio) % hop
    end_pad = start_pad + (hop - r if r else 0)
    x = F.pad(audio, (start_pad, end_pad))
    return x.unfold(-1, win, hop) * torch.from_numpy(M4.kbdwin(win).astype(np.float32))


def decode_coefficients(log_spectro, norm, cfg):
    """denormalize + explicit decode + permute (pix2pixHD_model.py:229-232, 253-257), differentiable"""
    s = log_spectro.abs() * (norm['max'] - norm['min']) + norm['min']
    amp = 10.0 * torch.pow(10.0, 0.05 * s) - cfg['min_value']
    spec = (amp[..., 0, :, :] - amp[..., 1, :, :]) / (2 * cfg['alpha'] - 1)
    return spec.permute(0, 2, 1)


def to_frames(log_spectro, norm, cfg):
    """pix2pixHD_model.py:251-258 with `_idct` = IMDCT4's per-frame inverse (models/mdct.py:547-549, twiddles :539-540)
    scaled 8/n_fft: IDCT_2N is twice the frame IMDCT2 windows and overlap-adds (mdct.py:440), IMDCT4 overlap-adds
    (4/n_fft) * this frame (mdct.py:560).  Differentiable torch restatement, float64 inside, float32 out."""
    N = cfg['n_fft']
    X = decode_coefficients(log_spectro, norm, cfg).double()
    if cfg.get('transform', 'mdct4') == 'mdct2':        # `_idct` = IDCT_2N_native as shipped (dct_native.py:44-67)
        n = torch.arange(N, dtype=torch.float64)
        C = torch.cos(math.pi * (2 * n[None, :] + 1) * n[:, None] / (2 * N))       # [k, n]
        return (2.0 * X @ C - X[..., :1]).float()
    exp1 = torch.exp(-1j * (math.pi / (2 * N) + math.pi / 4) * torch.arange(1, N, 2, dtype=torch.float64))
    exp2 = torch.exp(-1j * math.pi / (2 * N) * torch.arange(0, 2 * N, 2, dtype=torch.float64))
    y = torch.real(torch.fft.fft(exp1 * X, n=N) * exp2)
    return (8.0 / N * y).float()


def match_loss(sr_frames, cfg):
    """pix2pixHD_model.py:408-415"""
    half = cfg['win_length'] // 2
    w = torch.from_numpy(M4.kbdwin(cfg['win_length']).astype(np.float32))
    a = sr_frames[..., :-1, half:] * w[:half]
    b = sr_frames[..., 1:, :half] * w[half:]
    return ((a - b) ** 2).mean() * cfg['lambda_mat']


def time_losses(sdT, lr_frames, hr_frames, sr_frames, cfg):
    """pix2pixHD_model.py:375-387 with discriminate_time_D (:314-320) -> (G_GAN_t, D_real_t, D_fake_t)"""
    d, amin = cfg['D'], cfg['min_value']
    w = torch.from_numpy(M4.kbdwin(cfg['win_length']).astype(np.float32))
    run = lambda x: discriminator(x, sdT, d['num_D'], d['n_layers_D'], False)
    sr = (math.sqrt(cfg['up_ratio'] - 1) * w * sr_frames).unsqueeze(1)
    lr, hr = lr_frames.unsqueeze(1), hr_frames.unsqueeze(1)
    db = lambda x: _db(x.abs(), amin)
    loss_D_fake_t = gan_loss(run(torch.cat((db(lr), db(sr.detach())), dim=1)), False) * cfg['lambda_time']
    loss_D_real_t = gan_loss(run(torch.cat((db(lr), db(hr)), dim=1)), True) * cfg['lambda_time']
    loss_G_t = gan_loss(run(torch.cat((lr, sr), dim=1)), True) * cfg['lambda_time']       # :386: raw frames, no dB
    return loss_G_t, loss_D_real_t, loss_D_fake_t


# ---- model + optimisation step ---------------------------------------------------------------------------
def model_losses(sdG, sdD, lr_spectro, hr_spectro, cfg):
    """pix2pixHD_model.py:348-398 -> (G_GAN, G_GAN_Feat, D_real, D_fake, sr_result)"""
    g = cfg['G']
    sr = generator(lr_spectro, sdG, g['netG'], g['n_downsample_global'], g['n_blocks_global'], g['n_local_enhancers'],
                   g['n_blocks_local'])
    d = cfg['D']
    run_D = lambda x: discriminator(x, sdD, d['num_D'], d['n_layers_D'], True)
    pred_fake_pool = run_D(torch.cat((lr_spectro, sr.detach()), dim=1))
    loss_D_fake = gan_loss(pred_fake_pool, False)
    pred_real = run_D(torch.cat((lr_spectro, hr_spectro), dim=1))
    loss_D_real = gan_loss(pred_real, True)
    pred_fake = run_D(torch.cat((lr_spectro, sr), dim=1))
    loss_G_GAN = gan_loss(pred_fake, True)
    loss_feat = feature_matching(pred_fake, pred_real, d['num_D'], d['n_layers_D'], cfg['lambda_feat'])
    return loss_G_GAN, loss_feat, loss_D_real, loss_D_fake, sr


class TrainState:
    """weights + two Adam optimisers (pix2pixHD_model.py:131, 140) stepping as train.py:155-184 does; with sdT (the
    time-domain discriminator's weights, cfg['use_time_D']) and / or cfg['use_match_loss'] the frame losses of
    pix2pixHD_model.py:375-387, 408-415 join loss_G / loss_D as train.py:159-160 adds them"""

    def __init__(self, sdG, sdD, cfg, lr=2e-4, beta1=0.5, sdT=None):
        self.cfg = cfg
        self.G = {k: v.clone().float().requires_grad_(True) for k, v in sdG.items()}
        self.D = {k: v.clone().float().requires_grad_(True) for k, v in sdD.items()}
        self.T = {k: v.clone().float().requires_grad_(True) for k, v in (sdT or {}).items()}
        self.opt_G = torch.optim.Adam(list(self.G.values()), lr=lr, betas=(beta1, 0.999))
        self.opt_D = torch.optim.Adam(list(self.D.values()) + list(self.T.values()), lr=lr, betas=(beta1, 0.999))

    def step(self, lr_audio, hr_audio, spectros=None):
        cfg = self.cfg
        lnorm = hnorm = None
        if spectros is None:
            with torch.no_grad():
                hs, _, hnorm = to_spectro(hr_audio, cfg, False)
                ls, _, lnorm = to_spectro(lr_audio, cfg, True)
        else:
            ls, hs = spectros
        g_gan, g_feat, d_real, d_fake, sr = model_losses(self.G, self.D, ls, hs, cfg)
        loss_D = (d_fake + d_real) * 0.5
        loss_G = g_gan + g_feat
        extra = {}
        if cfg.get('use_time_D') or cfg.get('use_match_loss'):
            sr_frames = to_frames(sr, lnorm, cfg)
            if cfg.get('use_match_loss'):
                extra['G_mat'] = match_loss(sr_frames, cfg)
                loss_G = loss_G + extra['G_mat']
            if cfg.get('use_time_D'):
                g_t, d_real_t, d_fake_t = time_losses(self.T, lnorm['frames'], hnorm['frames'], sr_frames, cfg)
                extra.update(G_GAN_t=g_t, D_real_t=d_real_t, D_fake_t=d_fake_t)
                loss_G = loss_G + g_t
                loss_D = loss_D + (d_fake_t + d_real_t) * 0.5
        self.opt_G.zero_grad()
        loss_G.backward()
        gG = {k: v.grad.clone() for k, v in self.G.items()}
        self.opt_G.step()
        self.opt_D.zero_grad()
        loss_D.backward()
        gD = {k: v.grad.clone() for k, v in self.D.items()}
        self.last_gT = {k: v.grad.clone() for k, v in self.T.items()}
        self.last_extra = {k: float(v.detach()) for k, v in extra.items()}
        self.opt_D.step()
        return [float(g_gan.detach()), float(g_feat.detach()), float(d_real.detach()), float(d_fake.detach())], sr.detach(), gG, gD


def random_state_dicts(ngf=48, ndf=64, input_nc=2, output_nc=2, n_down=4, n_blocks_global=3, n_blocks_local=2, num_D=2,
                       n_layers_D=3, seed=1234):
    """G3L2-shaped LocalEnhancer + multiscale D weights drawn N(0, 0.02) (weights_init, networks.py:10-16); biases
    uniform as torch.nn.Conv2d does.  For timing and smoke runs only (not the reference's RNG stream)."""
    g = torch.Generator().manual_seed(seed)
    sdG, sdD = {}, {}

    def add(sd, key, cout, cin, k, transposed=False):
        shape = (cin, cout, k, k) if transposed else (cout, cin, k, k)
        sd[key + '.weight'] = torch.randn(shape, generator=g) * 0.02
        bound = 1 / math.sqrt((cout if transposed else cin) * k * k)
        sd[key + '.bias'] = (torch.rand(cout, generator=g) * 2 - 1) * bound

    c = ngf * 2
    add(sdG, 'model.1', c, input_nc, 7)
    i = 4
    for _ in range(n_down):
        add(sdG, f'model.{i}', c * 2, c, 3)
        c *= 2
        i += 3
    for _ in range(n_blocks_global):
        add(sdG, f'model.{i}.conv_block.1', c, c, 3)
        add(sdG, f'model.{i}.conv_block.5', c, c, 3)
        i += 1
    for _ in range(n_down):
        add(sdG, f'model.{i}', c // 2, c, 3, transposed=True)
        c //= 2
        i += 3
    add(sdG, 'model1_1.1', ngf, input_nc, 7)
    add(sdG, 'model1_1.4', ngf * 2, ngf, 3)
    for b in range(n_blocks_local):
        add(sdG, f'model1_2.{b}.conv_block.1', ngf * 2, ngf * 2, 3)
        add(sdG, f'model1_2.{b}.conv_block.5', ngf * 2, ngf * 2, 3)
    add(sdG, f'model1_2.{n_blocks_local}', ngf, ngf * 2, 3, transposed=True)
    add(sdG, f'model1_2.{n_blocks_local + 4}', output_nc, ngf, 7)
    for s in range(num_D):
        nf, prev = ndf, input_nc + output_nc
        for j in range(n_layers_D + 2):
            cout = 1 if j == n_layers_D + 1 else nf
            add(sdD, f'scale{s}_layer{j}.0', cout, prev, 4)
            prev, nf = cout, min(nf * 2, 512)
    return sdG, sdD


def random_time_D_state_dict(ndf=64, num_D=2, n_layers_D=3, seed=4321):
    """time_D = define_D(2, ndf, n_layers_D, ..., num_D, getIntermFeat=False) (pix2pixHD_model.py:64-65): keys
    layer{s}.{0,2,5,8,11}.*; same N(0, 0.02) draw as random_state_dicts.  For timing and smoke runs only."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    idx = [0] + [2 + 3 * k for k in range(n_layers_D + 1)]
    for s in range(num_D):
        nf, prev = ndf, 2
        for j, i in enumerate(idx):
            cout = 1 if j == n_layers_D + 1 else nf
            sd[f'layer{s}.{i}.weight'] = torch.randn((cout, prev, 4, 4), generator=g) * 0.02
            bound = 1 / math.sqrt(prev * 16)
            sd[f'layer{s}.{i}.bias'] = (torch.rand(cout, generator=g) * 2 - 1) * bound
            prev, nf = cout, min(nf * 2, 512)
    return sd
