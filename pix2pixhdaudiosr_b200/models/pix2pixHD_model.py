"""Drop-in `Pix2PixHDModel` (reference: models/pix2pixHD_model.py:15-455, 511-569): audio -> lapped-transform
spectrogram -> dB / +- phase encoding -> generator -> multiscale discriminator -> LSGAN + feature-matching
losses, with the same `forward(lr_audio, inst, hr_audio, feat, infer)` / `inference(lr_audio, inst)` signatures,
loss order and `optimizer_G` / `optimizer_D` attributes, executed by the sm_100a kernels behind include/p2p_b200.h.

Differences from the reference, all deliberate:
* the transform is MDCT4/IMDCT4 by default (README.md:133 invites it); `opt.transform = 'mdct2'` selects what
  the shipped model wires (MDCT2 / IMDCT2 / IDCT_2N_native, :35-40) -- then every line below is the reference's;
* `opt.fp16` selects bf16 storage with fp32 accumulation / master weights instead of fp16 autocast + GradScaler;
* the three per-step `.cpu().numpy()` copies of pix2pixHD_model.py:418-432 only happen when `infer` is set;
* the generator's pass through D runs with detached D weights (the reference computes and discards those
  weight gradients, train.py:164-176);
* `use_time_D` / `use_match_loss` (pix2pixHD_model.py:375-387, 408-415) run on MDCT4's own frames: the shipped model
  takes them from MDCT2(return_ola) / IDCT_2N; here `norm_param['frames']` are the windowed frames of mdct.py:497-503
  and `to_frames` is IMDCT4's per-frame inverse before window and overlap-add (mdct.py:547-549) times 8/n_fft, i.e.
  twice the frame IMDCT4 windows and overlap-adds -- the relation IDCT_2N has to IMDCT2 (mdct.py:440);
* `use_hifigan_D`, the instance-feature encoder and the VGG loss are not built.
"""
import ctypes
import math

import torch

from .. import _lib, ops
from ..optim import FusedAdam
from ..util.util import kbdwin
from . import networks
from .base_model import BaseModel
from ..dct.dct_native import DCT_2N_native, IDCT_2N_native
from .mdct import IMDCT2, IMDCT4, MDCT2, MDCT4, _Plan, _Plan2, _Synthesis2Fn, _SynthesisFn


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class Pix2PixHDModel(BaseModel):
    def name(self):
        return 'Pix2PixHDModel'

    def init_loss_filter(self, use_gan_feat_loss, use_vgg_loss, use_match_loss, use_time_loss):
        flags = (True, use_gan_feat_loss, use_vgg_loss, use_match_loss, use_time_loss, use_time_loss, use_time_loss, True, True)

        def loss_filter(*losses):
            return [l for (l, f) in zip(losses, flags) if f]
        return loss_filter

    def initialize(self, opt):
        BaseModel.initialize(self, opt)
        for flag in ('use_hifigan_D', 'instance_feat', 'label_feat'):
            if getattr(opt, flag, False):
                raise NotImplementedError('--%s is outside the path built so far (DESIGN.md §5)' % flag)
        self.use_frames = bool(opt.isTrain and (opt.use_time_D or opt.use_match_loss))
        if self.use_frames:
            if not opt.explicit_encoding:
                raise ValueError('use_time_D / use_match_loss need explicit_encoding: to_frames returns None otherwise '
                                 '(pix2pixHD_model.py:251-258)')
            if opt.win_length != opt.n_fft:
                raise NotImplementedError('use_time_D / use_match_loss: win_length must equal n_fft')
        self.transform = getattr(opt, 'transform', 'mdct4')
        if self.transform not in ('mdct4', 'mdct2'):
            raise ValueError("transform %r: 'mdct4' (SURVEY H1) or 'mdct2' (what the shipped model wires, :35-40)" % self.transform)
        if not opt.no_vgg_loss and opt.isTrain:
            raise NotImplementedError('the VGG loss is deprecated for audio (README.md:18): pass no_vgg_loss=True')
        if opt.pool_size > 0:
            raise NotImplementedError('pool_size > 0 (fake-image history) is unused by every audio recipe')
        precision = getattr(opt, 'precision', None) or ('bf16' if opt.fp16 else 'fp32')
        self.act_dtype = {'fp32': torch.float32, 'bf16': torch.bfloat16}[precision]
        input_nc = opt.label_nc if opt.label_nc != 0 else opt.input_nc

        # lapped transform (pix2pixHD_model.py:35-40)
        self.up_ratio = opt.hr_sampling_rate / opt.lr_sampling_rate
        self.window = kbdwin(opt.win_length)
        if self.transform == 'mdct2':       # :37-40 as shipped: DCT-II frames of n_fft bins
            self._dct, self._idct = DCT_2N_native(), IDCT_2N_native()
            self._mdct = MDCT2(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=self.window,
                               device=self.device, dct_op=self._dct)
            self._imdct = IMDCT2(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=self.window,
                                 device=self.device, idct_op=self._idct)
        else:
            self._mdct = MDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=self.window,
                               device=self.device)
            self._imdct = IMDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=self.window,
                                 device=self.device)
        # to_frames: the synthesis kernel with hop = win = n_fft and a window of ones gives the F frames side by side;
        # a second plan carries sqrt(up_ratio-1) * window (:375)
        self._frame_plan = self._frame_plan_scaled = None

        # networks (:42-60)
        netG_input_nc = input_nc + (0 if opt.no_instance else 1)
        networks.set_default_precision(precision)
        try:
            self.netG = networks.define_G(netG_input_nc, opt.output_nc, opt.ngf, opt.netG, opt.n_downsample_global,
                                          opt.n_blocks_global, opt.n_local_enhancers, opt.n_blocks_local, opt.norm,
                                          gpu_ids=self.gpu_ids)
            if self.isTrain:
                netD_input_nc = input_nc + opt.output_nc + (0 if opt.no_instance else 1)
                self.netD = networks.define_D(netD_input_nc, opt.ndf, opt.n_layers_D, opt.norm, opt.no_lsgan, opt.num_D,
                                              not opt.no_ganFeat_loss, gpu_ids=self.gpu_ids)
                if opt.use_time_D:      # :64-65
                    # its [frames, n_fft] maps are wide: run it on the transposed image (same results, same state_dict)
                    self.time_D = networks.define_D(2, opt.ndf, opt.n_layers_D, opt.norm, opt.no_lsgan, opt.num_D, False,
                                                    gpu_ids=self.gpu_ids, transposed_image=True)
        finally:
            networks.set_default_precision('fp32')
        if opt.verbose:
            print('---------- Networks initialized -------------')

        if not self.isTrain or opt.continue_train or opt.load_pretrain:
            pretrained_path = '' if not self.isTrain else opt.load_pretrain
            self.load_network(self.netG, 'G', opt.which_epoch, pretrained_path)
            if self.isTrain:
                self.load_network(self.netD, 'D', opt.which_epoch, pretrained_path)
                if opt.use_time_D:
                    self.load_network(self.time_D, 'time_D', opt.which_epoch, pretrained_path)

        if self.isTrain:
            self.old_lr = opt.lr
            self.loss_filter = self.init_loss_filter(not opt.no_ganFeat_loss, not opt.no_vgg_loss, opt.use_match_loss,
                                                     opt.use_hifigan_D or opt.use_time_D)
            self.criterionGAN = networks.GANLoss(use_lsgan=not opt.no_lsgan)
            self.loss_names = self.loss_filter('G_GAN', 'G_GAN_Feat', 'G_VGG', 'G_mat', 'G_GAN_t', 'D_real_t', 'D_fake_t',
                                               'D_real', 'D_fake')
            if opt.niter_fix_global > 0:
                params = [v for k, v in self.netG.named_parameters() if k.startswith('model' + str(opt.n_local_enhancers))]
                print('------------- Only training the local enhancer network (for %d epochs) ------------' % opt.niter_fix_global)
            else:
                params = list(self.netG.parameters())
            print('Total number of parameters of G: %d' % (sum([param.numel() for param in params])))
            d_params = list(self.netD.parameters())
            if opt.use_time_D:          # :137-138
                d_params += list(self.time_D.parameters())
            print('Total number of parameters of D: %d' % (sum([param.numel() for param in d_params])))
            if len(self.gpu_ids) > 0:
                self.optimizer_G = FusedAdam(params, lr=opt.lr, betas=(opt.beta1, 0.999))
                self.optimizer_D = FusedAdam(d_params, lr=opt.lr, betas=(opt.beta1, 0.999))
            else:   # inspectable on a CPU-only host; nothing can step without the GPU
                self.optimizer_G = self.optimizer_D = None

    # ------------------------------------------------------------------------------------------------
    # spectrogram encode / decode (pix2pixHD_model.py:142-258)
    # ------------------------------------------------------------------------------------------------
    def _mask_noise(self, B, mask_size, F, C, device):
        """the high-band filler of :201-221, drawn channels-last"""
        mode = self.opt.mask_mode
        if mode is None:
            return None
        noise = torch.randn(B, mask_size, F, C, device=device)
        lo, hi = noise.min(), noise.max()
        if mode == 'mode0':
            return noise / (hi - lo)
        if mode == 'mode1':
            sign = 2 * torch.randint(low=0, high=2, size=noise.size(), device=device) - 1
            return (noise - lo) / (hi - lo) * sign
        if mode == 'mode2':
            return (noise - lo) / (hi - lo)
        raise ValueError('unknown mask_mode %r' % (mode,))

    def to_spectro(self, audio, mask=False):
        """audio [B,T] -> (log_spectro [B,C,H,W] in [0,1], pha [B,1,H,W], norm_param); H = bins, W = frames.
        The returned images are channels-last in memory (what the conv kernels read) behind NCHW shapes."""
        opt = self.opt
        audio = audio.to(self.device)
        spec = self._mdct(audio)                                        # [B, F, M] fp32, kernel: mdct4_analysis
        if spec.dim() != 3:
            raise ValueError('to_spectro expects audio shaped [B, T]')
        B, F, M = spec.shape
        frames = None
        if getattr(opt, 'use_time_D', False):                           # :143-145: the windowed frames, [B, F, win]
            start_pad, _ = self._mdct.pads(int(len(audio)))
            frames = ops.frame_window(audio, self._mdct.window.to(audio.device, torch.float32), F, opt.hop_length, start_pad)
        C = 2 if opt.explicit_encoding else 1
        dev = spec.device
        logs = torch.empty((B, M, F, C), dtype=torch.float32, device=dev)
        pha = torch.empty((B, 1, M, F), dtype=torch.float32, device=dev)
        work = torch.empty(8, dtype=torch.float32, device=dev)
        norm = torch.empty(4, dtype=torch.float32, device=dev)
        lib = _lib.lib()
        _lib.check(lib.p2p_spectro_encode(_p(spec), _p(logs), _p(pha), _p(work), B, F, M, C, opt.alpha, opt.min_value, _stream()))
        if not opt.explicit_encoding and opt.phase_encoding_mode is not None:      # :179-192
            mode = opt.phase_encoding_mode
            if mode == 'uni_dist':
                pha = pha * torch.rand(pha.size(), device=dev)
            elif mode == 'norm_dist':
                n = torch.randn(pha.size(), device=dev)
                pha = pha * ((n - n.min()) / (n.max() - n.min()))
            elif mode == 'norm_dist2':
                pha = pha * torch.randn(pha.size(), device=dev).abs()
            elif mode == 'scale':
                pha = pha * 0.5
        mask_size, noise = 0, None
        if mask:
            mask_size = int(M * (1 - 1 / self.up_ratio))
            noise = self._mask_noise(B, mask_size, F, C, dev) if mask_size > 0 else None
        _lib.check(lib.p2p_spectro_normalize(_p(logs), _p(work), _p(norm), _p(noise), B, F, M, C, mask_size,
                                             int(bool(mask) and mask_size > 0), _stream()))
        norm_param = {'max': norm[0], 'min': norm[1], 'mean': norm[2], 'std': norm[3], 'frames': frames}
        return logs.permute(0, 3, 1, 2), pha, norm_param

    def _decode(self, log_spectro, norm_param, pha, gain):
        """denormalize (:229-232) + explicit decode (:236-237) + permute(0,2,1) (:246-248) -> [B, F, M] coefficients"""
        opt = self.opt
        x = log_spectro.detach()
        if x.dtype not in (torch.float32, torch.bfloat16):
            x = x.float()
        B, C, M, F = x.shape
        norm = torch.stack((norm_param['max'], norm_param['min'])).to(x.device, torch.float32)
        coef = torch.empty((B, F, M), dtype=torch.float32, device=x.device)
        sb, sc, sh, sw = x.stride()
        if pha is not None:
            pha = pha.to(x.device, torch.float32).contiguous()
        _lib.check(_lib.lib().p2p_spectro_decode(_p(x), sb, sc, sh, sw, _p(norm), _p(pha), _p(coef), B, C, F, M, opt.alpha,
                                                 opt.min_value, gain, ops._DT[x.dtype], _stream()))
        return coef

    def denormalize(self, log_spectro, norm_param):
        """|x| (max - min) + min, then DB_to_amplitude(., 10, 0.5) - min_value (element-wise; used by callers only)"""
        lo, hi = norm_param['min'].to(log_spectro.device), norm_param['max'].to(log_spectro.device)
        spectro = torch.abs(log_spectro.float()) * (hi - lo) + lo
        return 10.0 * torch.pow(10.0, 0.05 * spectro) - self.opt.min_value

    def to_audio(self, log_spectro, norm_param, pha=None):
        """spectrogram image -> audio [B,1,1,T] (:234-249), one decode kernel + the IMDCT4 synthesis kernel"""
        log_spectro = log_spectro.to(self.device)
        if not self.opt.explicit_encoding:
            if self.up_ratio > 1 and pha is not None:
                size = pha.size(-2)
                keep = int(size * (1 / self.up_ratio))
                psudo_pha = (2 * torch.randint(low=0, high=2, size=pha.size(), device=pha.device) - 1).to(pha.dtype)
                pha = torch.cat((pha[..., :keep, :], psudo_pha[..., keep:, :]), dim=-2)
            else:
                pha = None
        else:
            pha = None
        with torch.no_grad():
            coef = self._decode(log_spectro, norm_param, pha, math.sqrt(self.up_ratio - 1))
            return self._imdct(coef)

    def _decode_with_grad(self, log_spectro, norm_param):
        """denormalize + explicit decode + permute (:229-232, 253-257) -> coefficients [B, F, M], differentiable"""
        log_spectro = log_spectro.to(self.device)
        norm = torch.stack((norm_param['max'], norm_param['min'])).to(log_spectro.device, torch.float32)
        return ops.SpectroDecodeFn.apply(log_spectro, norm, self.opt.alpha, self.opt.min_value, 1.0)

    def _frames_from_coef(self, coef, plan):
        """synthesis kernel at hop = win = n_fft: no overlap, the F frames land side by side -> [B, F, n_fft]"""
        B, F, _ = coef.shape
        N = self.opt.n_fft
        if self.transform == 'mdct2':       # IDCT_2N_native on every frame (:257)
            return _Synthesis2Fn.apply(coef, plan, F * N, 0, 1.0, 0.5).view(B, F, N)
        return _SynthesisFn.apply(coef, plan, F * N, 0, 8.0 / N).view(B, F, N)

    def _frame_plans(self):
        if self._frame_plan is None:
            N = self.opt.n_fft
            P = _Plan2 if self.transform == 'mdct2' else _Plan
            self._frame_plan = P(N, N, N, torch.ones(N))
            self._frame_plan_scaled = P(N, N, N, math.sqrt(self.up_ratio - 1) * self.window.detach().cpu().float())
        return self._frame_plan, self._frame_plan_scaled

    def to_frames(self, log_spectro, norm_param):
        """spectrogram image -> per-frame time signals [B, F, n_fft] with a gradient (:251-258): decode kernel +
        synthesis kernel (no window, no overlap-add)"""
        if not self.opt.explicit_encoding:
            return None
        return self._frames_from_coef(self._decode_with_grad(log_spectro, norm_param), self._frame_plans()[0])

    def discriminate_time_D(self, label_frames, test_frames):
        """time-domain discriminator on dB(|label|), dB(|test.detach()|) stacked on channels (:314-320); frames [B,1,F,N]"""
        amin = self.opt.min_value
        label = ops.FramesDbFn.apply(label_frames, None, 1.0, amin)
        test = ops.FramesDbFn.apply(test_frames.detach(), None, 1.0, amin)
        return self.time_D.forward(torch.cat((label, test), dim=1))

    def encode_input(self, lr_audio, inst_map=None, hr_audio=None, feat_map=None):
        if hr_audio is not None:
            with torch.no_grad():
                hr_spectro, hr_pha, hr_norm_param = self.to_spectro(hr_audio, mask=False)
        else:
            hr_spectro = hr_pha = hr_norm_param = None
        with torch.no_grad():
            lr_spectro, lr_pha, lr_norm_param = self.to_spectro(lr_audio, mask=True)
        if not self.opt.no_instance:
            lr_spectro = torch.cat((lr_spectro, inst_map.to(self.device)), dim=1)
        return lr_spectro, lr_pha, hr_spectro, hr_pha, feat_map, inst_map, hr_norm_param, lr_norm_param

    def discriminate_F(self, input_label, test_image, use_pool=False):
        """frequency-domain discriminator on (label, detached image) (:304-312; the pool is the identity)"""
        return self.netD.forward(torch.cat((input_label, test_image.detach().float()), dim=1))

    # ------------------------------------------------------------------------------------------------
    def forward(self, lr_audio, inst, hr_audio, feat, infer=False):
        opt = self.opt
        lr_spectro, lr_pha, hr_spectro, hr_pha, _, _, hr_norm_param, lr_norm_param = self.encode_input(lr_audio, inst, hr_audio, feat)
        sr_result = self.netG.forward(lr_spectro)

        if getattr(self, '_share_fake_pass', False) and getattr(self, '_batch_D', False):
            # train.train_step with train.BATCH_D: the shared pass below AND the pass on the real image run as ONE discriminator
            # forward over cat(real, fake) along the batch (InstanceNorm2d is per image: each image sees the arithmetic of
            # its own pass); the features are split back into their halves.  loss_G's backward walks the fake half only
            # (ops.set_backward_phase(fake_half_only=True)), loss_D's the whole batch.
            B = lr_spectro.shape[0]
            real_in = torch.cat((lr_spectro, hr_spectro.detach().float()), dim=1)
            fake_in = ops.GradGateFn.apply(torch.cat((lr_spectro, sr_result.float()), dim=1))
            old = ops.set_batch_halves(B)
            try:
                both = self.netD.forward(torch.cat((real_in, fake_in), dim=0), gated=True)
            finally:
                ops.set_batch_halves(old)
            split = [[ops.BatchHalvesFn.apply(f, B) for f in scale] for scale in both]
            pred_real = [[rf[0] for rf in scale] for scale in split]
            pred_fake = [[rf[1] for rf in scale] for scale in split]
            loss_D_fake = self.criterionGAN(pred_fake, False)
            loss_D_real = self.criterionGAN(pred_real, True)
            loss_G_GAN = self.criterionGAN(pred_fake, True)
        elif getattr(self, '_share_fake_pass', False):
            # train.train_step: D's forward on (label, fake) runs ONCE and serves both losses -- the reference's first and third
            # D passes (:351-361) compute the same numbers (pool_size 0).  The gate in front of D is closed while loss_D is
            # back-propagated (the reference detaches the fake image there), and D's weight gradients are not formed while
            # loss_G is (ops.set_backward_phase).  The caller must back-propagate loss_G with retain_graph=True.
            fake_in = ops.GradGateFn.apply(torch.cat((lr_spectro, sr_result.float()), dim=1))
            pred_fake = self.netD.forward(fake_in, gated=True)
            loss_D_fake = self.criterionGAN(pred_fake, False)
            pred_real = self.discriminate_F(lr_spectro, hr_spectro)
            loss_D_real = self.criterionGAN(pred_real, True)
            loss_G_GAN = self.criterionGAN(pred_fake, True)
        else:
            pred_fake_pool = self.discriminate_F(lr_spectro, sr_result, use_pool=True)
            loss_D_fake = self.criterionGAN(pred_fake_pool, False)
            pred_real = self.discriminate_F(lr_spectro, hr_spectro)
            loss_D_real = self.criterionGAN(pred_real, True)
            # the generator's pass: gradients flow to G through D's activations; D's own weight gradients are not formed
            pred_fake = self.netD.forward(torch.cat((lr_spectro, sr_result.float()), dim=1), needs_wgrad=False)
            loss_G_GAN = self.criterionGAN(pred_fake, True)

        # time-domain GAN on frames (:375-387)
        loss_G_GAN_time = loss_D_real_time = loss_D_fake_time = 0
        sr_coef = self._decode_with_grad(sr_result, lr_norm_param) if self.use_frames else None
        if opt.use_time_D:
            # sqrt(up_ratio-1) * window * to_frames(sr): the window rides in the synthesis kernel
            sr_frames = self._frames_from_coef(sr_coef, self._frame_plans()[1]).unsqueeze(1)
            lr_frames = lr_norm_param['frames'].unsqueeze(1)
            hr_frames = hr_norm_param['frames'].unsqueeze(1)
            pred_fake_frames = self.discriminate_time_D(lr_frames, sr_frames)
            loss_D_fake_time = self.criterionGAN(pred_fake_frames, False) * opt.lambda_time
            pred_real_frames = self.discriminate_time_D(lr_frames, hr_frames)
            loss_D_real_time = self.criterionGAN(pred_real_frames, True) * opt.lambda_time
            # the generator's pass (:386 feeds the frames themselves, without the dB map): gradients reach G through
            # time_D's activations only; its weight gradients are discarded upstream (train.py:164-176)
            _pred_fake_frames = self.time_D.forward(torch.cat((lr_frames, sr_frames), dim=1), needs_wgrad=False)
            loss_G_GAN_time = self.criterionGAN(_pred_fake_frames, True) * opt.lambda_time

        loss_G_GAN_Feat = 0
        if not opt.no_ganFeat_loss:
            loss_G_GAN_Feat = networks.feature_matching_loss(pred_fake, pred_real, opt.num_D, opt.n_layers_D, opt.lambda_feat)

        # frame-match loss (:408-415): adjacent frames must agree on their overlap
        loss_G_match = 0
        if opt.explicit_encoding and opt.use_match_loss:
            raw_frames = self._frames_from_coef(sr_coef, self._frame_plans()[0])
            loss_G_match = ops.FrameMatchFn.apply(raw_frames, self._mdct.window.to(raw_frames.device, torch.float32),
                                                  float(opt.lambda_mat))

        if infer:   # the visualiser's host copies (:418-432) -- only when the caller asked for the image
            self.current_lable = lr_spectro.detach()[0].float().mean(0).cpu().numpy() if opt.explicit_encoding else lr_spectro.detach().cpu().numpy()[0, 0]
            self.current_generated = sr_result.detach()[0].float().mean(0).cpu().numpy() if opt.explicit_encoding else sr_result.detach().float().cpu().numpy()[0, 0]
            self.current_real = hr_spectro.detach()[0].float().mean(0).cpu().numpy() if opt.explicit_encoding else hr_spectro.detach().cpu().numpy()[0, 0]
        return [self.loss_filter(loss_G_GAN, loss_G_GAN_Feat, 0, loss_G_match, loss_G_GAN_time, loss_D_real_time,
                                 loss_D_fake_time, loss_D_real, loss_D_fake),
                None if not infer else sr_result]

    def inference(self, lr_audio, inst):
        lr_spectro, lr_pha, _, _, _, _, _, lr_norm_param = self.encode_input(lr_audio, inst, None)
        with torch.no_grad():
            sr_spectro = self.netG.forward(lr_spectro)
        return sr_spectro, lr_pha, lr_norm_param, lr_spectro

    def save(self, which_epoch):
        self.save_network(self.netG, 'G', which_epoch, self.gpu_ids)
        self.save_network(self.netD, 'D', which_epoch, self.gpu_ids)
        if getattr(self.opt, 'use_time_D', False):
            self.save_network(self.time_D, 'time_D', which_epoch, self.gpu_ids)

    def update_fixed_params(self):
        """after niter_fix_global epochs also train the global generator (:521-531)"""
        self.optimizer_G = FusedAdam(list(self.netG.parameters()), lr=self.opt.lr, betas=(self.opt.beta1, 0.999))

    def update_learning_rate(self):
        lrd = self.opt.lr / self.opt.niter_decay
        lr = self.old_lr - lrd
        for opt_ in (self.optimizer_D, self.optimizer_G):
            for g in opt_.param_groups:
                g['lr'] = lr
        self.old_lr = lr


class InferenceModel(Pix2PixHDModel):
    def forward(self, lr_audio, inst):
        return self.inference(lr_audio, inst)
