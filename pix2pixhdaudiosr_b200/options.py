"""The `opt` namespace the reference's models read (options/base_options.py:11-72, train_options.py:5-55,
audio_config.py).  Argument parsing itself is out of scope; `make_opt(**overrides)` returns the reference's
defaults for every field the hot path touches, `g3l2_opt()` the recipe of generated/GEN_VCTK_G3L2_48ngf/opt.txt."""
import argparse

_DEFAULTS = dict(
    name='label2city', gpu_ids=[0], checkpoints_dir='./checkpoints', model='pix2pixHD', norm='instance', verbose=False,
    fp16=False, seed=1234, batchSize=1, label_nc=35, input_nc=3, output_nc=3, resize_or_crop='scale_width', alpha=0.6,
    netG='global', ngf=64, n_downsample_global=4, n_blocks_global=9, n_blocks_local=3, n_local_enhancers=1,
    niter_fix_global=0, no_instance=False, instance_feat=False, label_feat=False, feat_num=3, load_features=False,
    mask=False, mask_mode=None, explicit_encoding=False, min_value=1e-7,
    isTrain=True, continue_train=False, load_pretrain='', which_epoch='latest', niter=100, niter_decay=100, beta1=0.5, lr=0.0002,
    phase_encoding_mode=None, num_D=2, n_layers_D=3, ndf=64, lambda_feat=10.0, lambda_mat=10.0, lambda_time=0.4,
    no_ganFeat_loss=False, no_vgg_loss=False, use_match_loss=False, no_lsgan=False, pool_size=0, use_hifigan_D=False,
    use_time_D=False, lr_sampling_rate=8000, hr_sampling_rate=48000, segment_length=32512, n_fft=512, hop_length=256,
    win_length=512, center=False, is_lr_input=False,
    # additions of this build: which lapped transform feeds the networks, and the arithmetic
    transform='mdct4', precision=None,
)


def make_opt(**overrides):
    d = dict(_DEFAULTS)
    unknown = set(overrides) - set(d)
    if unknown:
        raise TypeError("unknown option(s): %s" % sorted(unknown))
    d.update(overrides)
    return argparse.Namespace(**d)


def g3l2_opt(**overrides):
    """G3L2 ngf 48 (generated/GEN_VCTK_G3L2_48ngf/opt.txt) on MDCT4 spectrograms: n_fft 1024 / hop 512 gives the
    recipe's 512 frequency rows; segment 127 * 512 samples gives its 128 frames (shape S128)."""
    d = dict(netG='local', ngf=48, n_downsample_global=4, n_blocks_global=3, n_blocks_local=2, n_local_enhancers=1,
             input_nc=2, output_nc=2, label_nc=0, no_instance=True, no_vgg_loss=True, explicit_encoding=True,
             num_D=2, n_layers_D=3, ndf=64, mask=True, mask_mode='mode2', n_fft=1024, hop_length=512, win_length=1024,
             segment_length=127 * 512, center=True, name='G3L2_48ngf', fp16=True)
    d.update(overrides)
    return make_opt(**d)
