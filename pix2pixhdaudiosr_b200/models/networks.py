"""Drop-in `define_G` / `define_D` / `GANLoss` and the generator / discriminator classes of the reference's
models/networks.py (file:line cited per class), executed by the sm_100a kernels behind include/p2p_b200.h.

* Same constructor signatures, module tree and therefore the same `state_dict()` keys and logical shapes
  (`model.1.weight`, `model.16.conv_block.5.bias`, `scale0_layer3.0.weight`, ...), so reference checkpoints load.
* Same initialisation stream: parameters are drawn with the same torch RNG calls in the same order
  (Conv default init, then `weights_init` N(0, 0.02)), so equal seeds give equal weights.
* Activations are NHWC inside; module inputs / outputs keep the reference's logical NCHW shape (outputs are
  channels-last views).  Weights are stored tap-major `[KH][KW][Cout][Cin]` behind logical-shape views.
* `precision`: 'fp32' (exact CUDA-core path) or 'bf16' (bf16 storage, fp32 accumulate) -- module attribute
  `act_dtype`, default from `set_default_precision`.
"""
import math

import torch
import torch.nn as nn

from .. import ops

_DEFAULT_DTYPE = torch.float32


def set_default_precision(name):
    global _DEFAULT_DTYPE
    _DEFAULT_DTYPE = {"fp32": torch.float32, "bf16": torch.bfloat16}[name]


# --------------------------------------------------------------------------------------------------
# leaf modules (markers carry no parameters; the executor below fuses them)
# --------------------------------------------------------------------------------------------------
class ReflectionPad2d(nn.Module):
    def __init__(self, padding):
        super().__init__()
        self.padding = int(padding)

    def extra_repr(self):
        return str(self.padding)


class InstanceNorm2d(nn.Module):
    """affine=False, track_running_stats=False, eps=1e-5 (networks.py:22): no parameters, no buffers."""

    def __init__(self, num_features, eps=1e-5):
        super().__init__()
        self.num_features, self.eps = num_features, eps


class ReLU(nn.Module):
    def __init__(self, inplace=True):
        super().__init__()


class LeakyReLU(nn.Module):
    def __init__(self, negative_slope=0.2, inplace=True):
        super().__init__()
        self.negative_slope = negative_slope


class Tanh(nn.Module):
    pass


class _ConvBase(nn.Module):
    def _init_params(self, logical_shape, fan_in, perm):
        # same RNG calls as torch.nn.Conv2d.reset_parameters (kaiming_uniform_(a=sqrt(5)); uniform_ bias)
        w = torch.empty(logical_shape)
        nn.init.kaiming_uniform_(w, a=math.sqrt(5))
        bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
        b = torch.empty(self.out_channels).uniform_(-bound, bound)
        phys = w.permute(*perm).contiguous()                      # [KH][KW][Cout][Cin]
        inv = [perm.index(i) for i in range(4)]
        self.weight = nn.Parameter(phys.permute(*inv))            # logical view of the physical buffer
        self.bias = nn.Parameter(b)
        self.w16 = None                                           # bf16 shadow of the physical buffer
        self.gphys = None                                         # FusedAdam: this weight's slice of the flat grads
        self._perm = perm
        self.weight._p2p_conv = self                              # lets optim.FusedAdam find the physical layout

    def physical_weight(self):
        return self.weight.permute(*self._perm)

    def sync_shadow(self):
        phys = self.physical_weight().detach()
        if self.w16 is not None and self.w16.device == phys.device and self.w16.shape == phys.shape:
            self.w16.copy_(phys)                                  # keeps a view into FusedAdam's flat bf16 buffer valid
        else:
            self.w16 = phys.to(torch.bfloat16).contiguous()


class Conv2d(_ConvBase):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = int(kernel_size), int(stride), int(padding)
        k = self.kernel_size
        self.tswap = False
        self._init_params((out_channels, in_channels, k, k), in_channels * k * k, (2, 3, 0, 1))

    def set_transposed_image(self):
        """this layer will be run on spatially transposed images (MultiscaleDiscriminator(transposed_image=True)): store
        the weight [KW][KH][Cout][Cin] behind the same logical [Cout, Cin, KH, KW] parameter (state_dict unchanged).
        Call before the optimiser re-homes the parameter."""
        if self.tswap:
            return
        phys = self.weight.data.permute(3, 2, 0, 1).contiguous()
        self.weight = nn.Parameter(phys.permute(2, 3, 1, 0))
        self._perm = (3, 2, 0, 1)
        self.weight._p2p_conv = self
        self.w16, self.gphys, self.tswap = None, None, True

    def cfg_for(self, reflect_pad=0):
        if reflect_pad:
            assert self.padding == 0
            return ops.ConvCfg(False, self.stride, reflect_pad, ops.PAD_REFLECT, tswap=self.tswap)
        return ops.ConvCfg(False, self.stride, self.padding, ops.PAD_ZERO, tswap=self.tswap)

    def run(self, x, reflect_pad=0, needs_wgrad=True, norm=None, act_mod=None, res=None, consumer=None, want_plain=False, gated=False):
        """conv [+ InstanceNorm] [+ activation] [+ residual] as one fused op; `consumer` = (conv module, ConvCfg) of the layer
        that reads the result (its padded input planes are then written directly), want_plain: the result must be a plain
        contiguous NHWC tensor (it is also handed out as a feature map)"""
        return _conv_block(self, x, self.cfg_for(reflect_pad), needs_wgrad, norm, act_mod, res, consumer, want_plain, gated)

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, stride={self.stride}, padding={self.padding}"


class ConvTranspose2d(_ConvBase):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, output_padding=0):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride = int(kernel_size), int(stride)
        self.padding, self.output_padding = int(padding), int(output_padding)
        k = self.kernel_size
        # torch computes fan_in of a [Cin, Cout, k, k] transposed-conv weight as size(1) * k * k
        self._init_params((in_channels, out_channels, k, k), out_channels * k * k, (2, 3, 1, 0))

    def cfg_for(self, reflect_pad=0):
        return ops.ConvCfg(True, self.stride, self.padding, ops.PAD_ZERO, self.output_padding)

    def run(self, x, needs_wgrad=True, norm=None, act_mod=None, res=None, consumer=None, want_plain=False):
        return _conv_block(self, x, self.cfg_for(), needs_wgrad, norm, act_mod, res, consumer, want_plain)

    def extra_repr(self):
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, stride={self.stride}, "
                f"padding={self.padding}, output_padding={self.output_padding}")


def _act_code(act_mod):
    if isinstance(act_mod, ReLU):
        return ops.ACT_RELU, 0.0
    if isinstance(act_mod, LeakyReLU):
        return ops.ACT_LRELU, act_mod.negative_slope
    assert act_mod is None
    return ops.ACT_NONE, 0.0


def _conv_block(conv, x, cfg, needs_wgrad, norm, act_mod, res, consumer=None, want_plain=False, gated=False):
    w = conv.weight if needs_wgrad else conv.weight.detach()
    b = conv.bias if needs_wgrad else conv.bias.detach()
    if x.dtype == torch.bfloat16 and (conv.w16 is None or conv.w16.device != x.device):
        conv.sync_shadow()
    act, slope = _act_code(act_mod)
    out_plan = None
    if consumer is not None:
        cmod, ccfg = consumer
        out_plan = lambda shape, dtype: ops.plan_input_layout(shape, dtype, cmod.weight, ccfg)      # noqa: E731
    z, layout = ops.ConvBlockFn.apply(x, w, b, conv.w16, res, cfg, norm is not None, act, slope,
                                      norm.eps if norm is not None else 0.0, conv.gphys if needs_wgrad else None,
                                      getattr(x, "_p2p_layout", None), out_plan, want_plain, gated)
    if layout is not None:
        z._p2p_layout = layout          # (key, planes): this tensor already exists as its consumer's padded input
    return z


def _consumer_of(mods):
    """the convolution a module (list) starts with -> (conv module, ConvCfg), or None"""
    if isinstance(mods, nn.Sequential):
        mods = list(mods)
    elif not isinstance(mods, (list, tuple)):
        mods = [mods]
    if not mods:
        return None
    m = mods[0]
    if isinstance(m, nn.Sequential):
        return _consumer_of(list(m) + list(mods[1:]))
    if isinstance(m, ResnetBlock):
        return _consumer_of(list(m.conv_block))
    if isinstance(m, ReflectionPad2d):
        if len(mods) > 1 and isinstance(mods[1], Conv2d):
            return mods[1], mods[1].cfg_for(m.padding)
        return None
    if isinstance(m, (Conv2d, ConvTranspose2d)):
        return m, m.cfg_for()
    return None


def weights_init(m):
    """networks.py:10-16 (only the Conv branch is reachable: norm layers here have no parameters)."""
    classname = m.__class__.__name__
    if classname.find('Conv') != -1:
        w = m.weight.data
        fresh = torch.empty(w.shape, device=w.device, dtype=w.dtype).normal_(0.0, 0.02)   # contiguous draw = reference order
        w.copy_(fresh)


def get_norm_layer(norm_type='instance'):
    if norm_type == 'instance':
        return InstanceNorm2d
    raise NotImplementedError('normalization layer [%s] is not found' % norm_type)   # 'batch' unused by every recipe


# --------------------------------------------------------------------------------------------------
# executor: walks a reference-shaped module list and launches fused ops
# --------------------------------------------------------------------------------------------------
def _norm_act(x, norm, act_mod, res=None):
    act, slope = _act_code(act_mod)
    return ops.InstNormActFn.apply(x, res, norm is not None, act, slope, norm.eps if norm is not None else 0.0)


def _take_norm_act(mods, i):
    """modules following a conv at index i-1: (norm or None, activation or None, next index)"""
    norm = act = None
    if i < len(mods) and isinstance(mods[i], InstanceNorm2d):
        norm = mods[i]
        i += 1
    if i < len(mods) and isinstance(mods[i], (ReLU, LeakyReLU)):
        act = mods[i]
        i += 1
    return norm, act, i


def run_layers(mods, x, needs_wgrad=True, res_last=None, next_consumer=None, want_plain=False, gated=False):
    """Execute a reference-shaped module list.  Conv [+IN] [+act] runs as one fused op; `res_last` is added in
    the last conv block of the list (LocalEnhancer's `model_downsample(input_i) + output_prev`).  `next_consumer`: the
    module(s) that will read this list's result (lets the last block write its consumer's padded input directly);
    `want_plain`: the list's result is also used as a plain tensor (a discriminator feature map)."""
    mods = list(mods)
    i = 0
    while i < len(mods):
        m = mods[i]
        if isinstance(m, nn.Sequential):
            x = run_layers(m, x, needs_wgrad, next_consumer=mods[i + 1:] if i + 1 < len(mods) else next_consumer,
                           want_plain=want_plain and i + 1 == len(mods), gated=gated and i == 0)
            i += 1
        elif isinstance(m, (ReflectionPad2d, Conv2d, ConvTranspose2d)):
            pad, first = 0, i == 0
            if isinstance(m, ReflectionPad2d):
                pad, m = m.padding, mods[i + 1]
                assert isinstance(m, Conv2d), "ReflectionPad2d must be followed by Conv2d"
                i += 1
            norm, act, i = _take_norm_act(mods, i + 1)
            last = i == len(mods)
            res = res_last if last else None
            consumer = _consumer_of(mods[i:]) if not last else (_consumer_of(next_consumer) if next_consumer is not None else None)
            plain = want_plain and last
            if isinstance(m, Conv2d):
                x = m.run(x, reflect_pad=pad, needs_wgrad=needs_wgrad, norm=norm, act_mod=act, res=res, consumer=consumer,
                          want_plain=plain, gated=gated and first)
            else:
                x = m.run(x, needs_wgrad=needs_wgrad, norm=norm, act_mod=act, res=res, consumer=consumer, want_plain=plain)
            if res is not None:
                res_last = None
        elif isinstance(m, InstanceNorm2d):
            norm, act, i = _take_norm_act(mods, i)
            x = _norm_act(x, norm, act)
        elif isinstance(m, (ReLU, LeakyReLU)):
            x = _norm_act(x, None, m)
            i += 1
        elif isinstance(m, Tanh):
            x = ops.TanhFn.apply(x)
            i += 1
        elif isinstance(m, ResnetBlock):
            last = i + 1 == len(mods)
            consumer = _consumer_of(mods[i + 1:]) if not last else (_consumer_of(next_consumer) if next_consumer is not None else None)
            x = m.run(x, needs_wgrad, consumer)
            i += 1
        else:
            raise NotImplementedError(f"no kernel for module {type(m).__name__}")
    assert res_last is None, "res_last needs a trailing conv block"
    return x


def _to_nhwc(x, dtype):
    """logical NCHW (any strides) -> contiguous NHWC in the activation dtype"""
    return x.permute(0, 2, 3, 1).to(dtype).contiguous()


def _to_nchw(x):
    return x.permute(0, 3, 1, 2)


def sync_shadows(net):
    """refresh the bf16 weight shadows of every conv of `net` (after loading / updating fp32 weights)"""
    for m in net.modules():
        if isinstance(m, _ConvBase):
            m.sync_shadow()


# --------------------------------------------------------------------------------------------------
# generators (networks.py:129-253)
# --------------------------------------------------------------------------------------------------
class ResnetBlock(nn.Module):
    """networks.py:214-253: x + IN(Conv3x3(RefPad1(ReLU(IN(Conv3x3(RefPad1(x)))))))"""

    def __init__(self, dim, padding_type='reflect', norm_layer=InstanceNorm2d, activation=None, use_dropout=False):
        super().__init__()
        if padding_type != 'reflect':
            raise NotImplementedError('padding [%s] is not implemented' % padding_type)
        if use_dropout:
            raise NotImplementedError('dropout is not used by any recipe')
        self.conv_block = nn.Sequential(ReflectionPad2d(1), Conv2d(dim, dim, kernel_size=3, padding=0), norm_layer(dim),
                                        ReLU(True), ReflectionPad2d(1), Conv2d(dim, dim, kernel_size=3, padding=0),
                                        norm_layer(dim))

    def run(self, x, needs_wgrad=True, consumer=None):
        cb = self.conv_block
        h = cb[1].run(x, reflect_pad=cb[0].padding, needs_wgrad=needs_wgrad, norm=cb[2], act_mod=cb[3],
                      consumer=(cb[5], cb[5].cfg_for(cb[4].padding)))
        return cb[5].run(h, reflect_pad=cb[4].padding, needs_wgrad=needs_wgrad, norm=cb[6], act_mod=None, res=x, consumer=consumer)

    def forward(self, x):
        return _to_nchw(self.run(_to_nhwc(x, _act_dtype(self))))


def _act_dtype(m):
    return getattr(m, 'act_dtype', _DEFAULT_DTYPE)


class GlobalGenerator(nn.Module):
    """networks.py:183-211"""

    def __init__(self, input_nc, output_nc, ngf=64, n_downsampling=3, n_blocks=9, norm_layer=InstanceNorm2d,
                 padding_type='reflect'):
        assert (n_blocks >= 0)
        super().__init__()
        model = [ReflectionPad2d(3), Conv2d(input_nc, ngf, kernel_size=7, padding=0), norm_layer(ngf), ReLU(True)]
        for i in range(n_downsampling):
            mult = 2 ** i
            model += [Conv2d(ngf * mult, ngf * mult * 2, kernel_size=3, stride=2, padding=1),
                      norm_layer(ngf * mult * 2), ReLU(True)]
        mult = 2 ** n_downsampling
        for i in range(n_blocks):
            model += [ResnetBlock(ngf * mult, padding_type=padding_type, norm_layer=norm_layer)]
        for i in range(n_downsampling):
            mult = 2 ** (n_downsampling - i)
            model += [ConvTranspose2d(ngf * mult, int(ngf * mult / 2), kernel_size=3, stride=2, padding=1, output_padding=1),
                      norm_layer(int(ngf * mult / 2)), ReLU(True)]
        model += [ReflectionPad2d(3), Conv2d(ngf, output_nc, kernel_size=7, padding=0), Tanh()]
        self.model = nn.Sequential(*model)
        self.act_dtype = _DEFAULT_DTYPE

    def forward(self, input):
        return _to_nchw(run_layers(self.model, _to_nhwc(input, self.act_dtype)))


class LocalEnhancer(nn.Module):
    """networks.py:129-181"""

    def __init__(self, input_nc, output_nc, ngf=32, n_downsample_global=3, n_blocks_global=9, n_local_enhancers=1,
                 n_blocks_local=3, norm_layer=InstanceNorm2d, padding_type='reflect'):
        super().__init__()
        self.n_local_enhancers = n_local_enhancers
        ngf_global = ngf * (2 ** n_local_enhancers)
        model_global = GlobalGenerator(input_nc, output_nc, ngf_global, n_downsample_global, n_blocks_global, norm_layer).model
        model_global = [model_global[i] for i in range(len(model_global) - 3)]   # drop the final convolution layers
        self.model = nn.Sequential(*model_global)
        for n in range(1, n_local_enhancers + 1):
            ngf_global = ngf * (2 ** (n_local_enhancers - n))
            model_downsample = [ReflectionPad2d(3), Conv2d(input_nc, ngf_global, kernel_size=7, padding=0),
                                norm_layer(ngf_global), ReLU(True),
                                Conv2d(ngf_global, ngf_global * 2, kernel_size=3, stride=2, padding=1),
                                norm_layer(ngf_global * 2), ReLU(True)]
            model_upsample = []
            for i in range(n_blocks_local):
                model_upsample += [ResnetBlock(ngf_global * 2, padding_type=padding_type, norm_layer=norm_layer)]
            model_upsample += [ConvTranspose2d(ngf_global * 2, ngf_global, kernel_size=3, stride=2, padding=1, output_padding=1),
                               norm_layer(ngf_global), ReLU(True)]
            if n == n_local_enhancers:
                model_upsample += [ReflectionPad2d(3), Conv2d(ngf, output_nc, kernel_size=7, padding=0), Tanh()]
            setattr(self, 'model' + str(n) + '_1', nn.Sequential(*model_downsample))
            setattr(self, 'model' + str(n) + '_2', nn.Sequential(*model_upsample))
        self.act_dtype = _DEFAULT_DTYPE

    def forward(self, input):
        x = _to_nhwc(input, self.act_dtype)
        input_downsampled = [x]
        for i in range(self.n_local_enhancers):
            input_downsampled.append(ops.AvgPool3s2Fn.apply(input_downsampled[-1]))
        output_prev = run_layers(self.model, input_downsampled[-1])
        for n in range(1, self.n_local_enhancers + 1):
            model_downsample = getattr(self, 'model' + str(n) + '_1')
            model_upsample = getattr(self, 'model' + str(n) + '_2')
            input_i = input_downsampled[self.n_local_enhancers - n]
            # model_downsample ends with IN + ReLU: fuse `+ output_prev` into that pass (networks.py:180)
            h = run_layers(model_downsample, input_i, res_last=output_prev, next_consumer=model_upsample)
            output_prev = run_layers(model_upsample, h)
        return _to_nchw(output_prev)


def define_G(input_nc, output_nc, ngf, netG, n_downsample_global=3, n_blocks_global=9, n_local_enhancers=1,
             n_blocks_local=3, norm='instance', gpu_ids=[]):
    """networks.py:27-44"""
    norm_layer = get_norm_layer(norm_type=norm)
    if netG == 'global':
        netG = GlobalGenerator(input_nc, output_nc, ngf, n_downsample_global, n_blocks_global, norm_layer)
    elif netG == 'local':
        netG = LocalEnhancer(input_nc, output_nc, ngf, n_downsample_global, n_blocks_global,
                             n_local_enhancers, n_blocks_local, norm_layer)
    else:
        raise NotImplementedError('generator [%s] not implemented (the encoder / instance-feature path is out of scope)' % netG)
    if len(gpu_ids) > 0:
        assert (torch.cuda.is_available())
        netG.cuda(gpu_ids[0])
    netG.apply(weights_init)
    sync_shadows(netG)
    return netG


# --------------------------------------------------------------------------------------------------
# discriminators (networks.py:292-383) and GANLoss (networks.py:68-110)
# --------------------------------------------------------------------------------------------------
class NLayerDiscriminator(nn.Module):
    """networks.py:334-383: 4x4 convs, stride 2 (x n_layers) then stride 1 (x2), padding 2, LeakyReLU(0.2), IN."""

    def __init__(self, input_nc, ndf=64, n_layers=3, norm_layer=InstanceNorm2d, use_sigmoid=False, getIntermFeat=False):
        super().__init__()
        if use_sigmoid:
            raise NotImplementedError('use_sigmoid (vanilla GAN) is not used by any recipe; LSGAN only')
        self.getIntermFeat = getIntermFeat
        self.n_layers = n_layers
        kw = 4
        padw = int(math.ceil((kw - 1.0) / 2))
        sequence = [[Conv2d(input_nc, ndf, kernel_size=kw, stride=2, padding=padw), LeakyReLU(0.2, True)]]
        nf = ndf
        for n in range(1, n_layers):
            nf_prev = nf
            nf = min(nf * 2, 512)
            sequence += [[Conv2d(nf_prev, nf, kernel_size=kw, stride=2, padding=padw), norm_layer(nf), LeakyReLU(0.2, True)]]
        nf_prev = nf
        nf = min(nf * 2, 512)
        sequence += [[Conv2d(nf_prev, nf, kernel_size=kw, stride=1, padding=padw), norm_layer(nf), LeakyReLU(0.2, True)]]
        sequence += [[Conv2d(nf, 1, kernel_size=kw, stride=1, padding=padw)]]
        if getIntermFeat:
            for n in range(len(sequence)):
                setattr(self, 'model' + str(n), nn.Sequential(*sequence[n]))
        else:
            sequence_stream = []
            for n in range(len(sequence)):
                sequence_stream += sequence[n]
            self.model = nn.Sequential(*sequence_stream)
        self.act_dtype = _DEFAULT_DTYPE

    def forward(self, input):
        x = _to_nhwc(input, self.act_dtype)
        if self.getIntermFeat:
            res = []
            for n in range(self.n_layers + 2):
                nxt = getattr(self, 'model' + str(n + 1)) if n < self.n_layers + 1 else None
                x = run_layers(getattr(self, 'model' + str(n)), x, next_consumer=nxt, want_plain=True)
                res.append(_to_nchw(x))
            return res
        return _to_nchw(run_layers(self.model, x))


class MultiscaleDiscriminator(nn.Module):
    """networks.py:292-331: num_D PatchGANs on an average-pooled pyramid; finest scale first in the result."""

    def __init__(self, input_nc, ndf=64, n_layers=3, norm_layer=InstanceNorm2d, use_sigmoid=False, num_D=3,
                 getIntermFeat=False, transposed_image=False):
        super().__init__()
        # transposed_image: run every layer on the spatially transposed image (exact: square kernels, symmetric stride,
        # padding and pooling).  For wide images (the time-domain D's [frames, n_fft] maps) this puts the short axis on
        # the fast dimension, which the shifted-window conv kernels need (halo = (K-1) * padded width rows).
        self.transposed_image = transposed_image
        self.num_D = num_D
        self.n_layers = n_layers
        self.getIntermFeat = getIntermFeat
        for i in range(num_D):
            netD = NLayerDiscriminator(input_nc, ndf, n_layers, norm_layer, use_sigmoid, getIntermFeat)
            if getIntermFeat:
                for j in range(n_layers + 2):
                    setattr(self, 'scale' + str(i) + '_layer' + str(j), getattr(netD, 'model' + str(j)))
            else:
                setattr(self, 'layer' + str(i), netD.model)
        self.act_dtype = _DEFAULT_DTYPE
        if transposed_image:
            for m in self.modules():
                if isinstance(m, Conv2d):
                    m.set_transposed_image()

    def _out(self, x):
        y = _to_nchw(x)
        return y.transpose(2, 3) if self.transposed_image else y

    def singleD_forward(self, model, x, needs_wgrad=True, gated=False):
        if self.getIntermFeat:
            result = []
            for i in range(len(model)):
                x = run_layers(model[i], x, needs_wgrad, next_consumer=model[i + 1] if i + 1 < len(model) else None, want_plain=True,
                               gated=gated and i == 0)
                result.append(self._out(x))
            return result
        return [self._out(run_layers(model, x, needs_wgrad, gated=gated))]

    def forward(self, input, needs_wgrad=True, gated=False):
        """needs_wgrad=False runs with detached weights: the generator's pass through D, whose weight gradients
        the reference computes and then discards (train.py:164-176)."""
        num_D = self.num_D
        result = []
        x = _to_nhwc(input.transpose(2, 3) if self.transposed_image else input, self.act_dtype)
        for i in range(num_D):
            if self.getIntermFeat:
                model = [getattr(self, 'scale' + str(num_D - 1 - i) + '_layer' + str(j)) for j in range(self.n_layers + 2)]
            else:
                model = getattr(self, 'layer' + str(num_D - 1 - i))
            result.append(self.singleD_forward(model, x, needs_wgrad, gated))
            if i != (num_D - 1):
                x = ops.AvgPool3s2Fn.apply(x, gated)
        return result


def define_D(input_nc, ndf, n_layers_D, norm='instance', use_sigmoid=False, num_D=1, getIntermFeat=False, gpu_ids=[],
             transposed_image=False):
    """networks.py:46-54 (transposed_image: see MultiscaleDiscriminator; results and state_dict are unchanged)"""
    norm_layer = get_norm_layer(norm_type=norm)
    netD = MultiscaleDiscriminator(input_nc, ndf, n_layers_D, norm_layer, use_sigmoid, num_D, getIntermFeat, transposed_image)
    if len(gpu_ids) > 0:
        assert (torch.cuda.is_available())
        netD.cuda(gpu_ids[0])
    netD.apply(weights_init)
    sync_shadows(netD)
    return netD


def _nhwc_view(t):
    """NCHW-shaped tensor -> NHWC contiguous tensor (free for the channels-last views our nets return)"""
    return t.permute(0, 2, 3, 1).contiguous()


class GANLoss(nn.Module):
    """networks.py:68-110, LSGAN branch: sum over scales of MSE(pred[-1], real_label or fake_label)."""

    def __init__(self, use_lsgan=True, target_real_label=1.0, target_fake_label=0.0, tensor=torch.FloatTensor):
        super().__init__()
        if not use_lsgan:
            raise NotImplementedError('BCE GAN loss (--no_lsgan) is not used by any recipe')
        self.real_label = target_real_label
        self.fake_label = target_fake_label

    def _one(self, pred, target_is_real):
        return ops.MseConstFn.apply(_nhwc_view(pred), self.real_label if target_is_real else self.fake_label, 1.0)

    def __call__(self, input, target_is_real):
        if isinstance(input[0], list):
            # one scalar for all scales: the loss kernels accumulate into it (no per-scale zero-fill / add kernels)
            return ops.MseConstSumFn.apply(self.real_label if target_is_real else self.fake_label, 1.0,
                                           *[_nhwc_view(input_i[-1]) for input_i in input])
        return self._one(input[-1], target_is_real)


def feature_matching_loss(pred_fake, pred_real, num_D, n_layers_D, lambda_feat):
    """pix2pixHD_model.py:390-398: sum_i sum_{j<len-1} D_weights * feat_weights * L1(fake_ij, real_ij.detach()) * lambda_feat"""
    feat_weights = 4.0 / (n_layers_D + 1)
    D_weights = 1.0 / num_D
    fakes = [_nhwc_view(pred_fake[i][j]) for i in range(num_D) for j in range(len(pred_fake[i]) - 1)]
    reals = [_nhwc_view(pred_real[i][j]).detach() for i in range(num_D) for j in range(len(pred_fake[i]) - 1)]
    if not fakes:
        return 0
    return ops.L1SumFn.apply(D_weights * feat_weights * lambda_feat, *fakes, *reals)     # one accumulator for every term
