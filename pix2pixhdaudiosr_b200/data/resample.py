"""Device-side data path (reference data/audio_dataset.py:41-90): the three `torchaudio.functional.resample`
calls per item and the crop / zero-pad to `segment_length`, on the GPU through `p2p_resample` (csrc/resample.cu).

`resample(waveform, orig_freq, new_freq, lowpass_filter_width=6, rolloff=0.99)` follows torchaudio 2.11's
`resample(..., resampling_method='sinc_interp_hann')` step by step (functional.py: _get_sinc_resample_kernel /
_apply_sinc_resample_kernel): gcd reduction, Hann-windowed sinc filter bank of `new` phases x (2*width + orig) taps
(built here in float64), strided correlation, cut to ceil(new * T / orig).  No CPU fallback."""
import ctypes
import math

import torch

from .. import _lib

_BANKS = {}


def sinc_filter_bank(orig, new, lowpass_filter_width=6, rolloff=0.99):
    """[new, 2*width+orig] float64 filter bank and `width`, for gcd-reduced rates"""
    base_freq = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base_freq)
    idx = torch.arange(-width, width + orig, dtype=torch.float64)[None, :] / orig
    t = torch.arange(0, -new, -1, dtype=torch.float64)[:, None] / new + idx
    t = (t * base_freq).clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base_freq / orig
    bank = torch.where(t == 0, torch.ones_like(t), t.sin() / t) * window * scale
    return bank, width


def _bank_on(device, orig, new, lowpass_filter_width, rolloff):
    key = (device.type, device.index, orig, new, lowpass_filter_width, rolloff)
    if key not in _BANKS:
        bank, width = sinc_filter_bank(orig, new, lowpass_filter_width, rolloff)
        _BANKS[key] = (bank.to(torch.float32).contiguous().to(device), width)
    return _BANKS[key]


def resample(waveform, orig_freq, new_freq, lowpass_filter_width=6, rolloff=0.99, resampling_method='sinc_interp_hann'):
    if orig_freq <= 0 or new_freq <= 0:
        raise ValueError('Original frequency and desired frequecy should be positive')      # torchaudio's message
    if resampling_method != 'sinc_interp_hann':
        raise NotImplementedError("only resampling_method='sinc_interp_hann' (torchaudio's default, what the dataset uses)")
    if not waveform.is_cuda:
        raise RuntimeError('resample: the B200 path needs a CUDA tensor, got %s (no CPU fallback)' % waveform.device)
    if orig_freq == new_freq:
        return waveform
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    shape = waveform.shape
    x = waveform.to(torch.float32).reshape(-1, shape[-1]).contiguous()
    B, T = x.shape
    T_out = int(math.ceil(new * T / orig))
    bank, width = _bank_on(x.device, orig, new, lowpass_filter_width, rolloff)
    out = torch.empty((B, T_out), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().p2p_resample(x.data_ptr(), bank.data_ptr(), out.data_ptr(), B, T, T_out, orig, new, width,
                                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out.reshape(*shape[:-1], T_out)


def seg_pad_audio(waveform, segment_length):
    """audio_dataset.py:82-90 for a batch [B, T]: crop to segment_length or zero-pad at the end"""
    T = waveform.shape[-1]
    if T >= segment_length:
        return waveform[..., :segment_length]
    return torch.nn.functional.pad(waveform, (0, segment_length - T))


def make_lr_hr(waveform, orig_sample_rate, lr_sampling_rate, hr_sampling_rate, segment_length):
    """audio_dataset.py:56-62 for a batch of raw segments already on the device: returns {'image': hr, 'label': lr}"""
    hr = resample(waveform, orig_sample_rate, hr_sampling_rate)
    lr = resample(waveform, orig_sample_rate, lr_sampling_rate)
    lr = resample(lr, lr_sampling_rate, hr_sampling_rate)
    return {'image': seg_pad_audio(hr, segment_length), 'label': seg_pad_audio(lr, segment_length), 'inst': 0, 'feat': 0}
