"""Device-side resampling: the filter bank against torchaudio's own on CPU, the kernel against
torchaudio.functional.resample (the reference's call, data/audio_dataset.py:56-58) on the GPU."""
import math

import pytest
import torch
import torchaudio.functional as aF

from cases import relerr
from pix2pixhdaudiosr_b200.data import resample as R

RATES = [(48000, 8000), (8000, 48000), (48000, 16000), (16000, 48000), (44100, 48000), (48000, 12000), (22050, 16000)]


@pytest.mark.parametrize("orig,new", RATES)
def test_filter_bank_reproduces_torchaudio_on_cpu(orig, new):
    """correlating with our float64 bank = torchaudio.functional.resample.  torchaudio builds its bank in the waveform's
    dtype (float32: the phase argument idx/orig + j/new carries ~1e-7 * width of rounding), so agreement is 1e-5 for the
    160-phase 44.1 -> 48 kHz bank and 1e-6 for the small ones"""
    g = math.gcd(orig, new)
    o, n = orig // g, new // g
    bank, width = R.sinc_filter_bank(o, n)
    torch.manual_seed(orig + new)
    x = torch.randn(2, 4000)
    want = aF.resample(x, orig, new)
    xp = torch.nn.functional.pad(x, (width, width + o)).double()
    y = torch.nn.functional.conv1d(xp[:, None], bank[:, None], stride=o).transpose(1, 2).reshape(2, -1)
    y = y[:, : math.ceil(n * x.shape[1] / o)]
    assert y.shape == want.shape
    assert relerr(y.numpy(), want.numpy()) < (2e-5 if max(o, n) > 32 else 2e-6)


def test_argument_errors():
    with pytest.raises(ValueError):
        R.resample(torch.zeros(1, 10), 0, 8000)
    with pytest.raises(RuntimeError):
        R.resample(torch.zeros(1, 10), 48000, 8000)           # CPU tensor: no fallback
    with pytest.raises(NotImplementedError):
        R.resample(torch.zeros(1, 10), 48000, 8000, resampling_method='sinc_interp_kaiser')


@pytest.mark.gpu
@pytest.mark.parametrize("orig,new", RATES)
@pytest.mark.parametrize("T", [1, 37, 4001, 65024])
def test_kernel_matches_torchaudio(orig, new, T):
    torch.manual_seed(T)
    x = 0.1 * torch.randn(3, T)
    want = aF.resample(x, orig, new)
    got = R.resample(x.cuda(), orig, new)
    assert tuple(got.shape) == tuple(want.shape)
    assert relerr(got.cpu().numpy(), want.numpy()) < 3e-5      # see test_filter_bank_reproduces_torchaudio_on_cpu


@pytest.mark.gpu
def test_dataset_item_matches_the_reference_pipeline():
    """audio_dataset.py:56-62 on a batch: HR = resample to 48 kHz, LR = down to 8 kHz and back up, cropped / padded"""
    torch.manual_seed(9)
    wav = 0.1 * torch.randn(4, 60000)                        # 44.1 kHz source, shorter than the segment after resampling
    seg = 65024
    item = R.make_lr_hr(wav.cuda(), 44100, 8000, 48000, seg)
    hr = aF.resample(wav, 44100, 48000)
    lr = aF.resample(aF.resample(wav, 44100, 8000), 8000, 48000)
    pad = lambda w: w[:, :seg] if w.shape[1] >= seg else torch.nn.functional.pad(w, (0, seg - w.shape[1]))
    assert tuple(item['image'].shape) == tuple(item['label'].shape) == (4, seg)
    assert relerr(item['image'].cpu().numpy(), pad(hr).numpy()) < 3e-5
    assert relerr(item['label'].cpu().numpy(), pad(lr).numpy()) < 3e-5
