"""Training driver bookkeeping: iter.txt / loss_log.txt in the reference's formats (train.py:26-40, 202-204,
util/visualizer.py:102-110), checkpoints that resume -- including the optimiser state the reference does not save."""
import os

import numpy as np
import pytest
import torch

from pix2pixhdaudiosr_b200 import train_loop as TL


def test_error_line_has_the_visualizer_format():
    line = TL.format_errors(3, 128, {'G_GAN': 0.81234, 'G_GAN_Feat': 5.5, 'G_VGG': 0, 'D_real': 0.25, 'D_fake': 0.3}, 0.0123)
    assert line == '(epoch: 3, iters: 128, time: 0.012) G_GAN: 0.812 G_GAN_Feat: 5.500 D_real: 0.250 D_fake: 0.300 '


def test_iter_file_round_trip(tmp_path):
    p = str(tmp_path / 'iter.txt')
    assert TL.read_iter(p, continue_train=False) == (1, 0)
    assert TL.read_iter(p, continue_train=True) == (1, 0)            # missing file: train.py:36-37
    np.savetxt(p, (7, 4096), delimiter=',', fmt='%d')                # what train.py:204 writes
    assert open(p).read().split() == ['7', '4096']
    assert TL.read_iter(p, continue_train=True) == (7, 4096)


def test_synthetic_dataset_is_deterministic():
    a = list(TL.synthetic_dataset(2, 3, 100, seed=5))
    b = list(TL.synthetic_dataset(2, 3, 100, seed=5))
    assert len(a) == 2 and a[0]['image'].shape == (3, 100)
    assert all(torch.equal(x['image'], y['image']) and torch.equal(x['label'], y['label']) for x, y in zip(a, b))
    assert not torch.equal(a[0]['image'], a[1]['image'])


@pytest.mark.gpu
def test_checkpoint_resume_continues_the_same_trajectory(tmp_path):
    """3 steps, save, rebuild the model with --continue_train: weights, Adam moments / step counter, iter.txt and
    loss_log.txt come back, and the next step of the resumed run equals the next step of the uninterrupted run"""
    from pix2pixhdaudiosr_b200.models.models import create_model
    from pix2pixhdaudiosr_b200.train import train_step
    from test_model_gpu import batch, tiny_opt
    kw = dict(checkpoints_dir=str(tmp_path), name='resume', use_time_D=True, use_match_loss=True, lambda_time=10.0)
    opt = tiny_opt(**kw)
    torch.manual_seed(opt.seed)
    model = create_model(opt)
    data = lambda epoch: ({'label': batch(i)[0], 'image': batch(i)[1]} for i in range(4))
    lines = []
    n = TL.run(opt, model, data, dataset_size=8, max_steps=3, print_freq=2, save_latest_freq=4, log=lines.append)
    assert n == 3
    save_dir = os.path.join(str(tmp_path), 'resume')
    for f in ('latest_net_G.pth', 'latest_net_D.pth', 'latest_net_time_D.pth', 'latest_optim.pth', 'iter.txt', 'loss_log.txt'):
        assert os.path.isfile(os.path.join(save_dir, f)), f
    assert TL.read_iter(os.path.join(save_dir, 'iter.txt'), True) == (1, 6)
    log = open(os.path.join(save_dir, 'loss_log.txt')).read().splitlines()
    assert log[0].startswith('================ Training Loss')
    assert any(l.startswith('(epoch: 1, iters: ') and 'G_GAN: ' in l and 'D_fake_t: ' in l for l in log[1:])
    sd = torch.load(os.path.join(save_dir, 'latest_net_G.pth'))
    assert all(torch.equal(v, model.netG.state_dict()[k].cpu()) for k, v in sd.items())       # reference-format state_dict

    resumed = create_model(tiny_opt(continue_train=True, **kw))
    assert TL.load_optimizers(resumed)
    for a, b in ((model.optimizer_G, resumed.optimizer_G), (model.optimizer_D, resumed.optimizer_D)):
        assert int(a.step_dev.item()) == int(b.step_dev.item()) == 3
        assert torch.equal(a.flat, b.flat) and torch.equal(a.exp_avg, b.exp_avg) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)
    lr, hr = batch(3)
    la, _ = train_step(model, lr, hr)
    lb, _ = train_step(resumed, lr, hr)
    for k in la:        # atomics reorder fp32 sums; G_mat is a 1e-5-sized difference of nearly equal numbers
        tol = 5e-2 if k == 'G_mat' else 1e-3
        assert abs(float(la[k]) - float(lb[k])) <= tol * max(abs(float(la[k])), 1e-6), k
    wa, wb = model.optimizer_G.flat, resumed.optimizer_G.flat
    assert float((wa - wb).abs().max()) <= 2e-4 * 2 + 1e-7
