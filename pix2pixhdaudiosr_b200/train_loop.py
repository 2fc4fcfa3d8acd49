"""Minimal training driver around `train_step` with the reference's bookkeeping (reference train.py:26-35,
128-260, util/visualizer.py:102-110, models/base_model.py:36-89):

  * `<checkpoints_dir>/<name>/iter.txt` holds `epoch,epoch_iter` (np.savetxt, fmt %d) and is read back with
    `--continue_train` exactly as train.py:33-38 does;
  * `<checkpoints_dir>/<name>/loss_log.txt` gets the visualizer's line `(epoch: %d, iters: %d, time: %.3f) name: %.3f ...`
    (zero losses skipped) every `print_freq` samples;
  * `latest_net_G.pth` / `latest_net_D.pth` [/ `latest_net_time_D.pth`] every `save_latest_freq` samples and at the end of
    every `save_epoch_freq`-th epoch together with `<epoch>_net_*.pth` -- plain reference-format state_dicts, so files
    move both ways (load_network's strict -> subset -> size-matched fallbacks are kept);
  * additionally (absent in the reference, which restarts Adam from zero moments on resume): `latest_optim.pth` with both
    optimisers' step counter and moments in the LOGICAL parameter layout, restored on `--continue_train`;
  * learning-rate decay after `niter` epochs (`update_learning_rate`) and `niter_fix_global` (`update_fixed_params`).

Data: any iterable of dicts with 'label' (LR audio [B,T]) and 'image' (HR audio [B,T]) per step -- the keys the
reference's data loader yields (train.py:150); `synthetic_dataset` stands in for it (no dataset ships with this repo).
Multi-GPU: launch with torchrun, one process per GPU; rank 0 writes the files."""
import os
import time

import numpy as np
import torch
import torch.distributed as dist

from .train import GradSync, broadcast_parameters, train_step


def synthetic_dataset(n_batches, batch, segment_length, seed=0):
    """deterministic stand-in for data/audio_dataset.py: HR = noise, LR = a crude low-pass of it"""
    for i in range(n_batches):
        g = torch.Generator().manual_seed(seed * 1_000_003 + i)
        hr = 0.1 * torch.randn(batch, segment_length, generator=g)
        yield {'label': 0.5 * (hr + torch.roll(hr, 1, dims=1)), 'image': hr, 'inst': 0, 'feat': 0}


def read_iter(iter_path, continue_train):
    """train.py:33-40"""
    if continue_train:
        try:
            start_epoch, epoch_iter = np.loadtxt(iter_path, delimiter=',', dtype=int)
        except Exception:
            start_epoch, epoch_iter = 1, 0
        print('Resuming from epoch %d at iteration %d' % (start_epoch, epoch_iter))
        return int(start_epoch), int(epoch_iter)
    return 1, 0


def format_errors(epoch, i, errors, t):
    """util/visualizer.py:103-107"""
    message = '(epoch: %d, iters: %d, time: %.3f) ' % (epoch, i, t)
    for k, v in errors.items():
        if v != 0:
            message += '%s: %.3f ' % (k, v)
    return message


def _param_names(model, opt_):
    """'G.model.1.weight'-style names of the optimiser's parameters, in its order (None when a parameter is not found)"""
    ids = {}
    for tag in ('netG', 'netD', 'time_D'):
        net = getattr(model, tag, None)
        if net is not None:
            for k, p in net.named_parameters():
                ids[id(p)] = '%s.%s' % (tag, k)
    return [ids.get(id(p)) for p, _ in opt_._views]


def _optim_state(opt_, names=None):
    """step + moments per parameter, in the logical (reference) layout and order of param_groups[0]['params']; `names` (one per
    parameter) lets a later load match the moments by name instead of by position"""
    out = {'step': int(opt_.step_dev.item()), 'lr': float(opt_.param_groups[0]['lr']), 'exp_avg': [], 'exp_avg_sq': [],
           'names': list(names) if names is not None else None}
    for p, g in opt_._views:
        o, n = g.storage_offset(), p.numel()
        for name, flat in (('exp_avg', opt_.exp_avg), ('exp_avg_sq', opt_.exp_avg_sq)):
            v = torch.as_strided(flat, g.size(), g.stride(), o)          # same logical view as the gradient's
            out[name].append(v.detach().cpu().contiguous())
        assert n == out['exp_avg'][-1].numel()
    return out


def _load_optim_state(opt_, sd, names=None):
    """moments are matched BY NAME and shape when both sides carry names (a checkpoint written after update_fixed_params holds
    all of netG's parameters, a rebuilt optimizer_G with niter_fix_global > 0 only the local enhancer's): what is missing is
    left at zero with a warning; without names the lists must agree in length and every shape must match"""
    opt_.step_count = int(sd['step'])
    opt_.step_dev.fill_(opt_.step_count)
    opt_.param_groups[0]['lr'] = float(sd['lr'])
    opt_.sync_lr()
    saved_names = sd.get('names')
    if names is not None and saved_names is not None and None not in names and None not in saved_names:
        by_name = {k: (m, v) for k, m, v in zip(saved_names, sd['exp_avg'], sd['exp_avg_sq'])}
        pairs = [by_name.get(k) for k in names]
        missing = [k for k, pr in zip(names, pairs) if pr is None]
        if missing:
            print('optimizer state: no saved moments for %d parameter(s) (e.g. %s): they restart from zero' % (len(missing), missing[0]))
    else:
        if len(sd['exp_avg']) != len(opt_._views):
            raise ValueError('optimizer state holds %d parameters, the optimizer %d (and no names to match them by)'
                             % (len(sd['exp_avg']), len(opt_._views)))
        pairs = list(zip(sd['exp_avg'], sd['exp_avg_sq']))
    for (p, g), pr in zip(opt_._views, pairs):
        if pr is None:
            continue
        m, v = pr
        if tuple(m.shape) != tuple(g.size()):
            raise ValueError('optimizer state: moment of shape %s for a parameter of shape %s' % (tuple(m.shape), tuple(g.size())))
        o = g.storage_offset()
        torch.as_strided(opt_.exp_avg, g.size(), g.stride(), o).copy_(m)
        torch.as_strided(opt_.exp_avg_sq, g.size(), g.stride(), o).copy_(v)


def save_optimizers(model, label='latest'):
    os.makedirs(model.save_dir, exist_ok=True)
    torch.save({'G': _optim_state(model.optimizer_G, _param_names(model, model.optimizer_G)),
                'D': _optim_state(model.optimizer_D, _param_names(model, model.optimizer_D)), 'old_lr': model.old_lr},
               os.path.join(model.save_dir, '%s_optim.pth' % label))


def load_optimizers(model, label='latest'):
    path = os.path.join(model.save_dir, '%s_optim.pth' % label)
    if not os.path.isfile(path):
        print('%s not exists yet! (Adam restarts from zero moments, as the reference does)' % path)
        return False
    sd = torch.load(path, map_location='cpu')
    _load_optim_state(model.optimizer_G, sd['G'], _param_names(model, model.optimizer_G))
    _load_optim_state(model.optimizer_D, sd['D'], _param_names(model, model.optimizer_D))
    model.old_lr = sd['old_lr']
    return True


def run(opt, model, make_dataset, dataset_size, max_steps=None, print_freq=100, save_latest_freq=500, save_epoch_freq=10,
        log=print):
    """the loop of train.py:128-260.  make_dataset(epoch) -> iterable of batches; dataset_size in samples.
    Returns the number of optimisation steps taken."""
    rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    save_dir = os.path.join(opt.checkpoints_dir, opt.name)
    iter_path = os.path.join(save_dir, 'iter.txt')
    log_name = os.path.join(save_dir, 'loss_log.txt')
    start_epoch, epoch_iter = read_iter(iter_path, opt.continue_train)
    if opt.continue_train:
        load_optimizers(model)
    # bf16 training shards the optimiser step over the ranks (train.GradSync); the exact fp32 path reads the fp32 master weights
    # in its kernels, so it keeps the plain all-reduce
    sync = GradSync(shard=model.act_dtype == torch.bfloat16)
    broadcast_parameters(model)
    if rank == 0:
        os.makedirs(save_dir, exist_ok=True)
        with open(log_name, 'a') as f:
            f.write('================ Training Loss (%s) ================\n' % time.strftime('%c'))
    total_steps = (start_epoch - 1) * dataset_size + epoch_iter
    print_delta, save_delta = total_steps % print_freq, total_steps % save_latest_freq
    steps, iter_start_time = 0, time.time()

    def save(epoch, epoch_iter_next, also_epoch=False):
        sync.gather_state(model)          # sharded mode: a collective every rank takes part in (complete masters / moments)
        if rank != 0:
            return
        model.save('latest')
        save_optimizers(model, 'latest')
        if also_epoch:
            model.save(epoch)
        np.savetxt(iter_path, epoch_iter_next, delimiter=',', fmt='%d')

    for epoch in range(start_epoch, opt.niter + opt.niter_decay + 1):
        epoch_start_time = time.time()
        if epoch != start_epoch:
            epoch_iter = epoch_iter % dataset_size
        for data in make_dataset(epoch):
            if total_steps % print_freq == print_delta:
                iter_start_time = time.time()
            total_steps += opt.batchSize
            epoch_iter += opt.batchSize
            loss_dict, _ = train_step(model, data['label'], data['image'], sync=sync)
            steps += 1
            if total_steps % print_freq == print_delta:
                errors = {k: (float(v) if torch.is_tensor(v) else v) for k, v in loss_dict.items()}     # the step's only host sync
                message = format_errors(epoch, epoch_iter, errors, (time.time() - iter_start_time) / print_freq)
                if rank == 0:
                    log(message)
                    with open(log_name, 'a') as f:
                        f.write('%s\n' % message)
            if total_steps % save_latest_freq == save_delta:
                log('saving the latest model (epoch %d, total_steps %d)' % (epoch, total_steps))
                save(epoch, (epoch, epoch_iter))
            if epoch_iter >= dataset_size or (max_steps is not None and steps >= max_steps):
                break
        if max_steps is not None and steps >= max_steps:
            save(epoch, (epoch, epoch_iter))
            return steps
        log('End of epoch %d / %d \t Time Taken: %d sec' % (epoch, opt.niter + opt.niter_decay, time.time() - epoch_start_time))
        if epoch % save_epoch_freq == 0:
            log('saving the model at the end of epoch %d, iters %d' % (epoch, total_steps))
            save(epoch, (epoch + 1, 0), also_epoch=True)
        if opt.niter_fix_global != 0 and epoch == opt.niter_fix_global:
            model.update_fixed_params()
        if epoch > opt.niter:
            model.update_learning_rate()
    return steps


def main(argv=None):
    """python -m pix2pixhdaudiosr_b200.train_loop --name run1 [--continue_train] [--steps N] ...   (synthetic data)"""
    import argparse
    from . import options
    from .models.models import create_model
    ap = argparse.ArgumentParser()
    ap.add_argument('--name', default='g3l2_synth')
    ap.add_argument('--checkpoints_dir', default='./checkpoints')
    ap.add_argument('--continue_train', action='store_true')
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--batchSize', type=int, default=8)
    ap.add_argument('--dataset_size', type=int, default=64, help='samples per synthetic epoch')
    ap.add_argument('--transform', default='mdct4', choices=['mdct4', 'mdct2'])
    ap.add_argument('--use_time_D', action='store_true')
    ap.add_argument('--use_match_loss', action='store_true')
    ap.add_argument('--print_freq', type=int, default=16)
    ap.add_argument('--save_latest_freq', type=int, default=64)
    a = ap.parse_args(argv)
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if int(os.environ.get('WORLD_SIZE', '1')) > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    kw = dict(name=a.name, checkpoints_dir=a.checkpoints_dir, continue_train=a.continue_train, batchSize=a.batchSize,
              gpu_ids=[local], use_time_D=a.use_time_D, use_match_loss=a.use_match_loss, transform=a.transform)
    if a.transform == 'mdct2':
        kw.update(n_fft=512, hop_length=256, win_length=512, segment_length=127 * 256)
    opt = options.g3l2_opt(**kw)
    torch.manual_seed(opt.seed)
    model = create_model(opt)
    rank = int(os.environ.get('RANK', '0'))
    n = run(opt, model, lambda epoch: synthetic_dataset(a.dataset_size // a.batchSize, a.batchSize, opt.segment_length,
                                                        seed=epoch * 131 + rank),
            a.dataset_size, max_steps=a.steps, print_freq=a.print_freq, save_latest_freq=a.save_latest_freq)
    print('%d optimisation steps' % n)


if __name__ == '__main__':
    main()
