"""The N>1 host logic on CPU: two `gloo` processes run the gradient exchange of pix2pixhdaudiosr_b200/train.py
(GradSync / broadcast_parameters) on stand-in flat buffers and must end with identical, averaged gradients and
identical parameters -- what each data-parallel replica feeds its fused Adam step with."""
import os
import socket
import types

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pix2pixhdaudiosr_b200.train import GradSync, broadcast_parameters
    torch.manual_seed(rank)
    mk = lambda n: types.SimpleNamespace(flat=torch.randn(n), flat_bf16=torch.zeros(n, dtype=torch.bfloat16),
                                         flat_grad=torch.randn(n), grad_scale=1.0, grad_written_hook=None, grad_epoch=0)
    model = types.SimpleNamespace(optimizer_G=mk(1003), optimizer_D=mk(77))
    local_grads = [model.optimizer_G.flat_grad.clone(), model.optimizer_D.flat_grad.clone()]
    broadcast_parameters(model)
    sync = GradSync()
    assert sync.enabled and sync.world == world
    # early path: two "large" weight gradients of G are reduced the moment they are written, the rest at start()
    sync.EARLY_MIN = 100
    sync.attach_early(model.optimizer_G)
    from pix2pixhdaudiosr_b200.ops import GradSink
    g = model.optimizer_G
    for o, n in ((640, 200), (128, 256), (8, 16)):       # the last one is below EARLY_MIN: it must ride with the rest
        GradSink(g.flat_grad[o:o + n], g, o).written()
    hG = sync.start(model.optimizer_G)
    assert len(hG) == 2 + 3                               # 2 early ranges + the 3 gaps around them
    hD = sync.start(model.optimizer_D)           # both collectives in flight, as in train_step
    sync.finish(model.optimizer_G, hG)
    sync.finish(model.optimizer_D, hD)
    out[rank] = dict(wG=model.optimizer_G.flat.clone(), w16=model.optimizer_G.flat_bf16.clone(),
                     gG=model.optimizer_G.flat_grad * model.optimizer_G.grad_scale,
                     gD=model.optimizer_D.flat_grad * model.optimizer_D.grad_scale, local=local_grads)
    dist.destroy_process_group()


def test_two_rank_gradient_average_and_parameter_broadcast():
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    a, b = out[0], out[1]
    assert torch.equal(a["wG"], b["wG"])                                   # replicas start identical (rank 0's)
    assert torch.equal(a["w16"], a["wG"].to(torch.bfloat16))               # bf16 shadow refreshed from the master copy
    for k, i in (("gG", 0), ("gD", 1)):
        want = (a["local"][i] + b["local"][i]) / world
        assert torch.allclose(a[k], want, atol=1e-6) and torch.equal(a[k], b[k])


def test_single_process_sync_is_a_no_op():
    from pix2pixhdaudiosr_b200.train import GradSync
    s = GradSync()
    assert not s.enabled and s.start(None) is None
    s.finish(None, None)


def _shard_worker(rank, world, port, out):
    """sharded mode on stand-in buffers: every rank steps only its slice (a plain SGD stand-in for the fused Adam kernel) and the
    bf16 shadow must come back complete and identical on all ranks"""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pix2pixhdaudiosr_b200.train import GradSync, replica_drift

    class Opt:
        def __init__(self, n):
            g = torch.Generator().manual_seed(5)
            self.flat = torch.randn(n, generator=g)
            self.flat_bf16 = self.flat.to(torch.bfloat16)
            torch.manual_seed(100 + rank)
            self.flat_grad = torch.randn(n)
            self.grad_scale, self.shard, self.grad_shard = 1.0, None, None

        def set_shard(self, r, w):
            self.shard = (r, w)
            self.grad_shard = torch.zeros(self.flat.numel() // w)

        def shard_range(self):
            r, w = self.shard
            n = self.flat.numel() // w
            return r * n, (r + 1) * n

        def step(self):
            lo, hi = self.shard_range()
            self.flat[lo:hi] -= 0.1 * self.grad_scale * self.grad_shard
            self.flat_bf16[lo:hi] = self.flat[lo:hi].to(torch.bfloat16)

    opt = Opt(256)
    local = opt.flat_grad.clone()
    sync = GradSync(shard=True)
    assert sync.shard and sync.rank == rank
    h = sync.start(opt)
    sync.finish(opt, h)
    opt.step()
    sync.finish_gather(sync.after_step(opt))
    model = types.SimpleNamespace(optimizer_G=opt, optimizer_D=opt)
    out[rank] = dict(w16=opt.flat_bf16.clone(), local=local, drift=replica_drift(model), w0=torch.randn(256, generator=torch.Generator().manual_seed(5)))
    dist.destroy_process_group()


def test_two_rank_sharded_step_and_gather():
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_shard_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    a, b = out[0], out[1]
    want = (a["w0"] - 0.1 * (a["local"] + b["local"]) / world).to(torch.bfloat16)
    assert torch.equal(a["w16"], b["w16"]) and a["drift"] == 0 and b["drift"] == 0
    assert torch.allclose(a["w16"].float(), want.float(), atol=1e-2)
