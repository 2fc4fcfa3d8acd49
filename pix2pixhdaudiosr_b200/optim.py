"""Fused Adam over flat buffers (reference: torch.optim.Adam(params, lr=opt.lr, betas=(opt.beta1, 0.999)) at
models/pix2pixHD_model.py:131, 140 and optimizer_X.zero_grad()/step() at train.py:164-184).

All parameters handed to the optimiser are re-homed into ONE fp32 master buffer (conv weights in their physical
tap-major layout), with matching flat gradient / moment buffers and a flat bf16 shadow that the tcgen05 kernels
read.  `step()` is a single `p2p_adam_step` launch that also refreshes the bf16 shadow; `zero_grad()` is one
memset; the data-parallel gradient exchange is one all-reduce of `flat_grad` (train.py of this package)."""
import os

import torch

from . import ops

# zero_grad() skips the slices of the flat gradient buffer that the weight-gradient kernels STORE into (first write of an
# epoch, ops.GradSink) when they are large: on G3L2-ngf48 the six 1536x1536x3x3 bottleneck weights are 127 M of the 156 M
# gradient words, a 0.5 GB memset per step.  P2P_LAZY_ZERO_GRAD=0 restores the single full memset.
LAZY_ZERO_GRAD = os.environ.get("P2P_LAZY_ZERO_GRAD", "1") == "1"
_LAZY_MIN = 1 << 20      # elements: smaller slices are not worth a separate memset range

_ALIGN = 8      # elements: keeps every bf16 slice 16-byte aligned for the vector / tensor-core loads
_SHARD_ALIGN = 128   # the flat buffers are a multiple of this: equal, 16-byte aligned shards for up to 16 data-parallel ranks


class FusedAdam:
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        params = [p for p in params]
        if not params:
            raise ValueError("optimizer got an empty parameter list")
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FusedAdam needs CUDA parameters (no CPU fallback): move the networks to the GPU first")
        self.param_groups = [{"params": params, "lr": lr, "betas": tuple(betas), "eps": eps}]
        sizes = [(p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN for p in params]
        total = (sum(sizes) + _SHARD_ALIGN - 1) // _SHARD_ALIGN * _SHARD_ALIGN
        self.shard = None                                                  # (rank, world): train.GradSync's sharded mode
        self.grad_shard = None
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(total, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(total, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_bf16 = torch.zeros(total, dtype=torch.bfloat16, device=dev)
        self.step_count = 0
        self.grad_written_hook = None                                      # set by train.GradSync.attach_early
        self.grad_epoch = 0                                                # bumped by zero_grad (see ops.GradSink)
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)      # the authoritative step counter
        self.lr_dev = torch.full((1,), float(lr), dtype=torch.float32, device=dev)
        self._lr_on_dev = float(lr)
        self._scratch = torch.zeros(2, dtype=torch.float32, device=dev)
        self.grad_scale = 1.0                                              # 1 / world size under data parallelism
        self._views = []
        self._sinks = []
        self._skipped = []                                                 # sinks left un-zeroed by the last zero_grad
        self._seg_key, self._segs = None, None
        o = 0
        for p, sz in zip(params, sizes):
            n = p.numel()
            conv = getattr(p, "_p2p_conv", None)
            if conv is not None:
                perm = conv._perm
                inv = [perm.index(i) for i in range(4)]
                phys = p.data.permute(*perm)
                assert phys.is_contiguous(), "conv weights must be stored tap-major"
                self.flat[o:o + n].copy_(phys.reshape(-1))
                p.data = self.flat[o:o + n].view(phys.shape).permute(*inv)
                gview = self.flat_grad[o:o + n].view(phys.shape).permute(*inv)
                conv.w16 = self.flat_bf16[o:o + n].view(phys.shape)
                conv.gphys = ops.GradSink(self.flat_grad[o:o + n].view(phys.shape), self, o)   # wgrad kernels write here directly
                self._sinks.append(conv.gphys)
            else:
                self.flat[o:o + n].copy_(p.data.reshape(-1))
                p.data = self.flat[o:o + n].view(p.shape)
                gview = self.flat_grad[o:o + n].view(p.shape)
            p.grad = gview
            self._views.append((p, gview))
            o += sz
        self.flat_bf16.copy_(self.flat)

    def zero_grad(self, set_to_none=False):
        # slices whose first write of the epoch that just ended was a STORE by the window weight-gradient kernel are left
        # alone (the same layer call stores again); ops._conv_backward zeroes such a slice itself if it ever has to
        # accumulate into it, and step() zeroes the ones that received no gradient at all
        skip = [k for k in self._sinks if k.stored_epoch == self.grad_epoch and k.view.numel() >= _LAZY_MIN] \
            if LAZY_ZERO_GRAD and self.grad_epoch > 0 else []
        if not skip:
            self.flat_grad.zero_()
        else:
            key = tuple(k.offset for k in skip)
            if key != self._seg_key:
                segs, lo = [], 0
                for k in sorted(skip, key=lambda q: q.offset):
                    if k.offset > lo:
                        segs.append((lo, k.offset))
                    lo = k.offset + k.view.numel()
                if lo < self.flat_grad.numel():
                    segs.append((lo, self.flat_grad.numel()))
                self._seg_key, self._segs = key, segs
            for lo, hi in self._segs:
                self.flat_grad[lo:hi].zero_()
        self._skipped = skip
        self.grad_epoch += 1
        for k in skip:
            k.unzeroed_epoch = self.grad_epoch
        for p, g in self._views:
            if p.grad is not g:
                p.grad = g

    def sync_lr(self):
        """push param_groups[0]['lr'] to the device (call after changing it; never inside a graph capture)"""
        lr = float(self.param_groups[0]["lr"])
        if lr != self._lr_on_dev:
            self.lr_dev.fill_(lr)
            self._lr_on_dev = lr

    @torch.no_grad()
    def step(self):
        g = self.param_groups[0]
        if not torch.cuda.is_current_stream_capturing():
            self.sync_lr()
        for k in self._skipped:                 # left un-zeroed and not written this epoch: its gradient is zero, not last step's
            if k.epoch != self.grad_epoch:
                k.view.zero_()
        self._skipped = []
        self.step_count += 1
        if self.shard is not None:
            # sharded data parallelism (train.GradSync(shard=True)): this rank owns 1/world of the flat buffers -- its
            # gradients arrived by reduce-scatter in grad_shard, it updates its slice of the master weights, the moments and the
            # bf16 shadow; the shadow slices are all-gathered afterwards.  Adam's HBM traffic per rank drops by the world size.
            lo, hi = self.shard_range()
            ops.adam_step_dev(self.flat[lo:hi], self.grad_shard, self.exp_avg[lo:hi], self.exp_avg_sq[lo:hi], self.flat_bf16[lo:hi],
                              self.lr_dev, g["betas"][0], g["betas"][1], g["eps"], self.grad_scale, self.step_dev, self._scratch)
            return
        ops.adam_step_dev(self.flat, self.flat_grad, self.exp_avg, self.exp_avg_sq, self.flat_bf16, self.lr_dev,
                          g["betas"][0], g["betas"][1], g["eps"], self.grad_scale, self.step_dev, self._scratch)

    def shard_range(self):
        rank, world = self.shard
        n = self.flat.numel() // world
        return rank * n, (rank + 1) * n

    def set_shard(self, rank, world):
        if self.flat.numel() % (world * 4) != 0:
            raise ValueError("flat buffers do not split into %d aligned shards" % world)
        self.shard = (rank, world)
        if self.grad_shard is None or self.grad_shard.numel() != self.flat.numel() // world:
            self.grad_shard = torch.zeros(self.flat.numel() // world, dtype=torch.float32, device=self.flat.device)

    def gather_state(self, group=None):
        """sharded mode: bring the master weights and both moments of every shard to every rank (before a checkpoint is
        written; between steps only the bf16 shadow is complete everywhere)"""
        if self.shard is None:
            return
        import torch.distributed as dist
        lo, hi = self.shard_range()
        for buf in (self.flat, self.exp_avg, self.exp_avg_sq):
            dist.all_gather_into_tensor(buf, buf[lo:hi].clone(), group=group)

    def state_dict(self):
        return {"step": int(self.step_dev.item()), "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq,
                "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}]}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.step_dev.fill_(self.step_count)
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.param_groups[0].update(sd["param_groups"][0])
