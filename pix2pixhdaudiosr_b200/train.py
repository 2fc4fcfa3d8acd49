"""One optimisation step of the reference's train.py:148-184, and its data-parallel form.

Reference order:  losses = model(lr, inst, hr, feat, infer)  ->  loss_D, loss_G  ->
    optimizer_G.zero_grad(); loss_G.backward(); optimizer_G.step()
    optimizer_D.zero_grad(); loss_D.backward(); optimizer_D.step()
The reference's only multi-GPU mode is torch.nn.DataParallel (models/models.py:17-18), unreachable with --fp16.
Here: one process per GPU (torchrun), each rank a full replica on its own batch shard; the ONLY exchange is the
average of the flat gradient buffers (one NCCL all-reduce per network over NVLink / NVSwitch).  G's all-reduce is
issued asynchronously right after loss_G.backward() and overlaps loss_D.backward(): the discriminator's backward
touches neither G's weights nor G's gradients, so applying optimizer_G.step() after it is arithmetically the
reference's order."""
import os

import torch
import torch.distributed as dist


def assemble_losses(model, losses):
    """train.py:155-160"""
    # (the mean of a 0-dim loss and the "+ 0" of an absent term are identities: not launched as kernels)
    losses = [x if isinstance(x, int) or x.dim() == 0 else torch.mean(x) for x in losses]
    loss_dict = dict(zip(model.loss_names, losses))

    def total(*names):
        terms = [loss_dict[k] for k in names if k in loss_dict and not isinstance(loss_dict[k], int)]
        return sum(terms[1:], terms[0]) if terms else 0
    loss_D = total('D_fake', 'D_real') * 0.5
    if 'D_fake_t' in loss_dict or 'D_real_t' in loss_dict:
        loss_D = loss_D + total('D_fake_t', 'D_real_t') * 0.5
    loss_G = total('G_GAN', 'G_mat', 'G_GAN_Feat', 'G_VGG', 'G_GAN_t')
    return loss_dict, loss_G, loss_D


class GradSync:
    """average the flat gradient buffers over the data-parallel group (no-op for a single process).

    attach_early(optimizer): for an optimizer whose conv weights receive exactly ONE gradient per step (the generator's),
    every large weight gradient is all-reduced as soon as its wgrad kernel is enqueued, so the transfer of the bulk of
    the bytes (the 1536-channel residual blocks) overlaps the rest of the backward pass; start() then only reduces
    what is left."""

    EARLY_MIN = 1 << 20            # elements; smaller gradients travel with the final all-reduce

    def __init__(self, group=None, shard=False):
        """shard=True (bf16 training): reduce-scatter the gradients, let every rank run Adam on its 1/world slice only and
        all-gather the bf16 weight shadow the kernels read -- the same bytes over NVLink as the all-reduce would move in fp32
        minus a quarter, and Adam's 30 bytes per parameter of HBM traffic divided by the world size.  Between steps the fp32
        master weights and the moments are complete only per shard (FusedAdam.gather_state before a checkpoint)."""
        self.enabled = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        self.group = group
        self.world = dist.get_world_size(group) if self.enabled else 1
        self.rank = dist.get_rank(group) if self.enabled else 0
        self.shard = bool(shard) and self.enabled
        self._early = {}           # optimizer id -> {"handles": [...], "ranges": [(offset, n), ...]}

    def attach_early(self, optimizer):
        if not self.enabled:
            return
        st = self._early[id(optimizer)] = {"handles": [], "ranges": []}

        def hook(sink):
            n = sink.view.numel()
            if n >= self.EARLY_MIN:
                st["handles"].append(dist.all_reduce(sink.view, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
                st["ranges"].append((sink.offset, n))
        optimizer.grad_written_hook = hook

    def start(self, optimizer):
        if not self.enabled:
            return None
        if self.shard:
            if optimizer.shard is None:
                optimizer.set_shard(self.rank, self.world)
            if dist.get_backend(self.group) == "gloo":      # CPU tests: gloo has no reduce-scatter -- all-reduce, keep the own slice
                h = dist.all_reduce(optimizer.flat_grad, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
                h.wait()
                lo, hi = optimizer.shard_range()
                optimizer.grad_shard.copy_(optimizer.flat_grad[lo:hi])
                return []
            return [dist.reduce_scatter_tensor(optimizer.grad_shard, optimizer.flat_grad, op=dist.ReduceOp.SUM, group=self.group,
                                               async_op=True)]
        st = self._early.get(id(optimizer))
        if st is None or not st["ranges"]:
            return [dist.all_reduce(optimizer.flat_grad, op=dist.ReduceOp.SUM, group=self.group, async_op=True)]
        handles, pos, total = st["handles"], 0, optimizer.flat_grad.numel()
        for o, n in sorted(st["ranges"]) + [(total, 0)]:       # the gaps between the ranges already in flight
            if o > pos:
                handles.append(dist.all_reduce(optimizer.flat_grad[pos:o], op=dist.ReduceOp.SUM, group=self.group, async_op=True))
            pos = max(pos, o + n)
        st["handles"], st["ranges"] = [], []
        return handles

    def finish(self, optimizer, handles):
        if handles is None:
            return
        for h in handles:
            h.wait()                            # orders the current stream after the collective
        optimizer.grad_scale = 1.0 / self.world     # the averaging is folded into the Adam kernel's gradient read


    def after_step(self, optimizer):
        """sharded mode: all-gather the bf16 shadow slices the ranks just updated (asynchronous; finish_gather waits)"""
        if not self.shard:
            return None
        lo, hi = optimizer.shard_range()
        return dist.all_gather_into_tensor(optimizer.flat_bf16, optimizer.flat_bf16[lo:hi].clone(), group=self.group, async_op=True)

    def finish_gather(self, handle):
        if handle is not None:
            handle.wait()

    def gather_state(self, model):
        """sharded mode: complete master weights / moments on every rank (call before saving a checkpoint)"""
        if self.shard:
            for opt in (model.optimizer_G, model.optimizer_D):
                opt.gather_state(self.group)


class _NoSync:
    """single process: every exchange is a no-op"""
    enabled = shard = False

    def start(self, optimizer):
        return None

    def finish(self, optimizer, handles):
        pass

    def after_step(self, optimizer):
        return None

    def finish_gather(self, handle):
        pass


_NOSYNC = _NoSync()


def replica_drift(model, group=None):
    """max - min over the ranks of a checksum of every bf16 weight the kernels read (0 = the replicas are bit-identical)"""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return 0
    out = 0
    for opt in (model.optimizer_G, model.optimizer_D):
        v = opt.flat_bf16.view(torch.int16).to(torch.int64)
        idx = torch.arange(v.numel(), device=v.device, dtype=torch.int64) % 8191 + 1
        c = (v * idx).sum().reshape(1)
        hi, lo = c.clone(), c.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        out = max(out, int((hi - lo).item()))
    return out


def broadcast_parameters(model, src=0, group=None):
    """identical replicas at start (the reference's DataParallel re-broadcasts every step)"""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return
    for opt in (model.optimizer_G, model.optimizer_D):
        dist.broadcast(opt.flat, src=src, group=group)
        opt.flat_bf16.copy_(opt.flat)


OVERLAP_G_STEP = os.environ.get("P2P_OVERLAP_GSTEP", "1") == "1"
SHARE_FAKE_PASS = os.environ.get("P2P_SHARE_FAKE_PASS", "1") == "1"
# with the shared pass: the discriminator's forward on the real image joins it -- ONE forward over cat(real, fake), ONE
# full-batch backward for loss_D, a fake-half backward for loss_G (Pix2PixHDModel.forward, ops.BatchHalvesFn)
BATCH_D = os.environ.get("P2P_BATCH_D", "1") == "1"
_SIDE = {}


def _side_stream(device):
    key = torch.device(device)
    if key not in _SIDE:
        _SIDE[key] = torch.cuda.Stream(device=key)
    return _SIDE[key]


def train_step(model, lr_audio, hr_audio, inst=None, feat=None, infer=False, sync=None):
    """returns (loss_dict, generated-or-None); loss values stay on the device (no host sync in the step).

    The generator's update -- gradient exchange, Adam (156 M parameters, HBM-bound), all-gather of the bf16 shadow -- runs on
    a side stream WHILE the main stream computes the discriminator's backward pass: loss_D.backward() reads neither G's
    weights nor G's gradients, so this is arithmetically the reference's order (train.py:164-184), and the bandwidth-bound
    optimiser kernels share the SMs with the tensor-core-bound convolutions instead of running after them."""
    shared = SHARE_FAKE_PASS and hasattr(model, 'netD') and getattr(model.opt, 'pool_size', 0) == 0
    model._share_fake_pass = shared          # one D forward on the generated image serves both losses (see Pix2PixHDModel.forward)
    batched = shared and BATCH_D
    model._batch_D = batched
    try:
        losses, generated = model(lr_audio, inst, hr_audio, feat, infer)
    finally:
        model._share_fake_pass = False
        model._batch_D = False
    loss_dict, loss_G, loss_D = assemble_losses(model, losses)
    opt_G, opt_D = model.optimizer_G, model.optimizer_D
    from . import ops
    opt_G.zero_grad()
    try:
        if shared:
            ops.set_backward_phase(owners=[opt_G], gate_open=True, fake_half_only=batched)
        ops.wgrad_overlap_begin()           # window weight gradients on a side stream (ops._SideWgrad), joined below
        loss_G.backward(retain_graph=shared)
    finally:
        ops.wgrad_overlap_join()
        ops.set_backward_phase(None, True)
    if sync is None:
        sync = _NOSYNC

    def update(opt):
        h = sync.start(opt)
        sync.finish(opt, h)
        opt.step()
        sync.finish_gather(sync.after_step(opt))

    overlap = OVERLAP_G_STEP and opt_G.flat.is_cuda
    if overlap:
        main = torch.cuda.current_stream(opt_G.flat.device)
        side = _side_stream(opt_G.flat.device)
        side.wait_stream(main)
        with torch.cuda.stream(side):
            update(opt_G)
    else:
        h_G = sync.start(opt_G)
    opt_D.zero_grad()
    try:
        if shared:
            ops.set_backward_phase(owners=[opt_D], gate_open=False)
        ops.wgrad_overlap_begin()
        loss_D.backward()
    finally:
        ops.wgrad_overlap_join()
        ops.set_backward_phase(None, True)
    if not overlap:
        sync.finish(opt_G, h_G)
        opt_G.step()
        g_G = sync.after_step(opt_G)
        update(opt_D)
        sync.finish_gather(g_G)
    else:
        update(opt_D)
        main.wait_stream(side)
    return {k: (v.detach() if torch.is_tensor(v) else v) for k, v in loss_dict.items()}, generated


class GraphedTrainStep:
    """The whole optimisation step (MDCT4 -> encode -> G -> 3 x D -> losses -> 2 backward passes -> gradient
    all-reduce -> 2 Adam steps) captured once into a CUDA graph and replayed: ~1500 kernel launches per step become
    one graph launch, so the step is bounded by the device, not by Python / launch latency.  Inputs are copied into
    static device buffers; the loss values of the last replay are in `self.losses` (device tensors)."""

    def __init__(self, model, lr_audio, hr_audio, sync=None, warmup=3):
        """NOTE: the `warmup` eager steps before the capture are REAL optimisation steps (they move the weights and Adam's
        state, like any other step).  The learning rate lives on the device: __call__ pushes param_groups['lr'] changes
        (update_learning_rate) before every replay.  After model.update_fixed_params() (a new optimizer_G with new flat
        buffers) the captured graph is stale: build a new GraphedTrainStep -- __call__ refuses to replay a stale graph."""
        self.model, self.sync = model, sync
        self._opt_ids = (id(model.optimizer_G), id(model.optimizer_D))
        dev = torch.device(model.device)
        self.lr_static = lr_audio.to(dev).clone()
        self.hr_static = hr_audio.to(dev).clone()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):             # allocator warm-up, cudaFuncSetAttribute calls, NCCL channel set-up
                train_step(model, self.lr_static, self.hr_static, sync=sync)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        for opt in (model.optimizer_G, model.optimizer_D):
            opt.sync_lr()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.losses, _ = train_step(model, self.lr_static, self.hr_static, sync=sync)

    def __call__(self, lr_audio, hr_audio):
        if (id(self.model.optimizer_G), id(self.model.optimizer_D)) != self._opt_ids:
            raise RuntimeError("the model's optimizers were replaced after the step was captured (update_fixed_params): "
                               "build a new GraphedTrainStep")
        for opt in (self.model.optimizer_G, self.model.optimizer_D):
            opt.sync_lr()
        if lr_audio.data_ptr() != self.lr_static.data_ptr():
            self.lr_static.copy_(lr_audio, non_blocking=True)
        if hr_audio.data_ptr() != self.hr_static.data_ptr():
            self.hr_static.copy_(hr_audio, non_blocking=True)
        self.graph.replay()
        return self.losses, None
