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
import torch
import torch.distributed as dist


def assemble_losses(model, losses):
    """train.py:155-160"""
    losses = [torch.mean(x) if not isinstance(x, int) else x for x in losses]
    loss_dict = dict(zip(model.loss_names, losses))
    loss_D = (loss_dict['D_fake'] + loss_dict['D_real']) * 0.5 + (loss_dict.get('D_fake_t', 0) + loss_dict.get('D_real_t', 0)) * 0.5
    loss_G = (loss_dict['G_GAN'] + loss_dict.get('G_mat', 0) + loss_dict.get('G_GAN_Feat', 0) + loss_dict.get('G_VGG', 0)
              + loss_dict.get('G_GAN_t', 0))
    return loss_dict, loss_G, loss_D


class GradSync:
    """average the flat gradient buffers over the data-parallel group (no-op for a single process).

    attach_early(optimizer): for an optimizer whose conv weights receive exactly ONE gradient per step (the generator's),
    every large weight gradient is all-reduced as soon as its wgrad kernel is enqueued, so the transfer of the bulk of
    the bytes (the 1536-channel residual blocks) overlaps the rest of the backward pass; start() then only reduces
    what is left."""

    EARLY_MIN = 1 << 20            # elements; smaller gradients travel with the final all-reduce

    def __init__(self, group=None):
        self.enabled = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        self.group = group
        self.world = dist.get_world_size(group) if self.enabled else 1
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


def broadcast_parameters(model, src=0, group=None):
    """identical replicas at start (the reference's DataParallel re-broadcasts every step)"""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return
    for opt in (model.optimizer_G, model.optimizer_D):
        dist.broadcast(opt.flat, src=src, group=group)
        opt.flat_bf16.copy_(opt.flat)


def train_step(model, lr_audio, hr_audio, inst=None, feat=None, infer=False, sync=None):
    """returns (loss_dict, generated-or-None); loss values stay on the device (no host sync in the step)"""
    losses, generated = model(lr_audio, inst, hr_audio, feat, infer)
    loss_dict, loss_G, loss_D = assemble_losses(model, losses)
    opt_G, opt_D = model.optimizer_G, model.optimizer_D
    opt_G.zero_grad()
    loss_G.backward()
    h_G = sync.start(opt_G) if sync is not None else None
    opt_D.zero_grad()
    loss_D.backward()
    h_D = sync.start(opt_D) if sync is not None else None
    if sync is not None:
        sync.finish(opt_G, h_G)
    opt_G.step()
    if sync is not None:
        sync.finish(opt_D, h_D)
    opt_D.step()
    return {k: (v.detach() if torch.is_tensor(v) else v) for k, v in loss_dict.items()}, generated


class GraphedTrainStep:
    """The whole optimisation step (MDCT4 -> encode -> G -> 3 x D -> losses -> 2 backward passes -> gradient
    all-reduce -> 2 Adam steps) captured once into a CUDA graph and replayed: ~1500 kernel launches per step become
    one graph launch, so the step is bounded by the device, not by Python / launch latency.  Inputs are copied into
    static device buffers; the loss values of the last replay are in `self.losses` (device tensors)."""

    def __init__(self, model, lr_audio, hr_audio, sync=None, warmup=3):
        self.model, self.sync = model, sync
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
        if lr_audio.data_ptr() != self.lr_static.data_ptr():
            self.lr_static.copy_(lr_audio, non_blocking=True)
        if hr_audio.data_ptr() != self.hr_static.data_ptr():
            self.hr_static.copy_(hr_audio, non_blocking=True)
        self.graph.replay()
        return self.losses, None
