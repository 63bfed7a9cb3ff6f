"""Training-step pieces around the block path, mirroring what the reference driver uses:
``nn.L1Loss()`` (MoCE-IR-main/src/train.py:51), ``optim.AdamW(self.parameters(), lr=2e-4)`` (:83) and DDP gradient
averaging (:131).  Everything numeric runs in hand-written kernels (csrc/backward.cu); torch only owns the buffers.

``AdamW`` keeps parameters, gradients and both moments in four flat fp32 buffers (the module's ``nn.Parameter``s become
views of the first), so one optimizer step is ONE kernel launch over 26 M elements and the DDP all-reduce is ONE NCCL
call on the flat gradient buffer.
"""
from __future__ import annotations

from typing import Iterable, Optional

import torch
import torch.nn as nn

from . import autograd_ops as A
from . import ops


class L1Loss(nn.Module):
    """Drop-in for ``nn.L1Loss()`` (mean reduction only, as the reference uses it)."""

    def __init__(self, reduction: str = "mean"):
        super().__init__()
        if reduction != "mean":
            raise ValueError("b200ir.L1Loss supports reduction='mean' only")

    def forward(self, pred, target):
        if pred.shape != target.shape:
            raise ValueError("L1Loss: shape mismatch")
        return A.l1_loss(pred, target)


class AdamW(torch.optim.Optimizer):
    """``torch.optim.AdamW`` semantics (decoupled weight decay, bias correction, eps outside the sqrt) on flat buffers.

    A ``torch.optim.Optimizer`` subclass with ONE parameter group, so that the reference recipe
    (``optim.AdamW(self.parameters(), lr=2e-4)`` + ``LinearWarmupCosineAnnealingLR``, MoCE-IR-main/src/train.py:83-86)
    attaches unchanged: schedulers write ``param_groups[0]['lr']``; ``step()`` mirrors lr / weight decay into a
    two-float device buffer that the update kernel reads, so they also change under a captured CUDA graph
    (``GraphedTrainStep``).  Parameters must be fp32 CUDA tensors; after construction each ``p.data`` is a view into
    ``self.flat_p``.  Parameters whose ``.grad`` is None at ``step()`` are skipped (no decay, no moment update), like
    torch.  Difference from torch: the bias-correction step count is one counter for all parameters (torch counts per
    parameter, which only differs for a parameter that receives gradients in some steps and not in others).
    """

    def __init__(self, params: Iterable[nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2):
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        if len(self.param_groups) != 1:
            raise ValueError("b200ir.AdamW supports a single parameter group (one flat buffer)")
        self.params = [p for p in self.param_groups[0]["params"] if p.requires_grad]
        if not self.params:
            raise ValueError("AdamW: no trainable parameters")
        dev = self.params[0].device
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or p.device != dev:
                raise ValueError("AdamW: parameters must be float32 CUDA tensors on one device (no CPU fallback)")
        self.offsets, n = [], 0
        for p in self.params:
            self.offsets.append(n)
            n += (p.numel() + 3) // 4 * 4          # keep every view 16-byte aligned
        self.numel = n
        self.flat_p = torch.zeros(n, device=dev, dtype=torch.float32)
        self.flat_g = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        self.grad_views = []
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.flat_p[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                self.grad_views.append(self.flat_g[o:o + p.numel()].view(p.shape))
        self.step_count = 0
        # the step count also lives on the device: the update kernel increments and reads it there, so that a captured
        # CUDA graph of the whole training step replays with the right bias corrections; lr / weight decay likewise
        self.step_dev = torch.zeros(1, device=dev, dtype=torch.int32)
        self.hyper = torch.tensor([lr, weight_decay], device=dev, dtype=torch.float32)
        self._hyper_last = (float(lr), float(weight_decay))
        self._present = None
        self._runs = [(0, n)]
        A.bump_param_epoch()

    # scalar hyper-parameters live in param_groups[0] (what torch schedulers edit)
    lr = property(lambda self: self.param_groups[0]["lr"], lambda self, v: self.param_groups[0].__setitem__("lr", v))
    betas = property(lambda self: self.param_groups[0]["betas"],
                     lambda self, v: self.param_groups[0].__setitem__("betas", tuple(v)))
    eps = property(lambda self: self.param_groups[0]["eps"], lambda self, v: self.param_groups[0].__setitem__("eps", v))
    weight_decay = property(lambda self: self.param_groups[0]["weight_decay"],
                            lambda self, v: self.param_groups[0].__setitem__("weight_decay", v))

    def zero_grad(self, set_to_none: bool = True):
        for p in self.params:
            p.grad = None

    def sync_hyper(self):
        """Mirror (lr, weight_decay) of param_groups[0] into the device buffer the kernel reads (8-byte copy, only when
        they changed; never inside a stream capture: GraphedTrainStep calls this before each replay)."""
        cur = (float(self.lr), float(self.weight_decay))
        if cur != self._hyper_last and not torch.cuda.is_current_stream_capturing():
            self.hyper.copy_(torch.tensor(cur, dtype=torch.float32))
            self._hyper_last = cur

    def gather_grads(self):
        """Collect ``p.grad`` into the flat gradient buffer (no-op for gradients that already live there) and note which
        parameters have one: the update runs over the contiguous runs of those."""
        src, dst, present = [], [], []
        odd = []
        for p, gv in zip(self.params, self.grad_views):
            g = p.grad
            present.append(g is not None)
            if g is not None and g.data_ptr() != gv.data_ptr():
                # torch._foreach_copy_ takes its multi-tensor route only if EVERY pair matches in strides and dtype; one
                # strided gradient (the depthwise taps are views into a [C, 10] fold) sends all ~500 tensors down the
                # per-tensor route: one memcpy / copy kernel each.  So the odd ones are copied on their own.
                if g.is_contiguous() and g.dtype == gv.dtype and g.shape == gv.shape:
                    src.append(g)
                    dst.append(gv)
                else:
                    odd.append((gv, g))
        if src:
            torch._foreach_copy_(dst, src)
        for gv, g in odd:
            gv.copy_(g.view(gv.shape) if g.shape != gv.shape and g.is_contiguous() else g)
        if present != self._present:
            self._present, self._runs, start = present, [], None
            for i, ok in enumerate(present):
                if ok and start is None:
                    start = self.offsets[i]
                if not ok and start is not None:
                    self._runs.append((start, self.offsets[i]))
                    start = None
            if start is not None:
                self._runs.append((start, self.numel))
        return self.flat_g

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0, gathered: bool = False):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not gathered:
            self.gather_grads()
        self.sync_hyper()
        if not self._runs:
            return loss
        self.step_count += 1
        b1, b2 = self.betas
        for i, (a, b) in enumerate(self._runs):
            ops.adamw_step(self.flat_p[a:b], self.flat_g[a:b], self.exp_avg[a:b], self.exp_avg_sq[a:b], float(self.lr),
                           b1, b2, self.eps, float(self.weight_decay), self.step_count, grad_scale,
                           step_dev=self.step_dev, tick=(i == 0), hyper_dev=self.hyper)
        A.bump_param_epoch()   # packed bf16 weight caches key on this (the kernel does not bump tensor versions)
        return loss

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "lr": self.lr,
                "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.step_dev.fill_(self.step_count)
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.lr, self.betas, self.eps, self.weight_decay = sd["lr"], tuple(sd["betas"]), sd["eps"], sd["weight_decay"]
        self.sync_hyper()


def allreduce_flat_grads(flat_g: torch.Tensor, group=None) -> float:
    """DDP gradient averaging (MoCE-IR-main/src/train.py:131): ONE sum all-reduce of the flat gradient buffer; returns
    the 1/world_size factor to hand to ``AdamW.step(grad_scale=...)`` (the division is fused into the update kernel)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return 1.0
    world = dist.get_world_size(group)
    if world == 1:
        return 1.0
    dist.all_reduce(flat_g, op=dist.ReduceOp.SUM, group=group)
    return 1.0 / world


def _backward_side(loss: torch.Tensor):
    """loss.backward() with the weight gradients on the side stream (autograd_ops._side), joined before returning"""
    A.SIDE_WGRAD = SIDE_WGRAD
    try:
        loss.backward()
    finally:
        A.SIDE_WGRAD = False
        A.join_side()


SIDE_WGRAD = True   # b200ir.training.SIDE_WGRAD = False restores the single-stream backward (A/B, debugging)


def train_step(net: nn.Module, opt: AdamW, degraded: torch.Tensor, clean: torch.Tensor, loss_fn: Optional[nn.Module] = None,
               group=None, balance_loss_weight: float = 0.0) -> torch.Tensor:
    """``training_step`` of the reference driver (MoCE-IR-main/src/train.py:59-72) + optimizer step:
    restored = net(degrad); loss = L1(restored, clean); backward; (all-reduce); AdamW.  Returns the loss (device scalar)."""
    loss_fn = loss_fn or L1Loss()
    opt.zero_grad()
    packing = degraded.is_cuda and getattr(net, "compute_dtype", None) == torch.bfloat16
    if packing:
        # every 1x1 weight pack of the step in ONE launch (the forward / backward find them cached).  The table is learnt:
        # packs that still miss their cache during a step (first step, a newly routed MoCE-IR expert) are recorded and join
        # the table for the following steps, whatever modules the model is made of.
        pre = net.__dict__.get("_b200ir_prepacker")
        if pre is not None and not pre.valid():
            pre = net.__dict__["_b200ir_prepacker"] = None
        if pre is not None:
            pre.run()
        elif not torch.cuda.is_current_stream_capturing():
            A.bump_param_epoch()      # first step: every pack misses its cache and is recorded
        A._REC = [] if not torch.cuda.is_current_stream_capturing() else None
    try:
        restored = net(degraded)
        loss = loss_fn(restored, clean)
        if balance_loss_weight:     # MoCE-IR: loss += balance_loss_weight * net.total_loss (MoCE-IR-main/src/train.py:62,71)
            loss = loss + balance_loss_weight * net.total_loss
        _backward_side(loss)
    finally:
        rec, A._REC = A._REC, None
    if packing and rec:
        old = pre.records if pre is not None else []
        net.__dict__["_b200ir_prepacker"] = A.Prepacker(records=old + rec)
    flat_g = opt.gather_grads()
    scale = allreduce_flat_grads(flat_g, group)
    opt.step(grad_scale=scale, gathered=True)
    return loss.detach()


class GraphedTrainStep:
    """The whole training step (forward, L1, backward, gradient gather, [all-reduce], AdamW) captured ONCE in a CUDA
    graph and replayed per batch: ~2000 kernel launches become one graph launch, which removes the CPU launch bound of
    small-patch training (cfg4: 128x128 patches).  Shapes are fixed by the example batch.

    With a process group of more than one rank the gradient all-reduce stays outside the graphs (two graphs: everything
    up to the flat gradient buffer, then AdamW) so that NCCL is called eagerly between them.
    """

    def __init__(self, net: nn.Module, opt: AdamW, degraded: torch.Tensor, clean: torch.Tensor,
                 loss_fn: Optional[nn.Module] = None, group=None, warmup: int = 2):
        import torch.distributed as dist
        self.net, self.opt, self.group = net, opt, group
        self.loss_fn = loss_fn or L1Loss()
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.degraded = degraded.detach().clone()
        self.clean = clean.detach().clone()
        # warm up eagerly on a side stream (lazy inits, cudaFuncSetAttribute, allocator), then restore the state
        snap = [t.clone() for t in (opt.flat_p, opt.exp_avg, opt.exp_avg_sq, opt.step_dev)]
        steps0 = opt.step_count
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                train_step(net, opt, self.degraded, self.clean, self.loss_fn, group)
        torch.cuda.current_stream().wait_stream(side)
        for dst, src in zip((opt.flat_p, opt.exp_avg, opt.exp_avg_sq, opt.step_dev), snap):
            dst.copy_(src)
        opt.step_count = steps0
        opt.zero_grad()
        A.bump_param_epoch()
        torch.cuda.synchronize()
        self.g1 = torch.cuda.CUDAGraph()
        self.g2 = None
        launches0 = ops.LAUNCHES
        pre = net.__dict__.get("_b200ir_prepacker")   # built by the warm-up steps (bf16 compute only)
        with torch.cuda.graph(self.g1):
            if pre is not None and pre.valid():
                pre.run()
            restored = net(self.degraded)
            loss = self.loss_fn(restored, self.clean)
            _backward_side(loss)
            opt.gather_grads()
            self.loss = loss.detach()
            if self.world == 1:
                opt.step(grad_scale=1.0, gathered=True)
        if self.world > 1:
            self.g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g2):
                opt.step(grad_scale=1.0 / self.world, gathered=True)
        # the capture itself did not execute anything: undo the host-side step bookkeeping of the captured step()
        opt.step_count = steps0
        self.launches_per_step = ops.LAUNCHES - launches0   # libb200ir kernels inside the graph(s)

    def __call__(self, degraded: torch.Tensor, clean: torch.Tensor) -> torch.Tensor:
        self.degraded.copy_(degraded, non_blocking=True)
        self.clean.copy_(clean, non_blocking=True)
        self.opt.sync_hyper()   # lr / weight decay changed by a scheduler since the last replay -> device scalars
        self.g1.replay()
        if self.g2 is not None:
            import torch.distributed as dist
            dist.all_reduce(self.opt.flat_g, op=dist.ReduceOp.SUM, group=self.group)
            self.g2.replay()
        self.opt.step_count += 1
        ops.LAUNCHES += self.launches_per_step
        A.bump_param_epoch()   # parameters changed under the packed-weight caches
        return self.loss


class GraphedForward:
    """Inference counterpart of GraphedTrainStep: the forward of ``net`` for one fixed input shape captured in a CUDA
    graph (the ~320 launches of a Restormer forward become one graph launch; measured 30.1 -> 29.6 ms at 16 x 256x256).
    ``out = fwd(x)`` copies x into the static input buffer, replays, and returns the static output tensor (overwritten
    by the next call).  Parameters are read in place; after changing them call ``refresh()`` (packed bf16 weight caches
    were baked into the graph)."""

    def __init__(self, net: nn.Module, example: torch.Tensor, warmup: int = 2):
        self.net = net
        self.x = example.detach().clone()
        self._capture(warmup)

    def _capture(self, warmup):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):
                self.net(self.x)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        launches0 = ops.LAUNCHES
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.y = self.net(self.x)
        self.launches_per_call = ops.LAUNCHES - launches0

    def refresh(self):
        self._capture(1)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if x.shape != self.x.shape:
            raise ValueError(f"GraphedForward was captured for {tuple(self.x.shape)}, got {tuple(x.shape)}")
        self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        ops.LAUNCHES += self.launches_per_call
        return self.y


class PipelinedForward:
    """Serving loop over HOST buffers: ``submit(host_in, host_out)`` enqueues the pinned-host -> device copy of a batch on
    a copy stream, the forward (one CUDA-graph replay, ``GraphedForward``) on the current stream and the device -> pinned-host
    copy of the restored batch on a second copy stream, so that the copies of batches i+1 and i-1 overlap the forward of
    batch i.  ``depth`` input / output staging buffers on the device; results are complete after ``synchronize()`` (or
    after waiting on the event ``submit`` returns).  The reference's serving loop is test.py's
    ``restored = net(degrad_patch)`` over a DataLoader (MoCE-IR-main/src/test.py:95-103) with blocking ``.cuda()`` / ``.cpu()``."""

    def __init__(self, net: nn.Module, example: torch.Tensor, depth: int = 2):
        self.fwd = GraphedForward(net, example)
        self.depth = depth
        self.h2d, self.d2h = torch.cuda.Stream(), torch.cuda.Stream()
        self.xin = [torch.empty_like(self.fwd.x) for _ in range(depth)]
        self.yout = [torch.empty_like(self.fwd.y) for _ in range(depth)]
        mk = lambda: [torch.cuda.Event() for _ in range(depth)]   # noqa: E731
        self.ev_in, self.ev_in_free, self.ev_out, self.ev_out_free = mk(), mk(), mk(), mk()
        self.i = 0

    def submit(self, host_in: torch.Tensor, host_out: torch.Tensor):
        if host_in.shape != self.fwd.x.shape or host_out.shape != self.fwd.y.shape:
            raise ValueError("PipelinedForward: host buffers must have the captured input / output shapes")
        k, first = self.i % self.depth, self.i < self.depth
        main = torch.cuda.current_stream()
        if not first:
            self.h2d.wait_event(self.ev_in_free[k])          # the forward that read xin[k] has copied it out
        with torch.cuda.stream(self.h2d):
            self.xin[k].copy_(host_in, non_blocking=True)
            self.ev_in[k].record(self.h2d)
        main.wait_event(self.ev_in[k])
        self.fwd.x.copy_(self.xin[k], non_blocking=True)
        self.ev_in_free[k].record(main)
        self.fwd.graph.replay()
        ops.LAUNCHES += self.fwd.launches_per_call
        if not first:
            main.wait_event(self.ev_out_free[k])             # the device -> host copy that read yout[k] is done
        self.yout[k].copy_(self.fwd.y, non_blocking=True)
        self.ev_out[k].record(main)
        self.d2h.wait_event(self.ev_out[k])
        with torch.cuda.stream(self.d2h):
            host_out.copy_(self.yout[k], non_blocking=True)
            self.ev_out_free[k].record(self.d2h)
        self.i += 1
        return self.ev_out_free[k]

    def join(self):
        """make the current stream wait for every copy enqueued so far (a timing event recorded after it covers them)"""
        main = torch.cuda.current_stream()
        main.wait_stream(self.h2d)
        main.wait_stream(self.d2h)

    def synchronize(self):
        self.join()
        torch.cuda.current_stream().synchronize()

