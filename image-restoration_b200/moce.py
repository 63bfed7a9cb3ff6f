"""Drop-in mirror of the reference's MoCE-IR model (``/root/reference/moce_ir.py``; SURVEY.md §8 f1 and BASELINE
configs[2]): same class names, constructor signatures, attribute / parameter / buffer names and shapes, so
``load_state_dict(reference.state_dict(), strict=True)`` is exact (Lightning checkpoints add a ``net.`` prefix,
MoCE-IR-main/src/train.py:33).  Every tensor-sized operation runs in kernels of libb200ir.so: the Restormer block pieces
(``modules.py``), the two-source MDTA of the decoder, depthwise 3x3 / 7x7, and ``csrc/moce_ops.cu`` for the FFT expert
(per-patch circular convolution instead of rfft2 * rfft2 -> irfft2), the gates and the global pools.  Only the router's
arithmetic on the ``[batch, num_experts]`` logits (softmax, top-k, scatter: a few dozen numbers) is plain torch.

Quirks of the reference that are kept on purpose (§3.4 of SURVEY.md):
  * routing noise is added in eval mode too (moce_ir.py:741) — ``MoCEIR.noise_device`` selects the generator it is drawn
    from ("cpu" reproduces a CPU run of the reference bit for bit in its expert choice; default = the activations' device);
  * ``ModExpert.feat_extract`` applies ``process`` ``depth`` times to the SAME input (:566-568): one evaluation;
  * eval-mode ``AdapterLayer`` with batch > 1 and top-k = 1 runs, for every sample, the experts chosen by ALL samples on
    the whole batch and scales their SUM by the sample's own gate (:668-676, ``top_k_indices.squeeze(0)``).
Training mode (round 2): ``net.train()`` switches the adapter to the reference's SparseDispatcher semantics (:71-141,
659-667: per-expert sample gather, gate-weighted scatter-accumulate) and the router to its auxiliary importance + load
losses (:733-800); ``net.total_loss`` is the balance loss of MoCE-IR-main/src/train.py:62.  Every tensor-sized forward
and backward op is a libb200ir kernel behind a ``torch.autograd.Function`` (``autograd_ops.py``); pinned to outputs,
losses and all parameter gradients of the unmodified reference (``oracle/gen_golden_moceir_train.py``).
"""
from __future__ import annotations

import math
from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L
from . import autograd_ops as A
from . import ops
from .modules import (Attention, CrossAttention, Downsample, EncoderResidualGroup, FeedForward, LayerNorm,
                      OverlapPatchEmbed, Upsample, _ln_args)


def _conv1x1(x, conv: nn.Conv2d, res=None, cache=None, key=None, out=None):
    return A.ln_conv1x1(x, conv.weight, conv.bias, L.LN_NONE, None, None, res, out=out, cache=cache, key=key)


def _linear(x, lin: nn.Linear):
    """fp32 [B, in] -> [B, out] through the 1x1-conv kernel (a [B, C] matrix is an NCHW tensor with one pixel)."""
    y = ops.ln_conv1x1(x.float().contiguous().view(x.shape[0], x.shape[1], 1, 1), A._p(lin.weight), A._p(lin.bias))
    return y.view(x.shape[0], -1)


class MySequential(nn.Sequential):
    """nn.Sequential whose layers take two inputs (moce_ir.py:31-49)."""

    def forward(self, x1, x2):
        for layer in self:
            x1 = layer(x1, x2)
        return x1


class HighPassConv2d(nn.Module):
    """Fixed 3x3 high-pass depthwise kernel (moce_ir.py:224-253); the weight is a parameter of the checkpoint."""

    def __init__(self, c, freeze):
        super().__init__()
        self.conv = nn.Conv2d(in_channels=c, out_channels=c, kernel_size=3, padding=1, bias=False, groups=c)
        kernel = torch.tensor([[[[-1, -1, -1], [-1, 8, -1], [-1, -1, -1]]]], dtype=torch.float32)
        self.conv.weight.data = kernel.repeat(c, 1, 1, 1)
        if freeze:
            self.conv.requires_grad_ = False   # (sic: the reference sets an attribute, the weight stays trainable)

    def forward(self, x):
        return A.dwconv3x3(x, self.conv.weight, None)


class FrequencyEmbedding(nn.Module):
    """high-pass dw3x3 -> GELU -> global mean -> MLP (moce_ir.py:1048-1075); GELU is folded into the pooling kernel."""

    def __init__(self, dim):
        super().__init__()
        self.high_conv = nn.Sequential(HighPassConv2d(dim, freeze=True), nn.GELU())
        self.mlp = nn.Sequential(nn.Linear(dim, 2 * dim), nn.GELU(), nn.Linear(2 * dim, dim))

    def forward(self, x):
        h = self.high_conv[0](x)
        if A._needs_grad(h, *self.mlp.parameters()):
            # training: the pool is a kernel with a backward kernel; the MLP works on [B, dim] vectors (a few hundred
            # numbers) and goes through torch so that autograd sees it
            m = A.spatial_mean(h, gelu=True)
            return self.mlp(m)
        m = ops.spatial_mean(h, gelu=True)                       # [B, dim] fp32 = gelu(h).mean((-2, -1))
        m = _linear(m, self.mlp[0])
        m = ops.ew(m.contiguous(), None, ops.EW_GELU)
        return _linear(m, self.mlp[2])


class FFTAttention(nn.Module):
    """The adapter's expert body (moce_ir.py:373-422): q = dw3x3(1x1(x)), k, v = dw7x7(1x1(x)); per patch
    irfft2(rfft2(q) * rfft2(k)) = circular convolution of q and k; LayerNorm; * v; 1x1."""

    def __init__(self, dim: int, **kwargs):
        super().__init__()
        self.patch_size = kwargs["patch_size"]
        self.q = nn.Conv2d(dim, dim, kernel_size=1, bias=False)
        self.q_dwconv = nn.Conv2d(dim, dim, kernel_size=3, stride=1, padding=1, groups=dim)
        self.kv = nn.Conv2d(dim, dim * 2, kernel_size=1, bias=False)
        self.kv_dwconv = nn.Conv2d(dim * 2, dim * 2, kernel_size=7, stride=1, padding=7 // 2, groups=dim * 2)
        self.norm = LayerNorm(dim, "WithBias")
        self.proj_out = nn.Conv2d(dim, dim, kernel_size=1, padding=0)

    def forward(self, x):
        c = x.shape[1]
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        q = A.dwconv3x3(_conv1x1(x, self.q, cache=cache, key="q"), self.q_dwconv.weight, self.q_dwconv.bias)
        kv = A.dwconv_kxk(_conv1x1(x, self.kv, cache=cache, key="kv"), self.kv_dwconv.weight, self.kv_dwconv.bias)
        out = A.circconv_patch(q, kv[:, :c], self.patch_size)
        ln_type, lw, lb = _ln_args(self.norm)
        out = A.ew_mul(A.layernorm(out, ln_type, lw, lb), kv[:, c:])
        return _conv1x1(out, self.proj_out, cache=cache, key="out")


class ModExpert(nn.Module):
    """Low-rank expert (moce_ir.py:520-582): down-project, FFT body gated by silu(proj(shared)), up-project + shortcut."""

    def __init__(self, dim: int, rank: int, func: nn.Module, depth: int, patch_size: int, kernel_size: int):
        super().__init__()
        self.depth = depth
        self.proj = nn.ModuleList([nn.Conv2d(dim, rank, kernel_size=1, padding=0, bias=False),
                                   nn.Conv2d(dim, rank, kernel_size=1, padding=0, bias=False),
                                   nn.Conv2d(rank, dim, kernel_size=1, padding=0, bias=False)])
        self.body = func(rank, kernel_size=kernel_size, patch_size=patch_size)

    def forward(self, x, shared):
        if x.shape[0] == 0:
            return x
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        t = self.body(_conv1x1(x, self.proj[0], cache=cache, key="p0"))
        t = A.ew_silu_mul(t, _conv1x1(shared, self.proj[1], cache=cache, key="p1"))
        return _conv1x1(t, self.proj[2], res=x, cache=cache, key="p2")      # `depth` repeats give the same tensor


class RoutingFunction(nn.Module):
    """Noisy top-k router over pooled features + the frequency embedding (moce_ir.py:684-800)."""

    def __init__(self, dim, freq_dim, num_experts, k, complexity, use_complexity_bias: bool = True,
                 complexity_scale: str = "max"):
        super().__init__()
        # index 1 of the reference is einops' Rearrange('b c 1 1 -> b c') (parameter-free): Identity keeps `gate.2.weight`
        self.gate = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Identity(), nn.Linear(dim, num_experts, bias=False))
        self.freq_gate = nn.Linear(freq_dim, num_experts, bias=False)
        if complexity_scale == "min":
            complexity = complexity / complexity.min()
        elif complexity_scale == "max":
            complexity = complexity / complexity.max()
        self.register_buffer('complexity', complexity)
        self.k = k
        self.tau = 1
        self.num_experts = num_experts
        self.noise_std = (1.0 / num_experts) * 1.0
        self.use_complexity_bias = use_complexity_bias
        self.noise_device = None

    def forward(self, x, freq_emb):
        if self.training:
            return self._forward_train(x, freq_emb)
        logits = _linear(ops.spatial_mean(x), self.gate[2]) + _linear(freq_emb, self.freq_gate)
        nd = self.noise_device or logits.device
        noise = torch.randn(logits.shape, dtype=logits.dtype, device=nd).to(logits.device) * self.noise_std
        gating_scores = (logits + noise).softmax(dim=-1)
        top_k_values, top_k_indices = torch.topk(gating_scores, self.k, dim=-1)
        gates = torch.zeros_like(logits).scatter_(1, top_k_indices, top_k_values)
        return gates, top_k_indices, top_k_values, 0

    def _forward_train(self, x, freq_emb):
        """Training mode (moce_ir.py:733-758): the same noisy top-k gates plus the auxiliary loss
        0.5 * importance + 0.5 * load.  The global pool is a kernel (with a backward kernel); everything after it works on
        the [batch, num_experts] logits and stays in torch so that autograd differentiates it like the reference's."""
        pooled = A.spatial_mean(x)                                      # AdaptiveAvgPool2d(1), fp32 [B, dim]
        logits = F.linear(pooled, self.gate[2].weight.float()) + F.linear(freq_emb.float(), self.freq_gate.weight.float())
        loss_imp = self.importance_loss(logits.softmax(dim=-1))
        nd = self.noise_device or logits.device
        noise = torch.randn(logits.shape, dtype=logits.dtype, device=nd).to(logits.device) * self.noise_std
        noisy_logits = logits + noise
        gating_scores = noisy_logits.softmax(dim=-1)
        top_k_values, top_k_indices = torch.topk(gating_scores, self.k, dim=-1)
        loss_load = self.load_loss(logits, noisy_logits, self.noise_std)
        aux_loss = 0.5 * loss_imp + 0.5 * loss_load
        gates = torch.zeros_like(logits).scatter(1, top_k_indices, top_k_values)
        return gates, top_k_indices, top_k_values, aux_loss

    def importance_loss(self, gating_scores):
        """squared coefficient of variation of the per-expert importance (moce_ir.py:760-769)"""
        importance = gating_scores.sum(dim=0)
        importance = importance * (self.complexity * self.tau) if self.use_complexity_bias else importance
        return (importance.std() / (importance.mean() + 1e-8)) ** 2

    def load_loss(self, logits, logits_noisy, noise_std):
        """squared coefficient of variation of the probability that an expert is in the top-k (moce_ir.py:771-800)"""
        thresholds = torch.topk(logits_noisy, self.k, dim=-1).indices[:, -1]
        threshold_per_item = torch.sum(F.one_hot(thresholds, self.num_experts) * logits_noisy, dim=-1)
        noise_required_to_win = (threshold_per_item.unsqueeze(-1) - logits) / noise_std
        p = 1. - torch.distributions.normal.Normal(0, 1).cdf(noise_required_to_win)
        p_mean = p.mean(dim=0)
        return (p_mean.std() / (p_mean.mean() + 1e-8)) ** 2


class AdapterLayer(nn.Module):
    """Mixture of complexity experts (moce_ir.py:584-682): eval-mode semantics in ``forward`` (see the module docstring),
    SparseDispatcher semantics in training mode (``_forward_train``)."""

    def __init__(self, dim: int, rank: int, num_experts: int = 4, top_k: int = 2, expert_layer: nn.Module = FFTAttention,
                 stage_depth: int = 1, depth_type: str = "lin", rank_type: str = "constant", freq_dim: int = 128,
                 with_complexity: bool = False, complexity_scale: str = "min"):
        super().__init__()
        self.tau = 1
        self.loss = None
        self.top_k = top_k
        self.noise_eps = 1e-2
        self.num_experts = num_experts
        patch_sizes = [2 ** (i + 2) for i in range(num_experts)]
        kernel_sizes = [3 + (2 * i) for i in range(num_experts)]
        if depth_type == "lin":
            depths = [stage_depth + i for i in range(num_experts)]
        elif depth_type == "double":
            depths = [stage_depth + (2 * i) for i in range(num_experts)]
        elif depth_type == "exp":
            depths = [2 ** (i) for i in range(num_experts)]
        elif depth_type == "fact":
            depths = [math.factorial(i + 1) for i in range(num_experts)]
        elif isinstance(depth_type, int):
            depths = [depth_type for _ in range(num_experts)]
        elif depth_type == "constant":
            depths = [stage_depth for i in range(num_experts)]
        else:
            raise NotImplementedError
        if rank_type == "constant":
            ranks = [rank for _ in range(num_experts)]
        elif rank_type == "lin":
            ranks = [rank + i for i in range(num_experts)]
        elif rank_type == "double":
            ranks = [rank + (2 * i) for i in range(num_experts)]
        elif rank_type == "exp":
            ranks = [rank ** (i + 1) for i in range(num_experts)]
        elif rank_type == "fact":
            ranks = [math.factorial(rank + i) for i in range(num_experts)]
        elif rank_type == "spread":
            ranks = [dim // (2 ** i) for i in range(num_experts)][::-1]
        else:
            raise NotImplementedError
        self.experts = nn.ModuleList([
            MySequential(*[ModExpert(dim, rank=r, func=expert_layer, depth=d, patch_size=p, kernel_size=k)])
            for (d, r, p, k) in zip(depths, ranks, patch_sizes, kernel_sizes)])
        self.proj_out = nn.Conv2d(dim, dim, kernel_size=1, padding=0, bias=False)
        expert_complexity = torch.tensor([sum(p.numel() for p in expert.parameters()) for expert in self.experts])
        self.routing = RoutingFunction(dim, freq_dim, num_experts=num_experts, k=top_k, complexity=expert_complexity,
                                       use_complexity_bias=with_complexity, complexity_scale=complexity_scale)

    def route(self, x, freq_emb):
        """Run the router and start the copy of its decision (top-k indices | gates > 0 per expert) to pinned host memory.
        A caller that has other work for the GPU which does not depend on the routing (DecoderBlock: the `shared`
        self-attention) enqueues it between route() and forward(..., routed=...): the host then waits only for this small
        copy, not for the whole stream, and the expert kernels are launched while the GPU is still busy."""
        gates, top_k_indices, _, aux = self.routing(x, freq_emb)
        self.loss = aux
        dec = torch.cat([top_k_indices.to(torch.int32), (gates.detach() > 0).to(torch.int32)], 1)
        host = self.__dict__.get("_b200ir_pin")
        if host is None or host.shape != dec.shape:
            host = self.__dict__["_b200ir_pin"] = torch.empty(dec.shape, dtype=torch.int32).pin_memory()
        host.copy_(dec, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        return gates, top_k_indices, host, ev

    def forward(self, x, freq_emb, shared, routed=None):
        if routed is None:
            routed = self.route(x, freq_emb)
        if self.training:
            return self._forward_train(x, shared, routed)
        gates, top_k_indices, host, ev = routed
        sel = top_k_indices.squeeze(0)
        if sel.dim() != 1 and not (sel.dim() == 2 and sel.shape[1] == 1):
            raise ValueError("eval-mode AdapterLayer supports batch 1 (any top-k) or top-k 1 (any batch), like the reference")
        ev.synchronize()                         # the routing result is on the host (the reference indexes with it too)
        chosen: List[int] = [int(i) for i in host[:, :self.top_k].flatten().tolist()]
        g = gates.gather(1, top_k_indices).float().contiguous()              # [B, k]
        out = None
        if sel.dim() == 1:                       # batch 1: expert j scaled by its own gate
            for j, e in enumerate(chosen):
                y = self.experts[e](x, shared)
                out = ops.scale_rows(y, g[:, j].contiguous(), out=out, accumulate=out is not None)
        else:                                    # batch > 1, k = 1: sum over every sample's expert, own gate per sample
            counts = {}
            for e in chosen:
                counts[e] = counts.get(e, 0) + 1
            for e, cnt in counts.items():
                y = self.experts[e](x, shared)
                out = ops.scale_rows(y, (g[:, 0] * float(cnt)).contiguous(), out=out, accumulate=out is not None)
        return _conv1x1(out, self.proj_out, cache=self.__dict__.setdefault("_b200ir_cache", {}), key="out")


def _adapter_forward_train(self, x, shared, routed):
    """Training mode (moce_ir.py:659-667): SparseDispatcher semantics.  Expert e receives the samples b with
    gates[b, e] > 0 (``dispatch``: a row gather), and the output is out[b] = sum_e gates[b, e] * expert_e(x[b], shared[b])
    (``combine``: a gate-weighted scatter-accumulate, experts in ascending order like the reference's index_add over the
    expert-sorted batch index).  Experts that receive no sample are not run and get no gradient, as in the reference."""
    gates, top_k_indices, host, ev = routed
    B = x.shape[0]
    ev.synchronize()                                            # like the reference's `.tolist()` of the part sizes
    sel_host = host[:, self.top_k:] > 0
    x, shared = x.contiguous(), shared.contiguous()
    ys, gs, idxs = [], [], []
    for e in range(self.num_experts):
        rows = torch.nonzero(sel_host[:, e]).flatten()
        n = int(rows.numel())
        if n == 0:
            continue
        idx = rows.to(device=x.device, dtype=torch.int32)
        if n == B:
            xe, se = x, shared
        else:
            xe, se = A.gather_rows(x, idx), A.gather_rows(shared, idx)
        ys.append(self.experts[e](xe, se))
        gs.append(gates[rows.to(x.device), e])
        idxs.append(idx)
    out = A.combine_rows(B, ys, gs, idxs)
    return _conv1x1(out, self.proj_out, cache=self.__dict__.setdefault("_b200ir_cache", {}), key="out")


AdapterLayer._forward_train = _adapter_forward_train


class DecoderBlock(nn.Module):
    """moce_ir.py:839-897: ONE LayerNorm feeds two 1x1 projections (run as one GEMM with 2*dim outputs), `shared`
    self-attention, the expert adapter, the cross-attention mixer (+ shortcut) and the GDFN."""

    def __init__(self, dim, num_heads, ffn_expansion_factor, bias, LayerNorm_type, expert_layer, complexity_scale=None,
                 rank=None, num_experts=None, top_k=None, depth_type=None, rank_type=None, stage_depth=None,
                 freq_dim: int = 128, with_complexity: bool = False):
        super().__init__()
        self.norms = nn.ModuleList([LayerNorm(dim, LayerNorm_type), LayerNorm(dim, LayerNorm_type)])
        self.proj = nn.ModuleList([nn.Conv2d(dim, dim, kernel_size=1, padding=0),
                                   nn.Conv2d(dim, dim, kernel_size=1, padding=0)])
        self.shared = Attention(dim, num_heads, bias)
        self.mixer = CrossAttention(dim, num_heads=num_heads, bias=bias)
        self.ffn = FeedForward(dim, ffn_expansion_factor, bias)
        self.adapter = AdapterLayer(dim, rank, top_k=top_k, num_experts=num_experts, expert_layer=expert_layer,
                                    freq_dim=freq_dim, depth_type=depth_type, rank_type=rank_type,
                                    stage_depth=stage_depth, with_complexity=with_complexity,
                                    complexity_scale=complexity_scale)

    def _proj_pair(self):
        """proj[0] | proj[1] as one [2*dim, dim] weight (rebuilt when a parameter changes)."""
        ps = (self.proj[0].weight, self.proj[1].weight, self.proj[0].bias, self.proj[1].bias)
        sig = A._sig(*ps)
        ent = self.__dict__.get("_b200ir_pair")
        if ent is None or ent[0] != sig:
            w = torch.cat([ps[0].detach(), ps[1].detach()], 0).contiguous()
            b = torch.cat([ps[2].detach(), ps[3].detach()], 0).contiguous()
            ent = (sig, w, b)
            self.__dict__["_b200ir_pair"] = ent
        return ent[1], ent[2]

    def forward(self, x, freq_emb=None):
        dim = x.shape[1]
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        ln_type, lw, lb = _ln_args(self.norms[0])
        if A._needs_grad(self.proj[0].weight, self.proj[1].weight, self.proj[0].bias, self.proj[1].bias):
            # training: the stacked weight has to stay connected to its two parameters
            w2 = torch.cat([self.proj[0].weight, self.proj[1].weight], 0)
            b2 = torch.cat([self.proj[0].bias, self.proj[1].bias], 0)
            both = A.ln_conv1x1(x, w2, b2, ln_type, lw, lb, None)
        else:
            w2, b2 = self._proj_pair()
            both = A.ln_conv1x1(x, w2, b2, ln_type, lw, lb, None, cache=cache, key="pair")     # [B, 2*dim, H, W]
        routed = self.adapter.route(both[:, dim:], freq_emb)     # the router does not need x_s: decide first, so that
        x_s = self.shared._run(both[:, :dim])                      # the host has the decision while the GPU runs `shared`
        x_a = self.adapter(both[:, dim:], freq_emb, x_s, routed=routed)
        x = self.mixer(x_a, x_s, res=x)
        x = self.ffn._run(x, norm=self.norms[1], res=x)
        return x, self.adapter.loss


class DecoderResidualGroup(nn.Module):
    def __init__(self, dim: int, num_heads: List[int], num_blocks: int, ffn_expansion: int, LayerNorm_type: str,
                 bias: bool, complexity_scale=None, rank=None, num_experts=None, expert_layer=None, top_k=None,
                 depth_type=None, stage_depth=None, rank_type=None, freq_dim: int = 128, with_complexity: bool = False):
        super().__init__()
        self.loss = None
        self.num_blocks = num_blocks
        self.layers = nn.ModuleList([
            DecoderBlock(dim, num_heads, ffn_expansion, bias, LayerNorm_type, expert_layer=expert_layer, rank=rank,
                         num_experts=num_experts, top_k=top_k, stage_depth=stage_depth, freq_dim=freq_dim,
                         complexity_scale=complexity_scale, depth_type=depth_type, rank_type=rank_type,
                         with_complexity=with_complexity) for _ in range(num_blocks)])

    def forward(self, x, freq_emb=None):
        self.loss = 0
        for blk in self.layers:
            x, loss = blk(x, freq_emb)
            self.loss += loss
        return x


class MoCEIR(nn.Module):
    """moce_ir.py:1080-1231.  Activations run in ``compute_dtype`` (fp32 = exact CUDA-core path, bf16 = tensor-core path),
    the image goes in and comes out in float32 like the reference."""

    def __init__(self, inp_channels=3, out_channels=3, dim=32, levels: int = 4, heads=[1, 1, 1, 1],
                 num_blocks=[1, 1, 1, 3], num_dec_blocks=[1, 1, 1], ffn_expansion_factor=2, num_refinement_blocks=1,
                 LayerNorm_type='WithBias', bias=False, rank=2, num_experts=4, depth_type="lin", stage_depth=[3, 2, 1],
                 rank_type="constant", topk=1, expert_layer=FFTAttention, with_complexity=False, complexity_scale="max"):
        super().__init__()
        self.levels = levels
        self.num_blocks = num_blocks
        self.num_dec_blocks = num_dec_blocks
        self.num_refinement_blocks = num_refinement_blocks
        dims = [dim * 2 ** i for i in range(levels)]
        ranks = [rank for i in range(levels - 1)]
        self.patch_embed = OverlapPatchEmbed(in_c=inp_channels, embed_dim=dim, bias=False)
        self.freq_embed = FrequencyEmbedding(dims[-1])
        self.enc = nn.ModuleList([])
        for i in range(levels - 1):
            self.enc.append(nn.ModuleList([
                EncoderResidualGroup(dim=dims[i], num_blocks=num_blocks[i], num_heads=heads[i],
                                     ffn_expansion=ffn_expansion_factor, LayerNorm_type=LayerNorm_type, bias=True),
                Downsample(dim * 2 ** i)]))
        self.latent = EncoderResidualGroup(dim=dims[-1], num_blocks=num_blocks[-1], num_heads=heads[-1],
                                           ffn_expansion=ffn_expansion_factor, LayerNorm_type=LayerNorm_type, bias=True)
        dims = dims[::-1]
        ranks = ranks[::-1]
        heads = heads[::-1]
        num_dec_blocks = num_dec_blocks[::-1]
        self.dec = nn.ModuleList([])
        for i in range(levels - 1):
            self.dec.append(nn.ModuleList([
                Upsample(dims[i]),
                nn.Conv2d(dims[i], dims[i + 1], kernel_size=1, bias=bias),
                DecoderResidualGroup(dim=dims[i + 1], num_blocks=num_dec_blocks[i], num_heads=heads[i + 1],
                                     ffn_expansion=ffn_expansion_factor, LayerNorm_type=LayerNorm_type, bias=bias,
                                     expert_layer=expert_layer, freq_dim=dims[0], with_complexity=with_complexity,
                                     rank=ranks[i], num_experts=num_experts, stage_depth=stage_depth[i],
                                     depth_type=depth_type, rank_type=rank_type, top_k=topk,
                                     complexity_scale=complexity_scale)]))
        heads = heads[::-1]
        self.refinement = EncoderResidualGroup(dim=dim, num_blocks=num_refinement_blocks, num_heads=heads[0],
                                               ffn_expansion=ffn_expansion_factor, LayerNorm_type=LayerNorm_type,
                                               bias=True)
        self.output = nn.Conv2d(dim, out_channels, kernel_size=3, stride=1, padding=1, bias=bias)
        self.total_loss = None
        from .modules import _default_compute_dtype
        self.compute_dtype = _default_compute_dtype()   # not a parameter / buffer: state_dict stays the reference's

    def set_compute_dtype(self, dtype):
        if dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("compute dtype must be torch.float32 or torch.bfloat16")
        self.compute_dtype = dtype
        return self

    def set_noise_device(self, device):
        """Generator the router's eval-time noise is drawn from: "cpu" follows a CPU run of the reference under the same
        torch.manual_seed; None = the activations' device (what the reference does on that device)."""
        for m in self.modules():
            if isinstance(m, RoutingFunction):
                m.noise_device = device
        return self

    def forward(self, x, labels=None):
        if x.dim() != 4 or not x.is_cuda:
            raise ValueError("b200ir.MoCEIR takes a CUDA NCHW tensor (no CPU fallback)")
        H, W = x.shape[-2:]
        if H % 2 ** (self.levels - 1) or W % 2 ** (self.levels - 1):
            raise ValueError(f"H and W must be multiples of {2 ** (self.levels - 1)}, got {H}x{W}")
        grad = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters()))
        x = x.contiguous()
        cd = self.compute_dtype
        feats = self.patch_embed(x, out_dtype=cd)
        self.total_loss = 0
        enc_feats = []
        for block, downsample in self.enc:
            feats = block(feats)
            enc_feats.append(feats)
            feats = downsample(feats)
        feats = self.latent(feats)
        freq_emb = self.freq_embed(feats)
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        for i, (upsample, fusion, block) in enumerate(self.dec):
            skip = enc_feats.pop()
            B, cs, h, w = skip.shape
            if grad:
                buf = torch.cat([upsample(feats), skip], 1)                     # autograd path (moce_ir.py:1219)
            else:
                buf = torch.empty((B, 2 * cs, h, w), device=x.device, dtype=cd)  # the same without a cat kernel
                upsample(feats, out=buf[:, :cs])
                ops.copy_cast(skip, buf[:, cs:])
            feats = _conv1x1(buf, fusion, cache=cache, key=("fuse", i))
            feats = block(feats, freq_emb)
            self.total_loss += block.loss
        feats = self.refinement(feats)
        out = A.conv3x3(feats, self.output.weight, self.output.bias, x, L.SHUFFLE_NONE, out_dtype=x.dtype)
        self.total_loss /= sum(self.num_dec_blocks)
        return out
