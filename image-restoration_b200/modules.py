"""Drop-in mirror of the reference's nn.Module surface for the Restormer block path.

Same class names, constructor signatures, attribute / parameter names and shapes as
``/root/reference/Restormer.py`` (LayerNorm :60-70, FeedForward :76-93, Attention :99-132,
TransformerBlock :137-150, OverlapPatchEmbed :156-165, Downsample :171-179, Upsample :181-189,
Restormer :193-284), so ``load_state_dict(reference.state_dict(), strict=True)`` is exact and the
reference's train.py / test.py run unchanged on top.  ``forward`` launches hand-written sm_100a
kernels through the C ABI (``ops``); there is no PyTorch/CPU fallback.

Compute dtype = dtype of the activation tensor handed in (float32: exact CUDA-core path;
bfloat16: bf16 storage, fp32 accumulation/statistics, tcgen05 GEMMs).  ``Restormer`` takes the
image in float32 like the reference and runs its trunk in ``self.compute_dtype``.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from . import _lib as L
from . import ops
from . import autograd_ops as A


def _ln_args(norm: "LayerNorm"):
    body = norm.body  # ours, or a reference LayerNorm body adopted by dropin.accelerate(): told apart by its parameters
    if getattr(body, "bias", None) is not None:
        return L.LN_WITHBIAS, body.weight, body.bias
    return L.LN_BIASFREE, body.weight, None


##########################################################################
## Layer Norm  (Restormer.py:25-70)
class BiasFree_LayerNorm(nn.Module):
    def __init__(self, normalized_shape):
        super().__init__()
        if isinstance(normalized_shape, int):
            normalized_shape = (normalized_shape,)
        normalized_shape = torch.Size(normalized_shape)
        assert len(normalized_shape) == 1
        self.weight = nn.Parameter(torch.ones(normalized_shape))
        self.normalized_shape = normalized_shape


class WithBias_LayerNorm(nn.Module):
    def __init__(self, normalized_shape):
        super().__init__()
        if isinstance(normalized_shape, int):
            normalized_shape = (normalized_shape,)
        normalized_shape = torch.Size(normalized_shape)
        assert len(normalized_shape) == 1
        self.weight = nn.Parameter(torch.ones(normalized_shape))
        self.bias = nn.Parameter(torch.zeros(normalized_shape))
        self.normalized_shape = normalized_shape


class LayerNorm(nn.Module):
    """Channel LayerNorm over an NCHW tensor (Restormer.py:60-70).  Inside TransformerBlock it is
    fused into the following 1x1 conv; called on its own it is one pass of the stand-alone LayerNorm kernel
    (b200ir_ln_fwd; backward b200ir_ln_bwd)."""

    def __init__(self, dim, LayerNorm_type):
        super().__init__()
        if LayerNorm_type == 'BiasFree':
            self.body = BiasFree_LayerNorm(dim)
        else:
            self.body = WithBias_LayerNorm(dim)

    def forward(self, x):
        ln_type, w, b = _ln_args(self)
        return A.layernorm(x, ln_type, w, b)


##########################################################################
## Gated-Dconv Feed-Forward Network (GDFN)  (Restormer.py:76-93)
class FeedForward(nn.Module):
    def __init__(self, dim, ffn_expansion_factor, bias):
        super().__init__()
        hidden_features = int(dim * ffn_expansion_factor)
        self.project_in = nn.Conv2d(dim, hidden_features * 2, kernel_size=1, bias=bias)
        self.dwconv = nn.Conv2d(hidden_features * 2, hidden_features * 2, kernel_size=3, stride=1, padding=1,
                                groups=hidden_features * 2, bias=bias)
        self.project_out = nn.Conv2d(hidden_features, dim, kernel_size=1, bias=bias)

    def _run(self, x, norm=None, res=None, out=None):
        ln_type, lw, lb = _ln_args(norm) if norm is not None else (L.LN_NONE, None, None)
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        hd = self.project_out.in_channels
        g = A.fused_ln_proj_dw(x, self.project_in.weight, self.project_in.bias, self.dwconv.weight, self.dwconv.bias,
                               ln_type, lw, lb, hd, True, cache=cache, key=("fin", id(norm)))
        if g is None:   # training / fp32 / shapes the fused kernel does not take: two kernels, 2*hd channels through HBM
            t = A.ln_conv1x1(x, self.project_in.weight, self.project_in.bias, ln_type, lw, lb, None,
                             cache=cache, key=("in", id(norm)))
            g = A.dwconv3x3_gelu_gate(t, self.dwconv.weight, self.dwconv.bias)
            del t
        return A.ln_conv1x1(g, self.project_out.weight, self.project_out.bias, L.LN_NONE, None, None, res, out=out,
                            cache=cache, key="out")

    def forward(self, x):
        return self._run(x)


##########################################################################
## Multi-DConv Head Transposed Self-Attention (MDTA)  (Restormer.py:99-132)
class Attention(nn.Module):
    def __init__(self, dim, num_heads, bias):
        super().__init__()
        self.num_heads = num_heads
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.qkv = nn.Conv2d(dim, dim * 3, kernel_size=1, bias=bias)
        self.qkv_dwconv = nn.Conv2d(dim * 3, dim * 3, kernel_size=3, stride=1, padding=1, groups=dim * 3, bias=bias)
        self.project_out = nn.Conv2d(dim, dim, kernel_size=1, bias=bias)

    def _run(self, x, norm=None, res=None, out=None):
        ln_type, lw, lb = _ln_args(norm) if norm is not None else (L.LN_NONE, None, None)
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        t = A.ln_conv1x1(x, self.qkv.weight, self.qkv.bias, ln_type, lw, lb, None, cache=cache, key=("qkv", id(norm)))
        qkv = A.dwconv3x3(t, self.qkv_dwconv.weight, self.qkv_dwconv.bias)
        del t
        return A.mdta_core(qkv, self.temperature, self.project_out.weight, self.project_out.bias, self.num_heads,
                           res, out=out, cache=cache)

    def forward(self, x):
        return self._run(x)


##########################################################################
## Gated MDTA of the reference's gated-attention experiment (experiment/baseline_restormer_new_train/
## 20251116_use_headwise_gate_model/MoCE-IR-main/src/net/moce_ir.py:84-171)
class GatedMDTA(nn.Module):
    """MDTA whose qkv projection carries gate channels after q (``[q | gate | k | v]``; ``num_heads`` channels for
    gate_type='headwise', ``dim`` for 'elementwise', none for None) and whose attention output is multiplied by
    sigmoid(gate) BEFORE project_out.  The gate sits between attn.v and project_out, so project_out cannot be folded
    into the attention matrix: attn.v comes from the same Gram / softmax+fold / per-sample GEMM kernels with an
    identity fold, then the sigmoid-gate kernel (csrc/gate.cu), then project_out (+ bias + residual) as one more GEMM."""

    def __init__(self, dim, num_heads, bias, gate_type=None):
        super().__init__()
        if gate_type not in (None, "headwise", "elementwise"):
            raise ValueError(f"Unsupported gate_type: {gate_type}, choose from None, 'headwise', 'elementwise'")
        self.num_heads, self.gate_type, self.dim, self.bias = num_heads, gate_type, dim, bias
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.qkv_out_channels = dim * 3 + {None: 0, "headwise": num_heads, "elementwise": dim}[gate_type]
        self.qkv = nn.Conv2d(dim, self.qkv_out_channels, kernel_size=1, bias=bias)
        self.qkv_dwconv = nn.Conv2d(self.qkv_out_channels, self.qkv_out_channels, kernel_size=3, stride=1, padding=1,
                                    groups=self.qkv_out_channels, bias=bias)
        self.project_out = nn.Conv2d(dim, dim, kernel_size=1, bias=bias)

    def _eye(self, device):
        e = self.__dict__.get("_b200ir_eye")
        if e is None or e.device != device:
            e = torch.eye(self.dim, device=device, dtype=torch.float32).view(self.dim, self.dim, 1, 1)
            self.__dict__["_b200ir_eye"] = e
        return e

    def _run(self, x, norm=None, res=None, out=None):
        ln_type, lw, lb = _ln_args(norm) if norm is not None else (L.LN_NONE, None, None)
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        C, G = self.dim, self.qkv_out_channels - 3 * self.dim
        t = A.ln_conv1x1(x, self.qkv.weight, self.qkv.bias, ln_type, lw, lb, None, cache=cache, key=("qkv", id(norm)))
        wd, bd = self.qkv_dwconv.weight, self.qkv_dwconv.bias
        if G == 0:
            qkv = A.dwconv3x3(t, wd, bd)
            del t
            return A.mdta_core(qkv, self.temperature, self.project_out.weight, self.project_out.bias, self.num_heads,
                               res, out=out, cache=cache)
        sl = lambda a, b: (t[:, a:b], wd[a:b], bd[a:b] if bd is not None else None)
        if A._needs_grad(t, wd, bd):
            q = A.dwconv3x3(*sl(0, C))
            gate = A.dwconv3x3(*sl(C, C + G))
            kv = A.dwconv3x3(*sl(C + G, 3 * C + G))
            qkv = torch.cat([q, kv], dim=1)
            del q, kv
        else:
            # inference: the q and k|v depthwise launches write their channel slices of ONE [q|k|v] buffer (no torch.cat)
            qkv = torch.empty((t.shape[0], 3 * C, t.shape[2], t.shape[3]), device=t.device, dtype=t.dtype)
            for (a, b), dst in (((0, C), qkv[:, :C]), ((C + G, 3 * C + G), qkv[:, C:])):
                xs, ws, bs = sl(a, b)
                ops.dwconv3x3(xs, A._p(ws), A._p(bs), out=dst)
            xs, ws, bs = sl(C, C + G)
            gate = ops.dwconv3x3(xs, A._p(ws), A._p(bs))
        del t
        u = A.mdta_core(qkv, self.temperature, self._eye(x.device), None, self.num_heads, None,
                        cache=cache)                      # attn . v (identity fold)
        del qkv
        ug = A.sigmoid_gate(u, gate)                      # * sigmoid(gate), head-wise or element-wise
        del u, gate
        return A.ln_conv1x1(ug, self.project_out.weight, self.project_out.bias, L.LN_NONE, None, None, res, out=out,
                            cache=cache, key="out")

    def forward(self, x):
        return self._run(x)


##########################################################################
## RectiFormer experiment variants (experiment/RectiFormer.py:231-325)
class GatedSpatialFFN(nn.Module):
    """Split-gate FFN: project_in -> chunk -> gelu(dwconv_content(content)) * sigmoid(dwconv_gate(gate)) -> project_out.
    The two half-width depthwise convs are one launch (their weights form the [content | gate] weight of the same
    kernel that runs GDFN, with the gate function gelu(a) * sigmoid(b))."""

    def __init__(self, dim, ffn_expansion_factor, bias):
        super().__init__()
        hidden_features = int(dim * ffn_expansion_factor)
        self.project_in = nn.Conv2d(dim, hidden_features * 2, kernel_size=1, bias=bias)
        self.dwconv_gate = nn.Conv2d(hidden_features, hidden_features, kernel_size=3, stride=1, padding=1,
                                     groups=hidden_features, bias=bias)
        self.dwconv_content = nn.Conv2d(hidden_features, hidden_features, kernel_size=3, stride=1, padding=1,
                                        groups=hidden_features, bias=bias)
        self.project_out = nn.Conv2d(hidden_features, dim, kernel_size=1, bias=bias)

    def _dw_pair(self):
        ps = (self.dwconv_content.weight, self.dwconv_gate.weight, self.dwconv_content.bias, self.dwconv_gate.bias)
        sig = A._sig(*ps)
        ent = self.__dict__.get("_b200ir_dw")
        if ent is None or ent[0] != sig:
            w = torch.cat([ps[0].detach(), ps[1].detach()], 0).float().contiguous()
            b = torch.cat([ps[2].detach(), ps[3].detach()], 0).float().contiguous() if ps[2] is not None else None
            ent = (sig, w, b)
            self.__dict__["_b200ir_dw"] = ent
        return ent[1], ent[2]

    def _run(self, x, norm=None, res=None, out=None):
        ln_type, lw, lb = _ln_args(norm) if norm is not None else (L.LN_NONE, None, None)
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        t = A.ln_conv1x1(x, self.project_in.weight, self.project_in.bias, ln_type, lw, lb, None, cache=cache,
                         key=("in", id(norm)))
        ps = (self.dwconv_content.weight, self.dwconv_gate.weight, self.dwconv_content.bias, self.dwconv_gate.bias)
        if A._needs_grad(t, *ps):
            # training: the stacked [content | gate] depthwise weight stays connected to its two parameters
            w = torch.cat([ps[0], ps[1]], 0)
            b = torch.cat([ps[2], ps[3]], 0) if ps[2] is not None else None
            g = A.dwconv3x3_gelu_gate(t, w, b, sigmoid_gate=True)
        else:
            w, b = self._dw_pair()
            g = ops.dwconv3x3_gelu_sigmoid_gate(t, w, b)
        del t
        return A.ln_conv1x1(g, self.project_out.weight, self.project_out.bias, L.LN_NONE, None, None, res, out=out,
                            cache=cache, key="out")

    def forward(self, x):
        return self._run(x)


class ElementwiseGatedAttention(nn.Module):
    """MDTA with an element-wise sigmoid gate from its own branch (depthwise 3x3 -> 1x1 on x) between attn.v and
    project_out, and an optional per-sample dynamic temperature (experiment/RectiFormer.py:270-325)."""

    def __init__(self, dim, num_heads, bias):
        super().__init__()
        self.num_heads = num_heads
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.dim = dim
        self.qkv = nn.Conv2d(dim, dim * 3, kernel_size=1, bias=bias)
        self.qkv_dwconv = nn.Conv2d(dim * 3, dim * 3, kernel_size=3, stride=1, padding=1, groups=dim * 3, bias=bias)
        self.gate_generator = nn.Sequential(
            nn.Conv2d(dim, dim, kernel_size=3, stride=1, padding=1, groups=dim, bias=bias),
            nn.Conv2d(dim, dim, kernel_size=1, bias=bias))
        self.project_out = nn.Conv2d(dim, dim, kernel_size=1, bias=bias)

    def _eye(self, device):
        e = self.__dict__.get("_b200ir_eye")
        if e is None or e.device != device:
            e = torch.eye(self.dim, device=device, dtype=torch.float32).view(self.dim, self.dim, 1, 1)
            self.__dict__["_b200ir_eye"] = e
        return e

    def forward(self, x, dynamic_temp=None, res=None):
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        t = A.ln_conv1x1(x, self.qkv.weight, self.qkv.bias, L.LN_NONE, None, None, None, cache=cache, key="qkv")
        qkv = A.dwconv3x3(t, self.qkv_dwconv.weight, self.qkv_dwconv.bias)
        del t
        gd = A.dwconv3x3(x, self.gate_generator[0].weight, self.gate_generator[0].bias)
        gate = A.ln_conv1x1(gd, self.gate_generator[1].weight, self.gate_generator[1].bias, L.LN_NONE, None, None, None,
                            cache=cache, key="gate")
        u = A.mdta_core(qkv, self.temperature, self._eye(x.device), None, self.num_heads, None, cache=cache,
                        dynamic_temp=dynamic_temp)                     # attn . v (identity fold)
        ug = A.sigmoid_gate(u, gate)
        return A.ln_conv1x1(ug, self.project_out.weight, self.project_out.bias, L.LN_NONE, None, None, res,
                            cache=cache, key="out")


class GatedTransformerBlock(nn.Module):
    """x = x + attn(norm1(x)); x = x + ffn(norm2(x)) with the gated attention (same file, :180-193)."""

    def __init__(self, dim, num_heads, ffn_expansion_factor, bias, LayerNorm_type, gate_type=None):
        super().__init__()
        self.norm1 = LayerNorm(dim, LayerNorm_type)
        self.attn = GatedMDTA(dim, num_heads, bias, gate_type=gate_type)
        self.norm2 = LayerNorm(dim, LayerNorm_type)
        self.ffn = FeedForward(dim, ffn_expansion_factor, bias)

    def forward(self, x, out=None):
        x = self.attn._run(x, norm=self.norm1, res=x)
        return self.ffn._run(x, norm=self.norm2, res=x, out=out)


##########################################################################
## MoCE-IR decoder mixer (moce_ir.py:325-368): MDTA with q from one tensor and k, v from another.
class CrossAttention(nn.Module):
    """q = q_dwconv(q(x)) (1x1 + depthwise 3x3), k, v = kv_dwconv(kv(y)).chunk(2) (1x1 + depthwise 7x7); the two
    depthwise kernels write the channel slices [0,C) and [C,3C) of ONE buffer, which is exactly the qkv layout the MDTA
    kernels (Gram+norms, softmax+fold, attn.v+project_out) already take.  With gradients the two halves are produced by
    the autograd functions and concatenated (one extra copy) before the same MDTA core."""

    def __init__(self, dim, num_heads, bias):
        super().__init__()
        self.num_heads = num_heads
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.q = nn.Conv2d(dim, dim, kernel_size=1, bias=bias)
        self.q_dwconv = nn.Conv2d(dim, dim, kernel_size=3, stride=1, padding=1, groups=dim, bias=bias)
        self.kv = nn.Conv2d(dim, dim * 2, kernel_size=1, bias=bias)
        self.kv_dwconv = nn.Conv2d(dim * 2, dim * 2, kernel_size=7, stride=1, padding=7 // 2, groups=dim * 2, bias=bias)
        self.project_out = nn.Conv2d(dim, dim, kernel_size=1, bias=bias)

    def forward(self, x, y, res=None):
        """`res` (optional, not in the reference signature): added to the output inside the last GEMM's epilogue
        (DecoderBlock's `mixer(x_a, x_s) + shortcut`, moce_ir.py:892)."""
        B, C, H, W = x.shape
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        tq = A.ln_conv1x1(x, self.q.weight, self.q.bias, L.LN_NONE, None, None, None, cache=cache, key="q")
        tkv = A.ln_conv1x1(y, self.kv.weight, self.kv.bias, L.LN_NONE, None, None, None, cache=cache, key="kv")
        if torch.is_grad_enabled() and (tq.requires_grad or tkv.requires_grad):
            qkv = torch.cat([A.dwconv3x3(tq, self.q_dwconv.weight, self.q_dwconv.bias),
                             A.dwconv_kxk(tkv, self.kv_dwconv.weight, self.kv_dwconv.bias)], dim=1)
        else:
            qkv = torch.empty((B, 3 * C, H, W), device=x.device, dtype=tq.dtype)
            ops.dwconv3x3(tq, A._p(self.q_dwconv.weight), A._p(self.q_dwconv.bias), out=qkv[:, :C])
            ops.dwconv_kxk(tkv, A._p(self.kv_dwconv.weight), A._p(self.kv_dwconv.bias), out=qkv[:, C:])
        del tq, tkv
        return A.mdta_core(qkv, self.temperature, self.project_out.weight, self.project_out.bias, self.num_heads,
                           res, cache=cache)


##########################################################################
class TransformerBlock(nn.Module):
    """x = x + attn(norm1(x)); x = x + ffn(norm2(x))   (Restormer.py:137-150), as 8 kernel launches:
    LN+qkv | dw3x3 | Gram+norms | softmax+fold | attn.v+proj+res | LN+project_in | dw3x3+GELU gate |
    project_out+res."""

    def __init__(self, dim, num_heads, ffn_expansion_factor, bias, LayerNorm_type):
        super().__init__()
        self.norm1 = LayerNorm(dim, LayerNorm_type)
        self.attn = Attention(dim, num_heads, bias)
        self.norm2 = LayerNorm(dim, LayerNorm_type)
        self.ffn = FeedForward(dim, ffn_expansion_factor, bias)

    def forward(self, x, out=None):
        x = self.attn._run(x, norm=self.norm1, res=x)
        x = self.ffn._run(x, norm=self.norm2, res=x, out=out)
        return x


##########################################################################
## MoCE-IR flavour of the same block (moce_ir.py:805-834, 903-932): identical math, different attribute names
## (`norms.0`, `norms.1`, `mixer`, `ffn`) — those names are the checkpoint contract of MoCE-IR models.
class EncoderBlock(nn.Module):
    def __init__(self, dim, num_heads, ffn_expansion_factor, bias, LayerNorm_type):
        super().__init__()
        self.norms = nn.ModuleList([LayerNorm(dim, LayerNorm_type), LayerNorm(dim, LayerNorm_type)])
        self.mixer = Attention(dim, num_heads, bias)
        self.ffn = FeedForward(dim, ffn_expansion_factor, bias)

    def forward(self, x, out=None):
        x = self.mixer._run(x, norm=self.norms[0], res=x)            # x + mixer(norms[0](x))   moce_ir.py:826
        return self.ffn._run(x, norm=self.norms[1], res=x, out=out)  # x + ffn(norms[1](x))     moce_ir.py:830


class EncoderResidualGroup(nn.Module):
    def __init__(self, dim, num_heads, num_blocks, ffn_expansion, LayerNorm_type, bias):
        super().__init__()
        self.loss = None
        self.num_blocks = num_blocks
        self.layers = nn.ModuleList([EncoderBlock(dim, num_heads, ffn_expansion, bias, LayerNorm_type)
                                     for _ in range(num_blocks)])

    def forward(self, x):
        self.loss = 0                                                # moce_ir.py:926 (no auxiliary loss in the encoder)
        for blk in self.layers:
            x = blk(x)
        return x


##########################################################################
## Overlapped image patch embedding with 3x3 Conv  (Restormer.py:156-165)
class OverlapPatchEmbed(nn.Module):
    def __init__(self, in_c=3, embed_dim=48, bias=False):
        super().__init__()
        self.proj = nn.Conv2d(in_c, embed_dim, kernel_size=3, stride=1, padding=1, bias=bias)

    def forward(self, x, out_dtype=None):
        return A.conv3x3(x, self.proj.weight, self.proj.bias, None, L.SHUFFLE_NONE, out_dtype=out_dtype)


##########################################################################
## Resizing modules  (Restormer.py:171-189)
class Downsample(nn.Module):
    def __init__(self, n_feat):
        super().__init__()
        self.body = nn.Sequential(nn.Conv2d(n_feat, n_feat // 2, kernel_size=3, stride=1, padding=1, bias=False),
                                  nn.PixelUnshuffle(2))

    def forward(self, x, out=None):
        return A.conv3x3(x, self.body[0].weight, None, None, L.UNSHUFFLE2, out=out,
                         cache=self.__dict__.setdefault("_b200ir_cache", {}), key="w")


class Upsample(nn.Module):
    def __init__(self, n_feat):
        super().__init__()
        self.body = nn.Sequential(nn.Conv2d(n_feat, n_feat * 2, kernel_size=3, stride=1, padding=1, bias=False),
                                  nn.PixelShuffle(2))

    def forward(self, x, out=None):
        return A.conv3x3(x, self.body[0].weight, None, None, L.SHUFFLE2, out=out,
                         cache=self.__dict__.setdefault("_b200ir_cache", {}), key="w")


##########################################################################
##---------- Restormer -----------------------  (Restormer.py:193-284)
def _default_compute_dtype():
    v = os.environ.get("B200IR_COMPUTE_DTYPE", "fp32").lower()
    return torch.bfloat16 if v in ("bf16", "bfloat16") else torch.float32


class Restormer(nn.Module):
    def __init__(self,
                 inp_channels=3,
                 out_channels=3,
                 dim=48,
                 num_blocks=[4, 6, 6, 8],
                 num_refinement_blocks=4,
                 heads=[1, 2, 4, 8],
                 ffn_expansion_factor=2.66,
                 bias=False,
                 LayerNorm_type='WithBias',
                 dual_pixel_task=False):
        super().__init__()

        def stack(d, h, n):
            return nn.Sequential(*[TransformerBlock(dim=d, num_heads=h, ffn_expansion_factor=ffn_expansion_factor,
                                                    bias=bias, LayerNorm_type=LayerNorm_type) for _ in range(n)])

        self.patch_embed = OverlapPatchEmbed(inp_channels, dim)
        self.encoder_level1 = stack(dim, heads[0], num_blocks[0])
        self.down1_2 = Downsample(dim)
        self.encoder_level2 = stack(int(dim * 2 ** 1), heads[1], num_blocks[1])
        self.down2_3 = Downsample(int(dim * 2 ** 1))
        self.encoder_level3 = stack(int(dim * 2 ** 2), heads[2], num_blocks[2])
        self.down3_4 = Downsample(int(dim * 2 ** 2))
        self.latent = stack(int(dim * 2 ** 3), heads[3], num_blocks[3])
        self.up4_3 = Upsample(int(dim * 2 ** 3))
        self.reduce_chan_level3 = nn.Conv2d(int(dim * 2 ** 3), int(dim * 2 ** 2), kernel_size=1, bias=bias)
        self.decoder_level3 = stack(int(dim * 2 ** 2), heads[2], num_blocks[2])
        self.up3_2 = Upsample(int(dim * 2 ** 2))
        self.reduce_chan_level2 = nn.Conv2d(int(dim * 2 ** 2), int(dim * 2 ** 1), kernel_size=1, bias=bias)
        self.decoder_level2 = stack(int(dim * 2 ** 1), heads[1], num_blocks[1])
        self.up2_1 = Upsample(int(dim * 2 ** 1))
        self.decoder_level1 = stack(int(dim * 2 ** 1), heads[0], num_blocks[0])
        self.refinement = stack(int(dim * 2 ** 1), heads[0], num_refinement_blocks)
        self.dual_pixel_task = dual_pixel_task
        if self.dual_pixel_task:
            self.skip_conv = nn.Conv2d(dim, int(dim * 2 ** 1), kernel_size=1, bias=bias)
        self.output = nn.Conv2d(int(dim * 2 ** 1), out_channels, kernel_size=3, stride=1, padding=1, bias=bias)

        self.compute_dtype = _default_compute_dtype()  # not a parameter/buffer: state_dict stays the reference's

    def set_compute_dtype(self, dtype):
        if dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("compute dtype must be torch.float32 or torch.bfloat16")
        self.compute_dtype = dtype
        return self

    @staticmethod
    def _run_stack(seq, x, out_last=None):
        n = len(seq)
        for i, blk in enumerate(seq):
            x = blk(x, out=out_last if i == n - 1 else None)
        return x

    def forward(self, inp_img):
        if inp_img.dim() != 4:
            raise ValueError("expected NCHW input")
        if not inp_img.is_cuda:
            raise ValueError("b200ir.Restormer runs on CUDA only (no CPU fallback)")
        B, _, H, W = inp_img.shape
        if H % 8 or W % 8:
            raise ValueError(f"H and W must be multiples of 8 (three PixelUnshuffle(2) levels), got {H}x{W}")
        inp_img = inp_img.contiguous()
        cd = self.compute_dtype
        grad = torch.is_grad_enabled() and (inp_img.requires_grad or any(p.requires_grad for p in self.parameters()))
        dev = inp_img.device
        dim = self.patch_embed.proj.out_channels

        # skip tensors are produced directly inside their concat buffers (no torch.cat copy) when no
        # autograd graph is needed; with autograd the plain cat is used.
        def cat_buf(c_up, c_skip, h, w):
            return None if grad else torch.empty((B, c_up + c_skip, h, w), device=dev, dtype=cd)

        cat1 = cat_buf(dim, dim, H, W)
        cat2 = cat_buf(dim * 2, dim * 2, H // 2, W // 2)
        cat3 = cat_buf(dim * 4, dim * 4, H // 4, W // 4)

        inp_enc_level1 = self.patch_embed(inp_img, out_dtype=cd)
        out_enc_level1 = self._run_stack(self.encoder_level1, inp_enc_level1,
                                         None if grad or len(self.encoder_level1) == 0 else cat1[:, dim:])
        inp_enc_level2 = self.down1_2(out_enc_level1)
        out_enc_level2 = self._run_stack(self.encoder_level2, inp_enc_level2,
                                         None if grad or len(self.encoder_level2) == 0 else cat2[:, dim * 2:])
        inp_enc_level3 = self.down2_3(out_enc_level2)
        out_enc_level3 = self._run_stack(self.encoder_level3, inp_enc_level3,
                                         None if grad or len(self.encoder_level3) == 0 else cat3[:, dim * 4:])
        inp_enc_level4 = self.down3_4(out_enc_level3)
        latent = self._mix(1, inp_img, self._run_stack(self.latent, inp_enc_level4))

        inp_dec_level3 = self._cat(self.up4_3, latent, out_enc_level3, cat3, dim * 4)
        cache = self.__dict__.setdefault("_b200ir_cache", {})
        inp_dec_level3 = A.ln_conv1x1(inp_dec_level3, self.reduce_chan_level3.weight, self.reduce_chan_level3.bias,
                                      L.LN_NONE, None, None, None, cache=cache, key="rc3")
        out_dec_level3 = self._mix(2, inp_img, self._run_stack(self.decoder_level3, inp_dec_level3))

        inp_dec_level2 = self._cat(self.up3_2, out_dec_level3, out_enc_level2, cat2, dim * 2)
        inp_dec_level2 = A.ln_conv1x1(inp_dec_level2, self.reduce_chan_level2.weight, self.reduce_chan_level2.bias,
                                      L.LN_NONE, None, None, None, cache=cache, key="rc2")
        out_dec_level2 = self._mix(3, inp_img, self._run_stack(self.decoder_level2, inp_dec_level2))

        inp_dec_level1 = self._cat(self.up2_1, out_dec_level2, out_enc_level1, cat1, dim)
        out_dec_level1 = self._run_stack(self.decoder_level1, inp_dec_level1)
        out_dec_level1 = self._run_stack(self.refinement, out_dec_level1)

        if self.dual_pixel_task:
            out_dec_level1 = A.ln_conv1x1(inp_enc_level1, self.skip_conv.weight, self.skip_conv.bias, L.LN_NONE, None,
                                          None, out_dec_level1, cache=cache, key="skip")
            return A.conv3x3(out_dec_level1, self.output.weight, self.output.bias, None, L.SHUFFLE_NONE,
                             out_dtype=inp_img.dtype)
        return A.conv3x3(out_dec_level1, self.output.weight, self.output.bias, inp_img, L.SHUFFLE_NONE,
                         out_dtype=inp_img.dtype)

    def _mix(self, stage, inp_img, feat):
        """Hook after the latent (1), decoder_level3 (2) and decoder_level2 (3) stacks: identity in Restormer; AdaIR's
        frequency modules sit here (AdaIR-main/net/model.py:463-485)."""
        return feat

    @staticmethod
    def _cat(up, x, skip, buf, c_up):
        """torch.cat([up(x), skip], 1) (Restormer.py:259-260,264-265,269-270) without the copy when possible."""
        if buf is None:
            return torch.cat([up(x), skip], 1)
        up(x, out=buf[:, :c_up])
        if skip.data_ptr() != buf[:, c_up:].data_ptr():  # skip was not produced in place (empty stack)
            ops.copy_cast(skip, buf[:, c_up:])
        return buf
