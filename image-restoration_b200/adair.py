"""AdaIR extras: mirrors of ``Chanel_Cross_Attention``, ``SpatialGate``, ``ChannelGate``, ``FreRefine`` and
``FreModule`` of ``/root/reference/AdaIR-main/net/model.py:177-372`` with the reference's parameter names, running on
libb200ir kernels (``csrc/adair_ops.cu`` + the two-source MDTA).  Inference path (AdaIR's TransformerBlocks themselves are
plain Restormer blocks and are taken by ``modules.TransformerBlock`` / ``dropin.accelerate``).  Under autograd the same
modules run through the Functions of ``autograd_ops`` (backward kernels: ``csrc/adair_bwd.cu``), as AdaIR-main/train.py
needs; only the degraded image itself (``FreModule.forward``'s ``x``) cannot require a gradient.

FreModule.fft keeps only the spectrum bins inside a centred rectangle of half-size (h_, w_) = int(H // 128 * t) as the
"low" band; it is synthesised here from those few bins directly instead of fft2 -> mask -> ifft2 (see adair_ops.cu).
The tiny tensors of the threshold / channel-gate MLPs ([B, dim] and smaller) go through the 1x1-conv kernel; their
ReLU / sigmoid / int() on a few dozen numbers are torch ops.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L
from . import autograd_ops as A
from . import ops
from .modules import CrossAttention, Restormer


def _mm(x2d, conv):
    """[B, Cin] fp32 through a bias-free 1x1 nn.Conv2d's weight -> [B, Cout] (under autograd: F.linear on these few
    numbers, so the MLP weights get their gradients from torch)."""
    if A._needs_grad(x2d, conv.weight, conv.bias):
        return F.linear(x2d.float(), conv.weight.view(conv.out_channels, conv.in_channels).float(),
                        None if conv.bias is None else conv.bias.float())
    w = A._p(conv.weight).reshape(conv.out_channels, conv.in_channels)
    y = ops.ln_conv1x1(x2d.float().contiguous().view(x2d.shape[0], x2d.shape[1], 1, 1), w, A._p(conv.bias))
    return y.view(x2d.shape[0], -1)


class Chanel_Cross_Attention(CrossAttention):
    """model.py:177-216: CrossAttention whose k, v depthwise conv is 3x3 and whose head count is spelled ``num_head``."""

    def __init__(self, dim, num_head, bias):
        super().__init__(dim, num_head, bias)
        self.num_head = num_head
        self.kv_dwconv = nn.Conv2d(dim * 2, dim * 2, kernel_size=3, stride=1, padding=1, groups=dim * 2, bias=bias)

    def forward(self, x, y, res=None):
        assert x.shape == y.shape, 'The shape of feature maps from image and features are not equal!'
        return super().forward(x, y, res=res)


class SpatialGate(nn.Module):
    def __init__(self):
        super().__init__()
        self.spatial = nn.Conv2d(2, 1, kernel_size=7, padding=3, bias=False)

    def forward(self, x):
        return ops.spatial_gate(x, A._p(self.spatial.weight))          # fp32 [B, H*W]


class ChannelGate(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.avg = nn.AdaptiveAvgPool2d((1, 1))
        self.max = nn.AdaptiveMaxPool2d((1, 1))
        self.mlp = nn.Sequential(nn.Conv2d(dim, dim // 16, 1, bias=False), nn.ReLU(), nn.Conv2d(dim // 16, dim, 1, bias=False))

    def _mlp(self, v):
        return _mm(torch.relu(_mm(v, self.mlp[0])), self.mlp[2])

    def forward(self, x):
        return torch.sigmoid(self._mlp(A.spatial_mean(x)) + self._mlp(A.spatial_max(x))).contiguous()       # fp32 [B, C]


class FreRefine(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.SpatialGate = SpatialGate()
        self.ChannelGate = ChannelGate(dim)
        self.proj = nn.Conv2d(dim, dim, kernel_size=1)

    def forward(self, low, high):
        out = A.fre_refine_gate(low, high, self.SpatialGate.spatial.weight, self.ChannelGate(low))
        return A.ln_conv1x1(out, self.proj.weight, self.proj.bias, L.LN_NONE, None, None, None,
                            cache=self.__dict__.setdefault("_b200ir_cache", {}), key="proj")


class FreModule(nn.Module):
    """Adaptive Frequency Learning Block (model.py:295-365)."""

    def __init__(self, dim, num_heads, bias, in_dim=3):
        super().__init__()
        self.conv = nn.Conv2d(in_dim, dim, kernel_size=3, stride=1, padding=1, bias=False)     # unused by forward (as in the reference)
        self.conv1 = nn.Conv2d(in_dim, dim, kernel_size=3, stride=1, padding=1, bias=False)
        self.score_gen = nn.Conv2d(2, 2, 7, padding=3)                                          # unused by forward
        self.para1 = nn.Parameter(torch.zeros(dim, 1, 1))
        self.para2 = nn.Parameter(torch.ones(dim, 1, 1))
        self.channel_cross_l = Chanel_Cross_Attention(dim, num_head=num_heads, bias=bias)
        self.channel_cross_h = Chanel_Cross_Attention(dim, num_head=num_heads, bias=bias)
        self.channel_cross_agg = Chanel_Cross_Attention(dim, num_head=num_heads, bias=bias)
        self.frequency_refine = FreRefine(dim)
        self.rate_conv = nn.Sequential(nn.Conv2d(dim, dim // 8, 1, bias=False), nn.GELU(), nn.Conv2d(dim // 8, 2, 1, bias=False))

    def fft(self, x, n=128, out_dtype=None):
        x = A.conv3x3(x, self.conv1.weight, None, None, L.SHUFFLE_NONE, out_dtype=out_dtype)
        B, C, h, w = x.shape
        with torch.no_grad():      # the thresholds pass through .int() (model.py:345-346): no gradient reaches rate_conv
            t = _mm(ops.ew(_mm(ops.spatial_mean(x), self.rate_conv[0]).contiguous(), None, ops.EW_GELU), self.rate_conv[2])
            thr = torch.sigmoid(t)                                                               # [B, 2]
            counts = torch.stack([(h // n * thr[:, 0]).int(), (w // n * thr[:, 1]).int()], 1).contiguous()
        return A.lowfreq_split(x, counts, 4 * (h // n) * (w // n))

    def forward(self, x, y):
        if torch.is_grad_enabled() and x.requires_grad:
            raise RuntimeError("b200ir.FreModule: no gradient with respect to the degraded image (bilinear-resize adjoint not built)")
        _, _, H, W = y.size()
        x = ops.bilinear_resize(x.contiguous(), H, W)
        high_feature, low_feature = self.fft(x, out_dtype=y.dtype)
        high_feature = self.channel_cross_l(high_feature, y)
        low_feature = self.channel_cross_h(low_feature, y)
        agg = self.frequency_refine(low_feature, high_feature)
        out = self.channel_cross_agg(y, agg)
        return A.affine2(out, y, self.para1, self.para2)


class AdaIR(Restormer):
    """AdaIR-main/net/model.py:375-497: the Restormer U-Net with a FreModule after the latent, decoder_level3 and
    decoder_level2 stacks (all three built with heads[2], as in the reference), same parameter names."""

    def __init__(self, inp_channels=3, out_channels=3, dim=48, num_blocks=[4, 6, 6, 8], num_refinement_blocks=4,
                 heads=[1, 2, 4, 8], ffn_expansion_factor=2.66, bias=False, LayerNorm_type='WithBias', decoder=True):
        super().__init__(inp_channels, out_channels, dim, num_blocks, num_refinement_blocks, heads, ffn_expansion_factor,
                         bias, LayerNorm_type, dual_pixel_task=False)
        self.decoder = decoder
        if self.decoder:
            self.fre1 = FreModule(dim * 2 ** 3, num_heads=heads[2], bias=bias)
            self.fre2 = FreModule(dim * 2 ** 2, num_heads=heads[2], bias=bias)
            self.fre3 = FreModule(dim * 2 ** 1, num_heads=heads[2], bias=bias)

    def _mix(self, stage, inp_img, feat):
        if not self.decoder:
            return feat
        return (self.fre1, self.fre2, self.fre3)[stage - 1](inp_img, feat)

    def forward(self, inp_img, noise_emb=None):
        return super().forward(inp_img)
