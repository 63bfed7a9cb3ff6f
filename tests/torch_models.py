"""TEST INFRASTRUCTURE: plain-PyTorch nn.Modules shaped like the reference's classes (same attribute / parameter names
as Restormer.py:60-150 and moce_ir.py:805-834), whose forwards call the CPU oracle's functional restatement.  They stand
in for "an existing reference model" in the tests of b200ir.accelerate(); nothing in the product imports this file."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import restormer_oracle as O


class _Body(nn.Module):
    def __init__(self, dim, with_bias):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        if with_bias:
            self.bias = nn.Parameter(torch.zeros(dim))


class RefLayerNorm(nn.Module):
    def __init__(self, dim, LayerNorm_type):
        super().__init__()
        self.body = _Body(dim, LayerNorm_type != "BiasFree")

    def forward(self, x):
        return O.layernorm(x, self.body.weight, getattr(self.body, "bias", None))


class RefFeedForward(nn.Module):
    def __init__(self, dim, f, bias):
        super().__init__()
        hd = int(dim * f)
        self.project_in = nn.Conv2d(dim, hd * 2, 1, bias=bias)
        self.dwconv = nn.Conv2d(hd * 2, hd * 2, 3, 1, 1, groups=hd * 2, bias=bias)
        self.project_out = nn.Conv2d(hd, dim, 1, bias=bias)

    def forward(self, x):
        return O.feedforward(x, dict(self.named_parameters()), "")


class RefAttention(nn.Module):
    def __init__(self, dim, num_heads, bias):
        super().__init__()
        self.num_heads = num_heads
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.qkv = nn.Conv2d(dim, dim * 3, 1, bias=bias)
        self.qkv_dwconv = nn.Conv2d(dim * 3, dim * 3, 3, 1, 1, groups=dim * 3, bias=bias)
        self.project_out = nn.Conv2d(dim, dim, 1, bias=bias)

    def forward(self, x):
        return O.attention(x, dict(self.named_parameters()), "", self.num_heads)


class RefGatedAttention(RefAttention):
    """Same attribute names, different forward (sigmoid output gate, as the experiment/*gate* variants do):
    accelerate(verify=True) must leave it alone."""

    def forward(self, x):
        return torch.sigmoid(x) * super().forward(x)


class RefTransformerBlock(nn.Module):
    def __init__(self, dim, heads, f, bias, ln, attn_cls=RefAttention):
        super().__init__()
        self.norm1 = RefLayerNorm(dim, ln)
        self.attn = attn_cls(dim, heads, bias)
        self.norm2 = RefLayerNorm(dim, ln)
        self.ffn = RefFeedForward(dim, f, bias)

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        return x + self.ffn(self.norm2(x))


class RefEncoderBlock(nn.Module):
    def __init__(self, dim, heads, f, bias, ln):
        super().__init__()
        self.norms = nn.ModuleList([RefLayerNorm(dim, ln), RefLayerNorm(dim, ln)])
        self.mixer = RefAttention(dim, heads, bias)
        self.ffn = RefFeedForward(dim, f, bias)

    def forward(self, x):
        x = x + self.mixer(self.norms[0](x))
        return x + self.ffn(self.norms[1](x))


class RefDown(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.body = nn.Sequential(nn.Conv2d(n, n // 2, 3, 1, 1, bias=False), nn.PixelUnshuffle(2))

    def forward(self, x):
        return self.body(x)


class RefUp(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.body = nn.Sequential(nn.Conv2d(n, n * 2, 3, 1, 1, bias=False), nn.PixelShuffle(2))

    def forward(self, x):
        return self.body(x)


class RefMiniNet(nn.Module):
    """A small U-shaped model mixing every pattern accelerate() knows, a decoder that uses Attention / FeedForward /
    LayerNorm separately (like moce_ir.DecoderBlock) and one gated variant."""

    def __init__(self, dim=16):
        super().__init__()
        self.embed = nn.Conv2d(3, dim, 3, 1, 1, bias=False)
        self.enc = nn.Sequential(RefTransformerBlock(dim, 1, 2.66, False, "WithBias"),
                                 RefEncoderBlock(dim, 2, 2.0, True, "BiasFree"))
        self.down = RefDown(dim)
        self.mid = nn.ModuleList([RefEncoderBlock(dim * 2, 2, 2.0, True, "WithBias")])
        self.gated = RefTransformerBlock(dim * 2, 2, 2.0, False, "WithBias", attn_cls=RefGatedAttention)
        self.up = RefUp(dim * 2)
        self.dec_norm = RefLayerNorm(dim, "WithBias")
        self.dec_attn = RefAttention(dim, 1, True)
        self.dec_ffn = RefFeedForward(dim, 2.0, True)
        self.out = nn.Conv2d(dim, 3, 3, 1, 1, bias=True)

    def forward(self, img):
        x = self.enc(self.embed(img))
        y = self.down(x)
        for m in self.mid:
            y = m(y)
        y = self.gated(y)
        x = x + self.up(y)
        n = self.dec_norm(x)
        x = x + self.dec_attn(n)
        x = x + self.dec_ffn(n)
        return self.out(x) + img


class RefCrossAttention(nn.Module):
    """Stand-in with the attribute names / shapes of MoCE-IR's CrossAttention (moce_ir.py:325-343); forward = the oracle."""

    def __init__(self, dim, num_heads, bias, kv_kernel=7):
        super().__init__()
        self.num_heads = num_heads
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.q = nn.Conv2d(dim, dim, 1, bias=bias)
        self.q_dwconv = nn.Conv2d(dim, dim, 3, 1, 1, groups=dim, bias=bias)
        self.kv = nn.Conv2d(dim, dim * 2, 1, bias=bias)
        self.kv_dwconv = nn.Conv2d(dim * 2, dim * 2, kv_kernel, 1, kv_kernel // 2, groups=dim * 2, bias=bias)
        self.project_out = nn.Conv2d(dim, dim, 1, bias=bias)

    def forward(self, x, y):
        return O.cross_attention(x, y, dict(self.named_parameters()), "", self.num_heads)


class RefDecoderMixerNet(nn.Module):
    """Two branches mixed by a cross attention plus a residual, the way moce_ir.DecoderBlock.forward uses it (:894)."""

    def __init__(self, dim=32, heads=2):
        super().__init__()
        self.a = nn.Conv2d(dim, dim, 1)
        self.b = nn.Conv2d(dim, dim, 1)
        self.mixer = RefCrossAttention(dim, heads, True)

    def forward(self, x):
        return self.mixer(self.a(x), self.b(x)) + x


class RefGatedMDTA(nn.Module):
    """Stand-in with the structure of the gated-attention experiment's GatedMDTA; forward = the oracle."""

    def __init__(self, dim, num_heads, bias, gate_type):
        super().__init__()
        self.num_heads, self.gate_type = num_heads, gate_type
        ct = dim * 3 + {None: 0, "headwise": num_heads, "elementwise": dim}[gate_type]
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.qkv = nn.Conv2d(dim, ct, 1, bias=bias)
        self.qkv_dwconv = nn.Conv2d(ct, ct, 3, 1, 1, groups=ct, bias=bias)
        self.project_out = nn.Conv2d(dim, dim, 1, bias=bias)

    def forward(self, x):
        return O.gated_attention(x, dict(self.named_parameters()), "", self.num_heads, self.gate_type)


class RefChanelCrossAttention(RefCrossAttention):
    """Stand-in with the structure of AdaIR's Chanel_Cross_Attention (AdaIR-main/net/model.py:177-216): CrossAttention
    with a 3x3 kv depthwise conv and the head count spelled ``num_head``."""

    def __init__(self, dim, num_head, bias):
        super().__init__(dim, num_head, bias, kv_kernel=3)
        del self.num_heads
        self.num_head = num_head

    def forward(self, x, y):
        return O.cross_attention(x, y, dict(self.named_parameters()), "", self.num_head)


class RefAdaIRFusion(nn.Module):
    """x -> cross attention between two projections of x, the way AdaIR's FreRefine-style fusion calls it."""

    def __init__(self, dim=32, heads=2):
        super().__init__()
        self.a = nn.Conv2d(dim, dim, 1)
        self.b = nn.Conv2d(dim, dim, 1)
        self.cca = RefChanelCrossAttention(dim, heads, True)

    def forward(self, x):
        return self.cca(self.a(x), self.b(x)) + x


class RefGatedSpatialFFN(nn.Module):
    """Plain-torch stand-in with the structure and arithmetic of RectiFormer's GatedSpatialFFN
    (experiment/RectiFormer.py:231-266): content, gate = project_in(x).chunk(2); gelu(dw_c(content)) * sigmoid(dw_g(gate))."""

    def __init__(self, dim, ffn_expansion_factor, bias):
        super().__init__()
        hid = int(dim * ffn_expansion_factor)
        self.project_in = nn.Conv2d(dim, hid * 2, 1, bias=bias)
        self.dwconv_gate = nn.Conv2d(hid, hid, 3, 1, 1, groups=hid, bias=bias)
        self.dwconv_content = nn.Conv2d(hid, hid, 3, 1, 1, groups=hid, bias=bias)
        self.project_out = nn.Conv2d(hid, dim, 1, bias=bias)

    def forward(self, x):
        content, gate = self.project_in(x).chunk(2, dim=1)
        return self.project_out(F.gelu(self.dwconv_content(content)) * torch.sigmoid(self.dwconv_gate(gate)))


class RefElementwiseGatedAttention(nn.Module):
    """Stand-in for RectiFormer's ElementwiseGatedAttention (experiment/RectiFormer.py:270-325): MDTA, output gated by
    sigmoid(gate_generator(x)) before project_out, optional per-sample dynamic temperature."""

    def __init__(self, dim, num_heads, bias):
        super().__init__()
        self.num_heads, self.dim = num_heads, dim
        self.temperature = nn.Parameter(torch.ones(num_heads, 1, 1))
        self.qkv = nn.Conv2d(dim, dim * 3, 1, bias=bias)
        self.qkv_dwconv = nn.Conv2d(dim * 3, dim * 3, 3, 1, 1, groups=dim * 3, bias=bias)
        self.gate_generator = nn.Sequential(nn.Conv2d(dim, dim, 3, 1, 1, groups=dim, bias=bias), nn.Conv2d(dim, dim, 1, bias=bias))
        self.project_out = nn.Conv2d(dim, dim, 1, bias=bias)

    def forward(self, x, dynamic_temp=None):
        b, c, h, w = x.shape
        q, k, v = self.qkv_dwconv(self.qkv(x)).chunk(3, dim=1)
        gate = self.gate_generator(x)
        sh = lambda t: t.reshape(b, self.num_heads, c // self.num_heads, h * w)
        q, k, v = F.normalize(sh(q), dim=-1), F.normalize(sh(k), dim=-1), sh(v)
        temp = self.temperature if dynamic_temp is None else self.temperature.unsqueeze(0) * dynamic_temp
        attn = ((q @ k.transpose(-2, -1)) * temp).softmax(dim=-1)
        out = (attn @ v).reshape(b, c, h, w) * torch.sigmoid(gate)
        return self.project_out(out)


class RefRectiBlock(nn.Module):
    """x + attn(norm1(x), dynamic_temp); x + ffn(norm2(x)) with the two RectiFormer variants (the way its block calls them)."""

    def __init__(self, dim=32, heads=2, f=2.0, bias=True):
        super().__init__()
        self.norm1, self.norm2 = RefLayerNorm(dim, "WithBias"), RefLayerNorm(dim, "WithBias")
        self.attn = RefElementwiseGatedAttention(dim, heads, bias)
        self.ffn = RefGatedSpatialFFN(dim, f, bias)

    def forward(self, x, dynamic_temp=None):
        x = x + self.attn(self.norm1(x), dynamic_temp)
        return x + self.ffn(self.norm2(x))


class _RefSpatialGate(nn.Module):
    def __init__(self):
        super().__init__()
        self.spatial = nn.Conv2d(2, 1, 7, padding=3, bias=False)

    def forward(self, x):
        pooled = torch.cat([x.amax(dim=1, keepdim=True), x.mean(dim=1, keepdim=True)], 1)
        return torch.sigmoid(self.spatial(pooled))


class _RefChannelGate(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.mlp = nn.Sequential(nn.Conv2d(dim, dim // 16, 1, bias=False), nn.ReLU(), nn.Conv2d(dim // 16, dim, 1, bias=False))

    def forward(self, x):
        return torch.sigmoid(self.mlp(x.mean(dim=(2, 3), keepdim=True)) + self.mlp(x.amax(dim=(2, 3), keepdim=True)))


class _RefFreRefine(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.SpatialGate = _RefSpatialGate()
        self.ChannelGate = _RefChannelGate(dim)
        self.proj = nn.Conv2d(dim, dim, 1)

    def forward(self, low, high):
        return self.proj(low * self.SpatialGate(high) + high * self.ChannelGate(low))


class RefFreModule(nn.Module):
    """Stand-in with the structure and arithmetic of AdaIR's FreModule (AdaIR-main/net/model.py:295-372): bilinear resize
    of the image, conv3x3, a centred spectrum rectangle of half size int(H // 128 * t) split off with fft2 / ifft2, three
    channel cross attentions, FreRefine, per-channel affine.  Checked against the reference in the build container
    (outputs and gradients equal to rounding)."""

    def __init__(self, dim, num_heads, bias, in_dim=3):
        super().__init__()
        self.conv = nn.Conv2d(in_dim, dim, 3, 1, 1, bias=False)
        self.conv1 = nn.Conv2d(in_dim, dim, 3, 1, 1, bias=False)
        self.score_gen = nn.Conv2d(2, 2, 7, padding=3)
        self.para1 = nn.Parameter(torch.zeros(dim, 1, 1))
        self.para2 = nn.Parameter(torch.ones(dim, 1, 1))
        self.channel_cross_l = RefChanelCrossAttention(dim, num_heads, bias)
        self.channel_cross_h = RefChanelCrossAttention(dim, num_heads, bias)
        self.channel_cross_agg = RefChanelCrossAttention(dim, num_heads, bias)
        self.frequency_refine = _RefFreRefine(dim)
        self.rate_conv = nn.Sequential(nn.Conv2d(dim, dim // 8, 1, bias=False), nn.GELU(), nn.Conv2d(dim // 8, 2, 1, bias=False))

    def bands(self, x, n=128):
        x = self.conv1(x)
        B, C, H, W = x.shape
        t = torch.sigmoid(self.rate_conv(x.mean(dim=(2, 3), keepdim=True)))
        keep = torch.zeros_like(x)
        for i in range(B):
            hh, ww = int(H // n * t[i, 0, 0, 0]), int(W // n * t[i, 1, 0, 0])
            keep[i, :, H // 2 - hh:H // 2 + hh, W // 2 - ww:W // 2 + ww] = 1
        spec = torch.roll(torch.fft.fft2(x, norm="forward"), (H // 2, W // 2), (2, 3))
        back = lambda s: torch.fft.ifft2(torch.roll(s, (-(H // 2), -(W // 2)), (2, 3)), norm="forward").abs()   # noqa: E731
        return back(spec * (1 - keep)), back(spec * keep)

    def forward(self, x, y):
        x = torch.nn.functional.interpolate(x, y.shape[-2:], mode="bilinear")
        high, low = self.bands(x)
        high, low = self.channel_cross_l(high, y), self.channel_cross_h(low, y)
        out = self.channel_cross_agg(y, self.frequency_refine(low, high))
        return out * self.para1 + y * self.para2


class RefAdaIRTail(nn.Module):
    """A feature extractor followed by a frequency module, the way AdaIR calls it (model.py:463-465)."""

    def __init__(self, dim=32, heads=2):
        super().__init__()
        self.embed = nn.Conv2d(3, dim, 3, 1, 1)
        self.fre = RefFreModule(dim, heads, True)

    def forward(self, img):
        return self.fre(img, self.embed(img))
