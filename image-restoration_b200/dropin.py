"""``accelerate(model)``: swap the Restormer-family pieces of an EXISTING reference model for the B200 kernels, in place.

The reference repo holds ~70 copies of the same block under different wrappers (Restormer.py, moce_ir.py / MoCE-IR-main,
AdaIR-main/net/model.py, the experiment/*Rectiformer variants — SURVEY.md §2.3).  Instead of mirroring every wrapper,
this walks any ``nn.Module`` tree and replaces, by STRUCTURE (attribute names + parameter shapes, never by class
identity, so nothing of the reference is imported here):

  * blocks  ``norm1, attn, norm2, ffn``          (Restormer.TransformerBlock, Restormer.py:137-150)
  * blocks  ``norms[0..1], mixer, ffn``          (moce_ir.EncoderBlock, moce_ir.py:805-834)
  * MDTA    ``temperature, qkv, qkv_dwconv, project_out``   (Attention, Restormer.py:99-132)
  * GDFN    ``project_in, dwconv, project_out``             (FeedForward, Restormer.py:76-93)
  * gated MDTA ``gate_type`` + Attention's names with ``3*dim + gate`` qkv channels   (experiment GatedMDTA)
  * cross-MDTA ``temperature, q, q_dwconv, kv, kv_dwconv, project_out``   (moce_ir.CrossAttention, moce_ir.py:325-368)
  * LayerNorm wrappers with a ``body`` holding ``weight`` (and ``bias``)   (Restormer.py:60-70)
  * ``Downsample`` / ``Upsample`` = Sequential(conv3x3 no-bias, PixelUnshuffle/PixelShuffle(2))  (Restormer.py:171-189)
  * AdaIR's frequency module ``conv1, rate_conv, channel_cross_{l,h,agg}, frequency_refine, para1, para2``
    (FreModule, AdaIR-main/net/model.py:295-372)

The replacement adopts the reference's own ``nn.Conv2d`` / LayerNorm-body sub-modules, so ``state_dict()`` keys, the
optimizer's parameter objects and checkpoints are unchanged.  Variants that carry the same attribute names but a
different forward (gates, dynamic temperature ...) are caught by ``verify=True``: every candidate is run against the
original on a random input (fp32, on the model's CUDA device) and only swapped if the outputs agree.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch
import torch.nn as nn

from . import adair as AD
from . import modules as M


def _conv_ok(c, k, groups=None):
    return (isinstance(c, nn.Conv2d) and c.kernel_size == (k, k) and c.stride == (1, 1) and c.dilation == (1, 1)
            and c.padding == (k // 2, k // 2) and c.padding_mode == "zeros"
            and (groups is None or c.groups == groups))


def _is_ln(m) -> bool:
    body = getattr(m, "body", None)
    if body is None or not isinstance(body, nn.Module) or isinstance(body, nn.Sequential):
        return False
    names = {n for n, _ in body.named_parameters()}
    w = getattr(body, "weight", None)
    return names in ({"weight"}, {"weight", "bias"}) and w is not None and w.dim() == 1 and \
        {n for n, _ in m.named_parameters()} == {"body." + n for n in names}


def _is_mdta(m) -> bool:
    need = ("temperature", "qkv", "qkv_dwconv", "project_out", "num_heads")
    if not all(hasattr(m, a) for a in need):
        return False
    pn = {n.split(".")[0] for n, _ in m.named_parameters()}
    if pn != {"temperature", "qkv", "qkv_dwconv", "project_out"}:
        return False
    c3 = m.qkv.out_channels
    return (_conv_ok(m.qkv, 1, 1) and _conv_ok(m.qkv_dwconv, 3, c3) and _conv_ok(m.project_out, 1, 1)
            and c3 == 3 * m.qkv.in_channels and m.qkv_dwconv.in_channels == c3
            and m.project_out.in_channels == m.qkv.in_channels == m.project_out.out_channels
            and tuple(m.temperature.shape) == (m.num_heads, 1, 1) and m.qkv.in_channels % m.num_heads == 0)


def _is_gdfn(m) -> bool:
    if not all(hasattr(m, a) for a in ("project_in", "dwconv", "project_out")):
        return False
    pn = {n.split(".")[0] for n, _ in m.named_parameters()}
    if pn != {"project_in", "dwconv", "project_out"}:
        return False
    h2 = m.project_in.out_channels
    return (_conv_ok(m.project_in, 1, 1) and _conv_ok(m.dwconv, 3, h2) and _conv_ok(m.project_out, 1, 1)
            and h2 % 2 == 0 and m.dwconv.in_channels == h2 and m.project_out.in_channels == h2 // 2
            and m.project_out.out_channels == m.project_in.in_channels)


def _is_gated_mdta(m) -> bool:
    """GatedMDTA of the gated-attention experiment: Attention's names, `gate_type` in {headwise, elementwise} and
    3*dim + (num_heads | dim) qkv channels laid out [q | gate | k | v]."""
    need = ("temperature", "qkv", "qkv_dwconv", "project_out", "num_heads", "gate_type")
    if not all(hasattr(m, a) for a in need) or m.gate_type not in ("headwise", "elementwise"):
        return False
    pn = {n.split(".")[0] for n, _ in m.named_parameters()}
    if pn != {"temperature", "qkv", "qkv_dwconv", "project_out"}:
        return False
    c = m.qkv.in_channels
    ct = 3 * c + (m.num_heads if m.gate_type == "headwise" else c)
    return (_conv_ok(m.qkv, 1, 1) and _conv_ok(m.qkv_dwconv, 3, ct) and _conv_ok(m.project_out, 1, 1)
            and m.qkv.out_channels == ct and m.project_out.in_channels == c == m.project_out.out_channels
            and tuple(m.temperature.shape) == (m.num_heads, 1, 1) and c % m.num_heads == 0)


def _heads_of(m):
    """Head count of an attention module: ``num_heads`` (Restormer, MoCE-IR) or ``num_head`` (AdaIR)."""
    for a in ("num_heads", "num_head"):
        v = getattr(m, a, None)
        if isinstance(v, int) and v > 0:
            return v
    return None


def _is_cross(m) -> bool:
    """MoCE-IR CrossAttention (moce_ir.py:325-343): q / q_dwconv(3x3) from one input, kv / kv_dwconv(7x7) from another;
    AdaIR Chanel_Cross_Attention (AdaIR-main/net/model.py:177-216) is the same module with a 3x3 kv_dwconv and the head
    count spelled ``num_head``."""
    need = ("temperature", "q", "q_dwconv", "kv", "kv_dwconv", "project_out")
    if not all(hasattr(m, a) for a in need) or _heads_of(m) is None:
        return False
    heads = _heads_of(m)
    pn = {n.split(".")[0] for n, _ in m.named_parameters()}
    if pn != {"temperature", "q", "q_dwconv", "kv", "kv_dwconv", "project_out"}:
        return False
    c = m.q.in_channels
    kk = m.kv_dwconv.kernel_size[0] if isinstance(m.kv_dwconv, nn.Conv2d) else 0
    return (_conv_ok(m.q, 1, 1) and _conv_ok(m.q_dwconv, 3, c) and _conv_ok(m.kv, 1, 1) and kk in (3, 5, 7)
            and _conv_ok(m.kv_dwconv, kk, 2 * c) and _conv_ok(m.project_out, 1, 1)
            and m.q.out_channels == c and m.kv.in_channels == c and m.kv.out_channels == 2 * c
            and m.project_out.in_channels == c == m.project_out.out_channels
            and tuple(m.temperature.shape) == (heads, 1, 1) and c % heads == 0)


def _is_split_gate_ffn(m) -> bool:
    """RectiFormer GatedSpatialFFN (experiment/RectiFormer.py:231-266): project_in -> chunk ->
    gelu(dwconv_content(content)) * sigmoid(dwconv_gate(gate)) -> project_out."""
    need = ("project_in", "dwconv_gate", "dwconv_content", "project_out")
    if not all(hasattr(m, a) for a in need):
        return False
    if {n.split(".")[0] for n, _ in m.named_parameters()} != set(need):
        return False
    hid = m.project_out.in_channels if isinstance(m.project_out, nn.Conv2d) else 0
    return (_conv_ok(m.project_in, 1, 1) and _conv_ok(m.dwconv_gate, 3, hid) and _conv_ok(m.dwconv_content, 3, hid)
            and _conv_ok(m.project_out, 1, 1) and m.project_in.out_channels == 2 * hid
            and m.project_in.in_channels == m.project_out.out_channels)


def _is_elementwise_gated(m) -> bool:
    """RectiFormer ElementwiseGatedAttention (experiment/RectiFormer.py:270-325): MDTA + a gate branch
    gate_generator = Sequential(dw3x3, 1x1) on the block input, optional per-sample dynamic temperature."""
    need = ("temperature", "qkv", "qkv_dwconv", "gate_generator", "project_out")
    if not all(hasattr(m, a) for a in need) or not isinstance(getattr(m, "num_heads", None), int):
        return False
    if {n.split(".")[0] for n, _ in m.named_parameters()} != set(need):
        return False
    gg = m.gate_generator
    if not isinstance(gg, nn.Sequential) or len(gg) != 2:
        return False
    c = m.qkv.in_channels if isinstance(m.qkv, nn.Conv2d) else 0
    return (_conv_ok(m.qkv, 1, 1) and _conv_ok(m.qkv_dwconv, 3, 3 * c) and m.qkv.out_channels == 3 * c
            and _conv_ok(gg[0], 3, c) and _conv_ok(gg[1], 1, 1) and gg[1].in_channels == c == gg[1].out_channels
            and _conv_ok(m.project_out, 1, 1) and m.project_out.in_channels == c == m.project_out.out_channels
            and tuple(m.temperature.shape) == (m.num_heads, 1, 1) and c % m.num_heads == 0)


def _is_fremodule(m) -> bool:
    """AdaIR FreModule (AdaIR-main/net/model.py:295-372) with its FreRefine / SpatialGate / ChannelGate children."""
    kids = set(dict(m.named_children()))
    if kids != {"conv", "conv1", "score_gen", "channel_cross_l", "channel_cross_h", "channel_cross_agg", "frequency_refine",
                "rate_conv"}:
        return False
    if {n for n, _ in m.named_parameters(recurse=False)} != {"para1", "para2"}:
        return False
    if not all(_is_cross(getattr(m, a)) for a in ("channel_cross_l", "channel_cross_h", "channel_cross_agg")):
        return False
    dim = m.channel_cross_l.q.in_channels
    fr = m.frequency_refine
    sg, cg = getattr(fr, "SpatialGate", None), getattr(fr, "ChannelGate", None)
    if sg is None or cg is None or not hasattr(fr, "proj") or not hasattr(sg, "spatial") or not hasattr(cg, "mlp"):
        return False
    rc, mlp = m.rate_conv, cg.mlp
    return (_conv_ok(m.conv1, 3, 1) and m.conv1.out_channels == dim and m.conv1.bias is None
            and tuple(m.para1.shape) == (dim, 1, 1) == tuple(m.para2.shape)
            and _conv_ok(sg.spatial, 7, 1) and sg.spatial.in_channels == 2 and sg.spatial.out_channels == 1
            and sg.spatial.bias is None
            and isinstance(mlp, nn.Sequential) and len(mlp) == 3 and _conv_ok(mlp[0], 1, 1) and isinstance(mlp[1], nn.ReLU)
            and _conv_ok(mlp[2], 1, 1) and mlp[0].bias is None and mlp[2].bias is None
            and mlp[0].in_channels == dim == mlp[2].out_channels
            and _conv_ok(fr.proj, 1, 1) and fr.proj.in_channels == dim == fr.proj.out_channels
            and isinstance(rc, nn.Sequential) and len(rc) == 3 and _conv_ok(rc[0], 1, 1) and isinstance(rc[1], nn.GELU)
            and _conv_ok(rc[2], 1, 1) and rc[0].bias is None and rc[2].bias is None and rc[0].in_channels == dim
            and rc[2].out_channels == 2)


def _is_resize(m):
    """Sequential(conv3x3 (no bias), PixelUnshuffle(2) | PixelShuffle(2)) wrapped as `.body` -> 'down' | 'up' | None."""
    body = getattr(m, "body", None)
    if not isinstance(body, nn.Sequential) or len(body) != 2 or not _conv_ok(body[0], 3, 1) or body[0].bias is not None:
        return None
    if {n for n, _ in m.named_parameters()} != {"body.0.weight"}:
        return None
    if isinstance(body[1], nn.PixelUnshuffle) and body[1].downscale_factor == 2:
        return "down"
    if isinstance(body[1], nn.PixelShuffle) and body[1].upscale_factor == 2:
        return "up"
    return None


def _bare(cls):
    obj = cls.__new__(cls)
    nn.Module.__init__(obj)
    return obj


def _adopt_ln(m):
    ln = _bare(M.LayerNorm)
    ln.body = m.body
    return ln


def _adopt_mdta(m):
    a = _bare(M.Attention)
    a.num_heads = m.num_heads
    a.temperature = m.temperature
    a.qkv, a.qkv_dwconv, a.project_out = m.qkv, m.qkv_dwconv, m.project_out
    return a


def _adopt_gated(m):
    a = _bare(M.GatedMDTA)
    a.num_heads, a.gate_type, a.dim, a.bias = m.num_heads, m.gate_type, m.qkv.in_channels, m.qkv.bias is not None
    a.qkv_out_channels = m.qkv.out_channels
    a.temperature = m.temperature
    a.qkv, a.qkv_dwconv, a.project_out = m.qkv, m.qkv_dwconv, m.project_out
    return a


def _adopt_cross(m):
    a = _bare(M.CrossAttention)
    a.num_heads = _heads_of(m)
    a.temperature = m.temperature
    a.q, a.q_dwconv, a.kv, a.kv_dwconv, a.project_out = m.q, m.q_dwconv, m.kv, m.kv_dwconv, m.project_out
    return a


def _adopt_fremodule(m):
    f = _bare(AD.FreModule)
    f.conv, f.conv1, f.score_gen = m.conv, m.conv1, m.score_gen      # (registration order = the reference's state_dict order)
    f.para1, f.para2 = m.para1, m.para2
    for a in ("channel_cross_l", "channel_cross_h", "channel_cross_agg"):
        setattr(f, a, _adopt_cross(getattr(m, a)))
    r = _bare(AD.FreRefine)
    r.SpatialGate = _bare(AD.SpatialGate)
    r.SpatialGate.spatial = m.frequency_refine.SpatialGate.spatial
    r.ChannelGate = _bare(AD.ChannelGate)
    for a in ("avg", "max"):
        if hasattr(m.frequency_refine.ChannelGate, a):
            setattr(r.ChannelGate, a, getattr(m.frequency_refine.ChannelGate, a))
    r.ChannelGate.mlp = m.frequency_refine.ChannelGate.mlp
    r.proj = m.frequency_refine.proj
    f.frequency_refine = r
    f.rate_conv = m.rate_conv
    return f


def _adopt_split_gate_ffn(m):
    f = _bare(M.GatedSpatialFFN)
    f.project_in, f.dwconv_gate, f.dwconv_content, f.project_out = m.project_in, m.dwconv_gate, m.dwconv_content, m.project_out
    return f


def _adopt_elementwise_gated(m):
    a = _bare(M.ElementwiseGatedAttention)
    a.num_heads, a.dim = m.num_heads, m.qkv.in_channels
    a.temperature = m.temperature
    a.qkv, a.qkv_dwconv, a.gate_generator, a.project_out = m.qkv, m.qkv_dwconv, m.gate_generator, m.project_out
    return a


def _adopt_gdfn(m):
    f = _bare(M.FeedForward)
    f.project_in, f.dwconv, f.project_out = m.project_in, m.dwconv, m.project_out
    return f


def _adopt(m):
    """-> (replacement, kind) or (None, None)."""
    if isinstance(m, (M.TransformerBlock, M.EncoderBlock, M.Attention, M.CrossAttention, M.GatedMDTA,
                      M.GatedTransformerBlock, M.FeedForward, M.LayerNorm, M.Downsample, M.Upsample,
                      M.GatedSpatialFFN, M.ElementwiseGatedAttention, AD.FreModule)):
        return None, None
    kids = dict(m.named_children())
    if _is_fremodule(m):
        return _adopt_fremodule(m), "frequency_module"
    if set(kids) == {"norm1", "attn", "norm2", "ffn"} and _is_ln(m.norm1) and _is_ln(m.norm2) and _is_mdta(m.attn) \
            and _is_gdfn(m.ffn) and not list(m.named_parameters(recurse=False)):
        b = _bare(M.TransformerBlock)
        b.norm1, b.attn, b.norm2, b.ffn = _adopt_ln(m.norm1), _adopt_mdta(m.attn), _adopt_ln(m.norm2), _adopt_gdfn(m.ffn)
        return b, "block"
    if set(kids) == {"norms", "mixer", "ffn"} and isinstance(m.norms, nn.ModuleList) and len(m.norms) == 2 \
            and all(_is_ln(n) for n in m.norms) and _is_mdta(m.mixer) and _is_gdfn(m.ffn) \
            and not list(m.named_parameters(recurse=False)):
        b = _bare(M.EncoderBlock)
        b.norms = nn.ModuleList([_adopt_ln(m.norms[0]), _adopt_ln(m.norms[1])])
        b.mixer, b.ffn = _adopt_mdta(m.mixer), _adopt_gdfn(m.ffn)
        return b, "encoder_block"
    if _is_mdta(m):
        return _adopt_mdta(m), "attention"
    if _is_cross(m):
        return _adopt_cross(m), "cross_attention"
    if _is_gated_mdta(m):
        return _adopt_gated(m), "gated_attention"
    if _is_elementwise_gated(m):
        return _adopt_elementwise_gated(m), "elementwise_gated_attention"
    if _is_split_gate_ffn(m):
        return _adopt_split_gate_ffn(m), "split_gate_ffn"
    if _is_gdfn(m):
        return _adopt_gdfn(m), "feedforward"
    if _is_ln(m):
        return _adopt_ln(m), "layernorm"
    rz = _is_resize(m)
    if rz:
        r = _bare(M.Downsample if rz == "down" else M.Upsample)
        r.body = m.body
        return r, rz + "sample"
    return None, None


def _in_channels(new, kind):
    if kind in ("block",):
        return new.attn.qkv.in_channels
    if kind == "encoder_block":
        return new.mixer.qkv.in_channels
    if kind in ("attention", "gated_attention", "elementwise_gated_attention"):
        return new.qkv.in_channels
    if kind == "split_gate_ffn":
        return new.project_in.in_channels
    if kind == "cross_attention":
        return new.q.in_channels
    if kind == "frequency_module":
        return new.conv1.in_channels
    if kind == "feedforward":
        return new.project_in.in_channels
    if kind == "layernorm":
        return new.body.weight.numel()
    return new.body[0].in_channels


@torch.no_grad()
def _agrees(old, new, kind, tol):
    p = next(new.parameters())
    if not p.is_cuda or p.dtype != torch.float32:
        raise ValueError("accelerate(verify=True) needs the model on a CUDA device in float32")
    g = torch.Generator(device="cpu").manual_seed(1234)
    x = (torch.rand(2, _in_channels(new, kind), 16, 24, generator=g) * 2 - 1).to(p.device)
    # the ORIGINAL runs through cuDNN / cuBLAS here: keep it in true fp32 for the comparison (TF32 is on by default for
    # cuDNN convolutions and would differ from either exact implementation by ~1e-3)
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        if kind == "cross_attention":
            y = (torch.rand(x.shape, generator=g) * 2 - 1).to(p.device)
            ref, got = old(x, y), new(x, y)
        elif kind == "frequency_module":      # (image, feature); the image is resized to the feature's size inside
            y = (torch.rand(2, new.conv1.out_channels, 8, 12, generator=g) * 2 - 1).to(p.device)
            ref, got = old(x, y), new(x, y)
        else:
            ref = old(x)
            got = new(x)
    except Exception:
        return False
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    return ref.shape == got.shape and float((ref - got).abs().max()) <= tol * max(1.0, float(ref.abs().max()))


def accelerate(model: nn.Module, verify: bool = True, tol: float = 1e-4) -> Tuple[nn.Module, Dict[str, int]]:
    """Replace every structurally matching sub-module of `model` in place.  Returns (model, {kind: count}); modules that
    match by structure but not by behaviour (verify) are counted under 'rejected' and left untouched."""
    stats: Dict[str, int] = {}
    work: List[Tuple[nn.Module, str, nn.Module]] = []

    def walk(parent):
        for name, child in list(parent.named_children()):
            new, kind = _adopt(child)
            if new is None:
                walk(child)
                continue
            new.train(child.training)
            if verify and not _agrees(child, new, kind, tol):
                stats["rejected"] = stats.get("rejected", 0) + 1
                walk(child)
                continue
            work.append((parent, name, new))
            stats[kind] = stats.get(kind, 0) + 1

    walk(model)
    for parent, name, new in work:
        if isinstance(parent, (nn.Sequential, nn.ModuleList)) and name.isdigit():
            parent[int(name)] = new
        else:
            setattr(parent, name, new)
    return model, stats
