"""CPU oracle for the Restormer-family transformer block hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may call it, and there only as the checker / the timed
CPU baseline, never as the thing shipped.

It is a *functional restatement* (state_dict in, tensor out; no nn.Module) of the
algorithm in the reference's ``Restormer.py``.  Each function cites the reference
file:line it follows.  The arithmetic of the reference lives in third-party
``torch`` (ATen) and ``einops`` (reference pins: ``AdaIR-main/env.yaml:146,168``
einops==0.6.0; torch unpinned, ``MoCE-IR-main/install.sh:2``); here the same
ATen CPU ops are used (``conv2d``, ``var``, ``softmax``, ``gelu``), with the
``einops.rearrange`` calls restated as ``reshape``/``permute``.

Parity pin: the reference publishes no golden vectors for this path (SURVEY.md
§4, §8c), so the oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF run in
the build container: ``oracle/gen_golden.py`` imports ``/root/reference/Restormer.py``
unmodified, feeds it seeded synthetic weights/inputs and stores the outputs under
``tests/golden/``; ``tests/test_oracle_golden.py`` checks this file against them.

Works in float32 (default, what the reference computes in) or float64 (tighter
yardstick for the CUDA fp32 path).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------------------
# LayerNorm over channels of an NCHW tensor            (Restormer.py:19-70)
# --------------------------------------------------------------------------------------
def layernorm(x: Tensor, weight: Tensor, bias: Optional[Tensor]) -> Tensor:
    """``LayerNorm.forward`` = to_4d(body(to_3d(x)))  (Restormer.py:68-70).

    WithBias (bias given):  (x-mu)/sqrt(var_biased+1e-5)*w+b      (Restormer.py:54-57)
    BiasFree (bias None):    x/sqrt(var_biased+1e-5)*w             (Restormer.py:37-39)
    -- note the BiasFree numerator is NOT centred while the variance is about the mean.
    The reduction runs over C for every (b, h, w); to_3d/to_4d (Restormer.py:19-23) only
    move C last and back, so here the reduction is done over dim=1 directly.
    """
    mu = x.mean(dim=1, keepdim=True)
    var = x.var(dim=1, keepdim=True, unbiased=False)
    w = weight.view(1, -1, 1, 1)
    if bias is None:
        return x / torch.sqrt(var + 1e-5) * w
    return (x - mu) / torch.sqrt(var + 1e-5) * w + bias.view(1, -1, 1, 1)


# --------------------------------------------------------------------------------------
# MDTA attention                                      (Restormer.py:99-132)
# --------------------------------------------------------------------------------------
def attention(x: Tensor, p: Dict[str, Tensor], prefix: str, num_heads: int) -> Tensor:
    """``Attention.forward`` (Restormer.py:111-132)."""
    b, c, h, w = x.shape
    g = lambda k: p.get(prefix + k)
    qkv = F.conv2d(x, g("qkv.weight"), g("qkv.bias"))                              # :114 1x1
    qkv = F.conv2d(qkv, g("qkv_dwconv.weight"), g("qkv_dwconv.bias"), padding=1, groups=3 * c)  # :114 dw3x3
    q, k, v = qkv.chunk(3, dim=1)                                                  # :115
    d = c // num_heads
    q = q.reshape(b, num_heads, d, h * w)                                          # :117 'b (head c) h w -> b head c (h w)'
    k = k.reshape(b, num_heads, d, h * w)                                          # :118
    v = v.reshape(b, num_heads, d, h * w)                                          # :119
    q = F.normalize(q, dim=-1)                                                     # :121 x / max(||x||, 1e-12)
    k = F.normalize(k, dim=-1)                                                     # :122
    attn = (q @ k.transpose(-2, -1)) * g("temperature")                            # :124 temperature [h,1,1]
    attn = attn.softmax(dim=-1)                                                    # :125
    out = attn @ v                                                                 # :127
    out = out.reshape(b, c, h, w)                                                  # :129
    return F.conv2d(out, g("project_out.weight"), g("project_out.bias"))           # :131


def cross_attention(x: Tensor, y: Tensor, p: Dict[str, Tensor], prefix: str, num_heads: int) -> Tensor:
    """MoCE-IR ``CrossAttention.forward`` (moce_ir.py:345-368): MDTA whose q comes from ``x`` (1x1 + depthwise 3x3) and
    whose k, v come from ``y`` (1x1 + depthwise 7x7, padding 3); no LayerNorm, no residual inside."""
    b, c, h, w = x.shape
    g = lambda k: p.get(prefix + k)
    q = F.conv2d(x, g("q.weight"), g("q.bias"))                                               # :348 1x1
    q = F.conv2d(q, g("q_dwconv.weight"), g("q_dwconv.bias"), padding=1, groups=c)            # :348 dw3x3
    kv = F.conv2d(y, g("kv.weight"), g("kv.bias"))                                            # :349 1x1
    kk = g("kv_dwconv.weight").shape[-1]
    kv = F.conv2d(kv, g("kv_dwconv.weight"), g("kv_dwconv.bias"), padding=kk // 2, groups=2 * c)   # :349 dw7x7
    k, v = kv.chunk(2, dim=1)                                                                 # :350
    d = c // num_heads
    q = F.normalize(q.reshape(b, num_heads, d, h * w), dim=-1)                                # :352,356
    k = F.normalize(k.reshape(b, num_heads, d, h * w), dim=-1)                                # :353,357
    v = v.reshape(b, num_heads, d, h * w)                                                     # :354
    attn = ((q @ k.transpose(-2, -1)) * g("temperature")).softmax(dim=-1)                     # :359-361
    out = (attn @ v).reshape(b, c, h, w)                                                      # :363-365
    return F.conv2d(out, g("project_out.weight"), g("project_out.bias"))                      # :367


def gated_attention(x: Tensor, p: Dict[str, Tensor], prefix: str, num_heads: int, gate_type: Optional[str]) -> Tensor:
    """``GatedMDTA.forward`` of the reference's gated-attention experiment (experiment/baseline_restormer_new_train/
    20251116_use_headwise_gate_model/MoCE-IR-main/src/net/moce_ir.py:117-171): MDTA whose qkv projection carries extra
    gate channels after q -- ``[q | gate | k | v]`` with ``num_heads`` ('headwise') or ``dim`` ('elementwise') gate
    channels -- and whose attention output is multiplied by ``sigmoid(gate)`` before ``project_out``."""
    b, c, h, w = x.shape
    g = lambda k: p.get(prefix + k)
    t = F.conv2d(x, g("qkv.weight"), g("qkv.bias"))                                             # :122 1x1
    t = F.conv2d(t, g("qkv_dwconv.weight"), g("qkv_dwconv.bias"), padding=1, groups=t.shape[1])   # :122 dw3x3
    ng = {None: 0, "headwise": num_heads, "elementwise": c}[gate_type]
    q, gate, k, v = t.split([c, ng, c, c], dim=1)                                                # :125-140
    d = c // num_heads
    q = F.normalize(q.reshape(b, num_heads, d, h * w), dim=-1)                                   # :144,149
    k = F.normalize(k.reshape(b, num_heads, d, h * w), dim=-1)                                   # :145,150
    v = v.reshape(b, num_heads, d, h * w)                                                        # :146
    attn = ((q @ k.transpose(-2, -1)) * g("temperature")).softmax(dim=-1)                        # :151-152
    out = attn @ v                                                                               # :155
    if gate_type == "headwise":
        out = out * torch.sigmoid(gate.reshape(b, num_heads, 1, h * w))                          # :160-166
    elif gate_type == "elementwise":
        out = out * torch.sigmoid(gate.reshape(b, num_heads, d, h * w))                          # :162-166
    return F.conv2d(out.reshape(b, c, h, w), g("project_out.weight"), g("project_out.bias"))     # :169-170


def synth_gated_attention_params(dim: int, heads: int, bias: bool, gate_type: Optional[str], seed: int) -> Dict[str, Tensor]:
    """Seeded synthetic state_dict of a GatedMDTA module (same names as Attention; qkv has 3*dim + gate channels)."""
    gen = torch.Generator().manual_seed(seed)
    ng = {None: 0, "headwise": heads, "elementwise": dim}[gate_type]
    ct = 3 * dim + ng
    sd = {"temperature": 0.5 + torch.rand(heads, 1, 1, generator=gen)}
    sd["qkv.weight"] = (torch.rand(ct, dim, 1, 1, generator=gen) - 0.5) * (2.0 / dim ** 0.5)
    sd["qkv_dwconv.weight"] = (torch.rand(ct, 1, 3, 3, generator=gen) - 0.5) * (2.0 / 3.0)
    sd["project_out.weight"] = (torch.rand(dim, dim, 1, 1, generator=gen) - 0.5) * (2.0 / dim ** 0.5)
    if bias:
        sd["qkv.bias"] = (torch.rand(ct, generator=gen) - 0.5) * 0.2
        sd["qkv_dwconv.bias"] = (torch.rand(ct, generator=gen) - 0.5) * 0.2
        sd["project_out.bias"] = (torch.rand(dim, generator=gen) - 0.5) * 0.2
    return sd


def synth_cross_attention_params(dim: int, heads: int, bias: bool, seed: int, kv_kernel: int = 7) -> Dict[str, Tensor]:
    """Seeded synthetic state_dict of a CrossAttention module (names of moce_ir.py:334-343)."""
    gen = torch.Generator().manual_seed(seed)
    sd = {"temperature": 0.5 + torch.rand(heads, 1, 1, generator=gen)}
    def conv(name, cout, cin, k, groups=1):
        fan = (cin // groups) * k * k
        sd[name + ".weight"] = (torch.rand(cout, cin // groups, k, k, generator=gen) - 0.5) * (2.0 / fan ** 0.5)
        if bias:
            sd[name + ".bias"] = (torch.rand(cout, generator=gen) - 0.5) * 0.2
    conv("q", dim, dim, 1)
    conv("q_dwconv", dim, dim, 3, groups=dim)
    conv("kv", 2 * dim, dim, 1)
    conv("kv_dwconv", 2 * dim, 2 * dim, kv_kernel, groups=2 * dim)
    conv("project_out", dim, dim, 1)
    return sd


# --------------------------------------------------------------------------------------
# GDFN feed-forward                                   (Restormer.py:76-93)
# --------------------------------------------------------------------------------------
def feedforward(x: Tensor, p: Dict[str, Tensor], prefix: str) -> Tensor:
    """``FeedForward.forward`` (Restormer.py:88-93)."""
    g = lambda k: p.get(prefix + k)
    w_in = g("project_in.weight")
    x = F.conv2d(x, w_in, g("project_in.bias"))                                    # :89
    x = F.conv2d(x, g("dwconv.weight"), g("dwconv.bias"), padding=1, groups=w_in.shape[0])  # :90
    x1, x2 = x.chunk(2, dim=1)                                                     # :90
    x = F.gelu(x1) * x2                                                            # :91 exact erf GELU
    return F.conv2d(x, g("project_out.weight"), g("project_out.bias"))            # :92


def transformer_block(x: Tensor, p: Dict[str, Tensor], prefix: str, num_heads: int) -> Tensor:
    """``TransformerBlock.forward`` (Restormer.py:146-150)."""
    n1b = p.get(prefix + "norm1.body.bias")
    n2b = p.get(prefix + "norm2.body.bias")
    x = x + attention(layernorm(x, p[prefix + "norm1.body.weight"], n1b), p, prefix + "attn.", num_heads)
    x = x + feedforward(layernorm(x, p[prefix + "norm2.body.weight"], n2b), p, prefix + "ffn.")
    return x


# --------------------------------------------------------------------------------------
# glue                                                (Restormer.py:156-189)
# --------------------------------------------------------------------------------------
def overlap_patch_embed(x: Tensor, p: Dict[str, Tensor], prefix: str) -> Tensor:
    """``OverlapPatchEmbed.forward`` (Restormer.py:162-165): conv3x3, pad 1."""
    return F.conv2d(x, p[prefix + "proj.weight"], p.get(prefix + "proj.bias"), padding=1)


def pixel_unshuffle2(t: Tensor) -> Tensor:
    """nn.PixelUnshuffle(2): out[b, c*4+i*2+j, y, x] = t[b, c, 2y+i, 2x+j] (used at Restormer.py:176)."""
    b, c, h, w = t.shape
    t = t.reshape(b, c, h // 2, 2, w // 2, 2).permute(0, 1, 3, 5, 2, 4)
    return t.reshape(b, c * 4, h // 2, w // 2)


def pixel_shuffle2(t: Tensor) -> Tensor:
    """nn.PixelShuffle(2): out[b, c, 2y+i, 2x+j] = t[b, c*4+i*2+j, y, x] (used at Restormer.py:186)."""
    b, c4, h, w = t.shape
    c = c4 // 4
    t = t.reshape(b, c, 2, 2, h, w).permute(0, 1, 4, 2, 5, 3)
    return t.reshape(b, c, h * 2, w * 2)


def downsample(x: Tensor, p: Dict[str, Tensor], prefix: str) -> Tensor:
    """``Downsample.forward`` (Restormer.py:171-179): conv3x3 C->C/2 (no bias) + PixelUnshuffle(2)."""
    return pixel_unshuffle2(F.conv2d(x, p[prefix + "body.0.weight"], None, padding=1))


def upsample(x: Tensor, p: Dict[str, Tensor], prefix: str) -> Tensor:
    """``Upsample.forward`` (Restormer.py:181-189): conv3x3 C->2C (no bias) + PixelShuffle(2)."""
    return pixel_shuffle2(F.conv2d(x, p[prefix + "body.0.weight"], None, padding=1))


# --------------------------------------------------------------------------------------
# whole network                                       (Restormer.py:193-284)
# --------------------------------------------------------------------------------------
DEFAULT_CFG = dict(inp_channels=3, out_channels=3, dim=48, num_blocks=[4, 6, 6, 8],
                   num_refinement_blocks=4, heads=[1, 2, 4, 8], ffn_expansion_factor=2.66,
                   bias=False, LayerNorm_type="WithBias", dual_pixel_task=False)


def _stack(x, p, prefix, n, heads):
    for i in range(n):
        x = transformer_block(x, p, f"{prefix}.{i}.", heads)
    return x


def restormer_forward(p: Dict[str, Tensor], inp_img: Tensor, cfg: Optional[dict] = None) -> Tensor:
    """``Restormer.forward`` (Restormer.py:245-284)."""
    c = dict(DEFAULT_CFG)
    if cfg:
        c.update(cfg)
    nb, heads = c["num_blocks"], c["heads"]
    e1_in = overlap_patch_embed(inp_img, p, "patch_embed.")                        # :247
    e1 = _stack(e1_in, p, "encoder_level1", nb[0], heads[0])                       # :248
    e2 = _stack(downsample(e1, p, "down1_2."), p, "encoder_level2", nb[1], heads[1])   # :250-251
    e3 = _stack(downsample(e2, p, "down2_3."), p, "encoder_level3", nb[2], heads[2])   # :253-254
    lat = _stack(downsample(e3, p, "down3_4."), p, "latent", nb[3], heads[3])          # :256-257
    d3 = torch.cat([upsample(lat, p, "up4_3."), e3], 1)                            # :259-260
    d3 = F.conv2d(d3, p["reduce_chan_level3.weight"], p.get("reduce_chan_level3.bias"))  # :261
    d3 = _stack(d3, p, "decoder_level3", nb[2], heads[2])                          # :262
    d2 = torch.cat([upsample(d3, p, "up3_2."), e2], 1)                             # :264-265
    d2 = F.conv2d(d2, p["reduce_chan_level2.weight"], p.get("reduce_chan_level2.bias"))  # :266
    d2 = _stack(d2, p, "decoder_level2", nb[1], heads[1])                          # :267
    d1 = torch.cat([upsample(d2, p, "up2_1."), e1], 1)                             # :269-270 (no reduce conv)
    d1 = _stack(d1, p, "decoder_level1", nb[0], heads[0])                          # :271
    d1 = _stack(d1, p, "refinement", c["num_refinement_blocks"], heads[0])         # :273
    if c["dual_pixel_task"]:                                                       # :276-278
        d1 = d1 + F.conv2d(e1_in, p["skip_conv.weight"], p.get("skip_conv.bias"))
        return F.conv2d(d1, p["output.weight"], p.get("output.bias"), padding=1)
    return F.conv2d(d1, p["output.weight"], p.get("output.bias"), padding=1) + inp_img  # :281


# --------------------------------------------------------------------------------------
# synthetic, seed-reproducible weights with the reference's names and shapes
# --------------------------------------------------------------------------------------
def restormer_param_shapes(cfg: Optional[dict] = None) -> "List[tuple]":
    """(name, shape, kind) for every tensor of ``Restormer.state_dict()`` in the reference's
    registration order (Restormer.py:194-243); kind in {conv, dw, ln_w, ln_b, temp, bias}."""
    c = dict(DEFAULT_CFG)
    if cfg:
        c.update(cfg)
    dim, nb, heads, f = c["dim"], c["num_blocks"], c["heads"], c["ffn_expansion_factor"]
    bias, lnb = c["bias"], c["LayerNorm_type"] != "BiasFree"
    out: List[tuple] = []

    def conv(name, co, ci, k, b, groups=1):
        out.append((name + ".weight", (co, ci // groups, k, k), "dw" if groups > 1 else "conv"))
        if b:
            out.append((name + ".bias", (co,), "bias"))

    def block(prefix, d, h):
        hd = int(d * f)
        out.append((prefix + "norm1.body.weight", (d,), "ln_w"))
        if lnb:
            out.append((prefix + "norm1.body.bias", (d,), "ln_b"))
        out.append((prefix + "attn.temperature", (h, 1, 1), "temp"))
        conv(prefix + "attn.qkv", 3 * d, d, 1, bias)
        conv(prefix + "attn.qkv_dwconv", 3 * d, 3 * d, 3, bias, groups=3 * d)
        conv(prefix + "attn.project_out", d, d, 1, bias)
        out.append((prefix + "norm2.body.weight", (d,), "ln_w"))
        if lnb:
            out.append((prefix + "norm2.body.bias", (d,), "ln_b"))
        conv(prefix + "ffn.project_in", 2 * hd, d, 1, bias)
        conv(prefix + "ffn.dwconv", 2 * hd, 2 * hd, 3, bias, groups=2 * hd)
        conv(prefix + "ffn.project_out", d, hd, 1, bias)

    def stack(name, n, d, h):
        for i in range(n):
            block(f"{name}.{i}.", d, h)

    conv("patch_embed.proj", dim, c["inp_channels"], 3, False)
    stack("encoder_level1", nb[0], dim, heads[0])
    conv("down1_2.body.0", dim // 2, dim, 3, False)
    stack("encoder_level2", nb[1], dim * 2, heads[1])
    conv("down2_3.body.0", dim, dim * 2, 3, False)
    stack("encoder_level3", nb[2], dim * 4, heads[2])
    conv("down3_4.body.0", dim * 2, dim * 4, 3, False)
    stack("latent", nb[3], dim * 8, heads[3])
    conv("up4_3.body.0", dim * 16, dim * 8, 3, False)
    conv("reduce_chan_level3", dim * 4, dim * 8, 1, bias)
    stack("decoder_level3", nb[2], dim * 4, heads[2])
    conv("up3_2.body.0", dim * 8, dim * 4, 3, False)
    conv("reduce_chan_level2", dim * 2, dim * 4, 1, bias)
    stack("decoder_level2", nb[1], dim * 2, heads[1])
    conv("up2_1.body.0", dim * 4, dim * 2, 3, False)
    stack("decoder_level1", nb[0], dim * 2, heads[0])
    stack("refinement", c["num_refinement_blocks"], dim * 2, heads[0])
    if c["dual_pixel_task"]:
        conv("skip_conv", dim * 2, dim, 1, bias)
    conv("output", c["out_channels"], dim * 2, 3, bias)
    return out


def synth_tensor(shape, kind: str, gen: torch.Generator) -> Tensor:
    """Seeded synthetic values.  Conv weights: uniform(+-1/sqrt(fan_in)) — the bound
    nn.Conv2d's default kaiming_uniform(a=sqrt(5)) uses (the init the reference gets at
    Restormer.py:194-243) — LN weights around 1, LN/conv biases small non-zero and
    temperatures in [0.5,1.5] so that every parameter is exercised (the reference's
    ones/zeros init would hide bias/temperature bugs)."""
    if kind in ("conv", "dw"):
        fan_in = shape[1] * shape[2] * shape[3]
        bound = 1.0 / math.sqrt(fan_in)
        return (torch.rand(shape, generator=gen) * 2 - 1) * bound
    if kind == "ln_w":
        return 1.0 + 0.2 * (torch.rand(shape, generator=gen) * 2 - 1)
    if kind == "temp":
        return 0.5 + torch.rand(shape, generator=gen)
    if kind in ("ln_b", "bias"):
        return 0.1 * (torch.rand(shape, generator=gen) * 2 - 1)
    raise ValueError(kind)


def synth_state_dict(cfg: Optional[dict] = None, seed: int = 0, dtype=torch.float32) -> Dict[str, Tensor]:
    """A full ``Restormer`` state_dict (reference names/shapes/order) from one seed."""
    gen = torch.Generator().manual_seed(seed)
    return {n: synth_tensor(s, k, gen).to(dtype) for n, s, k in restormer_param_shapes(cfg)}


def synth_block_params(dim: int, heads: int, f: float, bias: bool, ln_type: str, seed: int,
                       dtype=torch.float32) -> Dict[str, Tensor]:
    """state_dict of one ``TransformerBlock(dim, heads, f, bias, ln_type)`` (no prefix)."""
    cfg = dict(dim=dim, num_blocks=[1, 0, 0, 0], heads=[heads] * 4, ffn_expansion_factor=f, bias=bias,
               LayerNorm_type=ln_type, num_refinement_blocks=0)
    gen = torch.Generator().manual_seed(seed)
    out = {}
    for n, s, k in restormer_param_shapes(cfg):
        if n.startswith("encoder_level1.0."):
            out[n[len("encoder_level1.0."):]] = synth_tensor(s, k, gen).to(dtype)
    return out


def synth_image(b: int, h: int, w: int, seed: int = 1234, c: int = 3, dtype=torch.float32) -> Tensor:
    """Synthetic input, uniform [0,1) like ToTensor() output (MoCE-IR-main/src/data/dataset_utils.py:185-186)."""
    return torch.rand(b, c, h, w, generator=torch.Generator().manual_seed(seed)).to(dtype)


def psnr(a: Tensor, b: Tensor, peak: float = 1.0) -> float:
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    return float("inf") if mse == 0 else 10.0 * math.log10(peak * peak / mse)


# --------------------------------------------------------------------------------------
# training step                                       (MoCE-IR-main/src/train.py:51-88)
# --------------------------------------------------------------------------------------
def l1_loss(pred: Tensor, target: Tensor) -> Tensor:
    """``nn.L1Loss()`` with the default mean reduction (MoCE-IR-main/src/train.py:51,65)."""
    return (pred - target).abs().mean()


def block_grads(x: Tensor, p: Dict[str, Tensor], num_heads: int, gy: Tensor):
    """Gradients of ``sum(transformer_block(x) * gy)`` w.r.t. x and every parameter — what torch autograd gives the
    reference inside ``loss.backward()``.  Returns (y, dx, {name: dparam})."""
    x = x.detach().clone().requires_grad_(True)
    q = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    y = transformer_block(x, q, "", num_heads)
    names = sorted(q)
    gs = torch.autograd.grad(y, [x] + [q[n] for n in names], gy)
    return y.detach(), gs[0], dict(zip(names, gs[1:]))


def adamw_reference(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, lr=2e-4, betas=(0.9, 0.999), eps=1e-8,
                    weight_decay=1e-2):
    """One ``torch.optim.AdamW`` update (defaults of ``optim.AdamW(lr=2e-4)``, MoCE-IR-main/src/train.py:83) restated
    on flat tensors: decoupled decay, bias-corrected moments, eps outside sqrt(v_hat).  Returns (p, m, v)."""
    b1, b2 = betas
    p = p * (1 - lr * weight_decay)
    m = m + (1 - b1) * (g - m)
    v = b2 * v + (1 - b2) * g * g
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    p = p - (lr / bc1) * m / (v.sqrt() / math.sqrt(bc2) + eps)
    return p, m, v


def train_step_reference(sd: Dict[str, Tensor], cfg: Optional[dict], degraded: Tensor, clean: Tensor, steps: int = 1,
                         lr: float = 2e-4):
    """``training_step`` + ``configure_optimizers`` of the reference driver as a plain loop (Lightning is not
    installed): restored = net(degrad); loss = L1(restored, clean); AdamW(lr=2e-4) step
    (MoCE-IR-main/src/train.py:59-88; the LinearWarmupCosine scheduler is left out — lr is constant here).
    Returns (losses, grads of the last step, updated state_dict)."""
    q = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    names = sorted(q)
    opt = torch.optim.AdamW([q[n] for n in names], lr=lr)
    losses, grads = [], None
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        loss = l1_loss(restormer_forward(q, degraded, cfg), clean)
        loss.backward()
        grads = {n: q[n].grad.detach().clone() for n in names}
        opt.step()
        losses.append(loss.item())
    return losses, grads, {n: q[n].detach().clone() for n in names}
