"""Case tables shared by oracle/gen_golden.py (generator) and tests/test_oracle_golden.py."""
# (name, dim, heads, ffn_factor, bias, ln_type, B, H, W, seed)
BLOCK_CASES = [
    ("blk_c48_h1", 48, 1, 2.66, False, "WithBias", 2, 16, 24, 11),
    ("blk_c96_h2_bias", 96, 2, 2.66, True, "WithBias", 1, 16, 16, 12),
    ("blk_c96_h1_biasfree", 96, 1, 2.66, False, "BiasFree", 1, 8, 16, 13),
    ("blk_c64_h2_f2_bias", 64, 2, 2.0, True, "BiasFree", 1, 8, 8, 14),
    ("blk_c192_h4", 192, 4, 2.66, False, "WithBias", 1, 8, 8, 15),
]

# (name, cfg-overrides, B, H, W, weight seed, image seed)
NET_CASES = [
    ("net_default_64", {}, 1, 64, 64, 0, 1234),
    ("net_dim16_bias_biasfree_32", dict(dim=16, num_blocks=[1, 2, 2, 1], num_refinement_blocks=1, bias=True,
                                        LayerNorm_type="BiasFree"), 2, 32, 48, 3, 77),
    ("net_dim32_dualpixel_32", dict(dim=32, inp_channels=6, num_blocks=[1, 1, 1, 2], num_refinement_blocks=2,
                                    ffn_expansion_factor=2.0, dual_pixel_task=True), 1, 32, 32, 5, 78),
]



# backward / training-step fixtures (oracle/gen_golden_train.py)
# (name, dim, heads, ffn_factor, bias, ln_type, B, H, W, seed)
TRAIN_BLOCK_CASES = [
    ("train_blk_c48_h1", 48, 1, 2.66, False, "WithBias", 2, 8, 16, 31),
    ("train_blk_c32_h2_bias_biasfree", 32, 2, 2.0, True, "BiasFree", 1, 8, 8, 32),
]
# (name, cfg-overrides, B, H, W, weight seed, image seed, steps)
TRAIN_NET_CASE = ("train_net_dim8_16", dict(dim=8, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1, heads=[1, 2, 4, 8]),
                  2, 16, 16, 9, 41, 3)

# MoCE-IR encoder groups (oracle/gen_golden_moce.py): (name, dim, heads, ffn_factor, bias, ln, blocks, B, H, W, seed)
MOCE_CASES = [
    ("moce_encgrp_c32_h2", 32, 2, 2, True, "WithBias", 2, 2, 8, 16, 51),
    ("moce_encgrp_c48_h1_biasfree", 48, 1, 2, True, "BiasFree", 1, 1, 8, 8, 52),
]


# CrossAttention of the MoCE-IR decoder (moce_ir.py:325-368): name, dim, heads, bias, batch, H, W, seed
CROSS_CASES = [
    ("moce_cross_c32_h2", 32, 2, True, 2, 12, 20, 61),
    ("moce_cross_c48_h1_nobias", 48, 1, False, 1, 8, 8, 62),
]


# GatedMDTA of the gated-attention experiment: name, dim, heads, bias, gate_type, batch, H, W, seed
GATED_CASES = [
    ("gated_mdta_headwise_c32_h2", 32, 2, True, "headwise", 2, 12, 16, 71),
    ("gated_mdta_elementwise_c48_h1", 48, 1, False, "elementwise", 1, 8, 8, 72),
    ("gated_mdta_none_c32_h4", 32, 4, True, None, 1, 8, 16, 73),
]


def moce_block_params(dim, heads, f, bias, ln, nblk, seed):
    """Seeded synthetic state_dict of an EncoderResidualGroup: Restormer block parameters renamed to the MoCE-IR
    attribute names (norm1 -> norms.0, norm2 -> norms.1, attn -> mixer; moce_ir.py:817-823) under ``layers.{i}.``."""
    from oracle import restormer_oracle as O
    sd = {}
    for i in range(nblk):
        p = O.synth_block_params(dim, heads, float(f), bias, ln, seed + i)
        for k, v in p.items():
            k = k.replace("norm1.", "norms.0.").replace("norm2.", "norms.1.").replace("attn.", "mixer.")
            sd[f"layers.{i}.{k}"] = v
    return sd


# Whole MoCE-IR model (moce_ir.py:1080-1231; SURVEY.md §8 f1, BASELINE configs[2]): name, constructor kwargs, batch, H, W,
# weight seed, input seed, torch.manual_seed set right before the forward (routing noise, moce_ir.py:741)
MOCEIR_SMALL = dict(dim=16, levels=4, heads=[1, 2, 2, 4], num_blocks=[1, 1, 1, 2], num_dec_blocks=[1, 1, 2],
                    ffn_expansion_factor=2, num_refinement_blocks=1, LayerNorm_type="WithBias", bias=False, rank=2,
                    num_experts=4, depth_type="constant", stage_depth=[1, 1, 1], rank_type="spread", topk=1,
                    with_complexity=False, complexity_scale="max")
# BASELINE configs[2] / SURVEY.md 8d: the MoCE_IR model of MoCE-IR-main/src/options.py:54-84
MOCEIR_CFG3 = dict(dim=48, num_blocks=[4, 6, 6, 8], num_dec_blocks=[2, 4, 4], levels=4, heads=[1, 2, 4, 8],
                   num_refinement_blocks=4, topk=1, num_experts=4, rank=2, depth_type="constant", stage_depth=[1, 1, 1],
                   rank_type="spread", complexity_scale="max")
MOCEIR_CASES = [
    ("moceir_cfg3_b1_64", MOCEIR_CFG3, 1, 64, 64, 81, 82, 83),         # the real cfg3 network at a CPU-affordable size
    ("moceir_small_b1_64", MOCEIR_SMALL, 1, 64, 64, 71, 72, 73),
    ("moceir_small_b2_64x96", MOCEIR_SMALL, 2, 64, 96, 74, 75, 76),    # batch > 1: the eval-mode expert-sum quirk
    ("moceir_small_bias_b1_32", dict(MOCEIR_SMALL, bias=True, LayerNorm_type="BiasFree", topk=2), 1, 32, 32, 77, 78, 79),
]

# training mode (SparseDispatcher + auxiliary losses): batch 2 so that the two samples can go to different experts; top-1
# and top-2 routing, with and without the complexity bias of the importance loss
MOCEIR_TRAIN_CASES = [
    ("moceir_train_small_b2_64", MOCEIR_SMALL, 2, 64, 64, 171, 172, 173),
    ("moceir_train_top2_b3_32x64", dict(MOCEIR_SMALL, bias=True, topk=2, with_complexity=True), 3, 32, 64, 174, 175, 176),
]


def synth_like(template_sd, seed):
    """Seeded synthetic state_dict with the keys / shapes of `template_sd` (a freshly constructed model's state_dict):
    LayerNorm weights near 1, biases small, temperatures near 1, conv / linear weights ~ U(-1, 1) / sqrt(fan_in); buffers
    computed by the constructor (`complexity`) and the fixed high-pass kernel keep the constructor's value."""
    import torch
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k in sorted(template_sd):
        t = template_sd[k]
        if k.endswith("complexity") or k.endswith("high_conv.0.conv.weight"):
            out[k] = t.detach().clone()
        elif k.endswith("temperature"):
            out[k] = 1.0 + 0.3 * (torch.rand(t.shape, generator=g) - 0.5)
        elif t.dim() == 1:
            if k.endswith("body.weight"):
                out[k] = 1.0 + 0.2 * (torch.rand(t.shape, generator=g) - 0.5)
            else:
                out[k] = 0.2 * (torch.rand(t.shape, generator=g) - 0.5)
        else:
            fan_in = 1
            for d in t.shape[1:]:
                fan_in *= d
            out[k] = (torch.rand(t.shape, generator=g) - 0.5) * 2.0 / fan_in ** 0.5
    return out


# RectiFormer experiment variants (experiment/RectiFormer.py:231-325): name, kind, dim, heads, ffn factor, bias, B, H, W, seed
RECTI_CASES = [
    ("recti_ffn_c32", "ffn", 32, 0, 2.66, True, 2, 16, 24, 91),
    ("recti_ffn_c48_nobias", "ffn", 48, 0, 2.0, False, 1, 8, 40, 92),
    ("recti_attn_c32_h2", "attn", 32, 2, 0, True, 2, 16, 24, 93),
    ("recti_attn_dyn_c48_h1", "attn_dyn", 48, 1, 0, False, 3, 8, 16, 94),
]
# AdaIR Chanel_Cross_Attention (AdaIR-main/net/model.py:177-216): name, dim, heads, bias, B, H, W, seed
ADAIR_CCA_CASES = [
    ("adair_cca_c32_h2", 32, 2, True, 2, 12, 20, 95),
    ("adair_cca_c48_h1_nobias", 48, 1, False, 1, 8, 8, 96),
]

# AdaIR FreModule (AdaIR-main/net/model.py:295-365): name, dim, heads, bias, B, image (Hi, Wi), feature (H, W), seed.
# H, W >= 128 so that the spectrum mask is not always empty (h // 128 > 0)
FREMODULE_CASES = [
    ("adair_fre_c16_h1_512", 16, 1, False, 1, (512, 512), (512, 512), 97),   # h // 128 = 4: a non-empty spectrum mask
    ("adair_fre_c16_h1_resize", 16, 1, True, 2, (96, 160), (128, 256), 98),
    ("adair_fre_c32_h2_small", 32, 2, False, 1, (64, 64), (64, 64), 99),     # h // 128 == 0: low = 0, high = |x|
]

# FreModule under autograd (AdaIR-main/train.py): name, dim, heads, bias, B, image (Hi, Wi), feature (H, W), seed,
# rate_scale (multiplies the synthetic rate_conv.2.weight so that the thresholds sit well away from the .int() steps).
# loss = mean(out * r) with a seeded random r; stored: loss, the mask half-sizes, d loss / d feature (strided sample +
# checksums) and every parameter gradient.
FREMODULE_TRAIN_CASES = [
    ("adair_fre_train_c16_256", 16, 1, True, 2, (128, 128), (256, 256), 197, -200.0),   # h // 128 = 2: 2x2 kept bins
    ("adair_fre_train_c16_384", 16, 1, False, 1, (96, 96), (384, 256), 199, -400.0),    # 3 x 2 steps: a larger, uneven mask
    ("adair_fre_train_c32_h2_64", 32, 2, False, 2, (64, 64), (64, 64), 198, 1.0),       # empty mask: low = 0, high = |x|
]

# The whole AdaIR network (AdaIR-main/net/model.py:375-497): name, constructor kwargs, B, H, W, seed, training golden?,
# rate_scale (as above, applied to fre3.rate_conv.2.weight).  Training: loss = L1(restored, clean) as in
# AdaIR-main/train.py:38-46; every parameter gradient stored (scaled fp16; tensors above 4096 elements as the strided
# sample `grad_sample` takes).
ADAIR_NET_CFG = dict(dim=16, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1, heads=[1, 2, 2, 4])
ADAIR_NET_CASES = [
    ("adair_net_d16_b2_64", ADAIR_NET_CFG, 2, 64, 64, 211, True, 1.0),
    ("adair_net_d16_b1_512", ADAIR_NET_CFG, 1, 512, 512, 212, False, (30.0, -400.0)),   # fre3 at 256x256: a 2x2-bin mask
]


def grad_sample(g):
    """flattened gradient, every ceil(numel / 4096)-th element."""
    f = g.reshape(-1)
    return f[::max(1, -(-f.shape[0] // 4096))]


def fre_sample(t):
    """What the FreModule goldens keep of a [B, C, H, W] tensor (the full 512x512 tensors would be 16 MB): a strided
    sample and two global checksums."""
    import numpy as np
    a = t.detach().cpu().float().numpy()
    return a[:, ::3, ::7, ::5].copy(), np.array([a.mean(dtype=np.float64), np.abs(a).mean(dtype=np.float64)])
