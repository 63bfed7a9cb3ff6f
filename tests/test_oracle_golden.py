"""Pin the CPU oracle (oracle/restormer_oracle.py) against outputs of the UNMODIFIED reference
(tests/golden/*.npz, produced by oracle/gen_golden.py from /root/reference/Restormer.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import BLOCK_CASES, NET_CASES

TOL = 2e-5  # fp32 CPU, same ATen kernels, different op grouping (reshape vs einops copies)


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.mark.parametrize("case", BLOCK_CASES, ids=[c[0] for c in BLOCK_CASES])
def test_block_against_reference_output(case):
    name, dim, heads, f, bias, ln, b, h, w, seed = case
    g = _load(name)
    p = O.synth_block_params(dim, heads, f, bias, ln, seed)
    x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
    n1 = O.layernorm(x, p["norm1.body.weight"], p.get("norm1.body.bias"))
    assert np.abs(n1.numpy() - g["y_norm1"]).max() <= TOL
    assert np.abs(O.attention(n1, p, "attn.", heads).numpy() - g["y_attn"]).max() <= TOL
    n2 = O.layernorm(x, p["norm2.body.weight"], p.get("norm2.body.bias"))
    assert np.abs(O.feedforward(n2, p, "ffn.").numpy() - g["y_ffn"]).max() <= TOL
    assert np.abs(O.transformer_block(x, p, "", heads).numpy() - g["y_block"]).max() <= TOL


@pytest.mark.parametrize("name,fn,c,shape", [
    ("down_c48", O.downsample, 48, (24, 48, 3, 3)),
    ("up_c96", O.upsample, 96, (192, 96, 3, 3)),
    ("embed_3_48", O.overlap_patch_embed, 3, (48, 3, 3, 3)),
])
def test_glue_against_reference_output(name, fn, c, shape):
    gen = torch.Generator().manual_seed(21)
    wt = O.synth_tensor(shape, "conv", gen)
    key = "proj.weight" if name.startswith("embed") else "body.0.weight"
    x = O.synth_image(2, 8, 12, seed=22, c=c)
    y = fn(x, {key: wt}, "")
    assert np.abs(y.numpy() - _load(name)["y"]).max() <= TOL


@pytest.mark.parametrize("case", NET_CASES, ids=[c[0] for c in NET_CASES])
def test_net_against_reference_output(case):
    name, cfg, b, h, w, wseed, iseed = case
    full = dict(O.DEFAULT_CFG)
    full.update(cfg)
    sd = O.synth_state_dict(cfg, seed=wseed)
    x = O.synth_image(b, h, w, seed=iseed, c=full["inp_channels"])
    y = O.restormer_forward(sd, x, cfg)
    assert np.abs(y.numpy() - _load(name)["y"]).max() <= 1e-4


def test_param_inventory_default():
    shapes = O.restormer_param_shapes()
    assert len(shapes) == 494  # SURVEY.md §8a a19 [probed]
    n = sum(int(np.prod(s)) for _, s, _ in shapes)
    assert n == 26_126_644


def test_fp64_oracle_close_to_fp32():
    p = O.synth_block_params(48, 1, 2.66, False, "WithBias", 11)
    x = O.synth_image(1, 8, 8, seed=5, c=48)
    y32 = O.transformer_block(x, p, "", 1)
    y64 = O.transformer_block(x.double(), {k: v.double() for k, v in p.items()}, "", 1)
    assert (y32.double() - y64).abs().max() < 1e-5


# ---- backward / training-step oracle pinned against the reference's own autograd + torch.optim.AdamW ----------------
def test_oracle_block_grads_match_reference():
    import numpy as np
    from oracle.gen_golden_cases import TRAIN_BLOCK_CASES
    for name, dim, heads, f, bias, ln, b, h, w, seed in TRAIN_BLOCK_CASES:
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))
        p = O.synth_block_params(dim, heads, f, bias, ln, seed)
        x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200))
        _, dx, dp = O.block_grads(x, p, heads, gy)
        ref = torch.from_numpy(gold["dx"])
        assert (dx - ref).abs().max() <= 1e-5 * ref.abs().max().clamp_min(1)
        assert {"d_" + n for n in dp} == {k for k in gold.files if k != "dx"}
        for n, g in dp.items():
            ref = torch.from_numpy(gold["d_" + n])
            assert (g - ref).abs().max() <= 2e-5 * ref.abs().max().clamp_min(1e-3), (name, n)


def test_oracle_train_step_matches_reference():
    import numpy as np
    from oracle.gen_golden_cases import TRAIN_NET_CASE
    name, cfg, b, h, w, wseed, iseed, steps = TRAIN_NET_CASE
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    clean = O.synth_image(b, h, w, seed=iseed)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(iseed + 1)) * (25.0 / 255.0)
    degraded = (clean + noise).clamp(0, 1)
    losses, grads, new_sd = O.train_step_reference(O.synth_state_dict(cfg, seed=wseed), cfg, degraded, clean, steps=steps)
    assert np.allclose(losses, gold["losses"], rtol=0, atol=1e-6)
    for n in grads:
        ref = torch.from_numpy(gold["g_" + n])
        assert (grads[n] - ref).abs().max() <= 1e-4 * ref.abs().max().clamp_min(1e-6), n
        assert (new_sd[n] - torch.from_numpy(gold["p_" + n])).abs().max() <= 1e-6, n


def test_cross_attention_oracle_vs_reference_golden():
    """oracle.cross_attention (restated from moce_ir.py:345-368) against outputs of the reference's own CrossAttention
    module stored by oracle/gen_golden_moce.py."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle import restormer_oracle as O
    from oracle.gen_golden_cases import CROSS_CASES
    for name, dim, heads, bias, b, h, w, seed in CROSS_CASES:
        p = O.synth_cross_attention_params(dim, heads, bias, seed)
        x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
        y = O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1
        out = O.cross_attention(x, y, p, "", heads)
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))["y"]
        assert np.abs(out.numpy() - gold).max() <= 1e-6


def test_gated_attention_oracle_vs_reference_golden():
    """oracle.gated_attention (restated from the gated-attention experiment's GatedMDTA.forward) against the stored
    outputs of the reference module (oracle/gen_golden_gated.py)."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle import restormer_oracle as O
    from oracle.gen_golden_cases import GATED_CASES
    for name, dim, heads, bias, gate_type, b, h, w, seed in GATED_CASES:
        p = O.synth_gated_attention_params(dim, heads, bias, gate_type, seed)
        x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
        out = O.gated_attention(x, p, "", heads, gate_type)
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))["y"]
        assert np.abs(out.numpy() - gold).max() <= 1e-6
