"""tcgen05 GEMM (LayerNorm + 1x1 conv + bias + residual) against the fp32 oracle math on bf16-rounded inputs."""
import pytest
import torch
import torch.nn.functional as F

from oracle import restormer_oracle as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


CASES = [
    # Cin, Cout, (H, W), B, bias, res, per_sample
    (48, 144, (16, 24), 2, False, False, False),
    (48, 254, (32, 32), 1, True, False, False),
    (96, 288, (16, 16), 2, False, False, False),
    (96, 510, (24, 40), 2, True, True, False),
    (127, 48, (16, 24), 2, False, True, False),
    (255, 96, (8, 40), 1, True, True, False),
    (192, 576, (8, 16), 2, False, False, False),
    (384, 1152, (8, 8), 2, True, False, False),
    (384, 2042, (8, 8), 1, False, False, False),
    (510, 192, (16, 16), 1, False, True, False),
    (1021, 384, (8, 8), 2, True, True, False),
    (384, 192, (8, 24), 2, True, False, False),
    (96, 96, (40, 40), 3, True, True, True),
    (384, 384, (8, 8), 2, False, True, True),
    (64, 128, (9, 8), 2, True, False, False),
]


@pytest.mark.parametrize("ln", ["none", "WithBias", "BiasFree"])
@pytest.mark.parametrize("Cin,Cout,HW,B,bias,res,per_sample", CASES)
def test_gemm_tc(ln, Cin, Cout, HW, B, bias, res, per_sample):
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    if per_sample and ln != "none":
        pytest.skip("per-sample weights are only used without LayerNorm")
    g = torch.Generator().manual_seed(Cin * 7 + Cout)
    H, W = HW
    x = (torch.rand(B, Cin, H, W, generator=g) * 2 - 0.5).bfloat16()
    wshape = (B, Cout, Cin) if per_sample else (Cout, Cin)
    w = (torch.rand(wshape, generator=g) - 0.5) / Cin ** 0.5
    b = (torch.rand(Cout, generator=g) - 0.5) if bias else None
    lw = 1 + 0.2 * (torch.rand(Cin, generator=g) - 0.5)
    lb = 0.1 * (torch.rand(Cin, generator=g) - 0.5)
    r = (torch.rand(B, Cout, H, W, generator=g) - 0.5).bfloat16() if res else None
    xn = x.float()
    if ln != "none":
        xn = O.layernorm(xn, lw, lb if ln == "WithBias" else None)
    if per_sample:
        ref = torch.einsum("boc,bchw->bohw", w, xn)
        if bias:
            ref = ref + b.view(1, -1, 1, 1)
    else:
        ref = F.conv2d(xn, w.view(Cout, Cin, 1, 1), b)
    if res:
        ref = ref + r.float()
    code = {"none": L.LN_NONE, "WithBias": L.LN_WITHBIAS, "BiasFree": L.LN_BIASFREE}[ln]
    xd = x.cuda()
    assert ops.gemm_tc_usable(xd, Cout)
    packed = ops.pack_conv1x1(w.cuda(), b.cuda() if bias else None, code, lw.cuda(), lb.cuda(), per_sample=per_sample)
    y = ops.gemm_tc(xd, packed, Cout, code, r.cuda() if res else None, out_dtype=torch.float32)
    torch.cuda.synchronize()
    y = y.cpu()
    # weights are rounded to bf16 inside the tensor-core path: relative error ~2^-9/sqrt(3) per product
    assert rel_err(y, ref) < 4e-3, rel_err(y, ref)
    # and against the CUDA-core kernel on identical inputs (same bf16 activations, fp32 weights)
    y2 = ops.ln_conv1x1(xd, w.cuda(), b.cuda() if bias else None, code, lw.cuda(), lb.cuda(), r.cuda() if res else None,
                        out_dtype=torch.float32, per_sample_w=per_sample).cpu()
    assert rel_err(y, y2) < 4e-3


def test_gemm_tc_large_tail_and_multi_cta():
    """Many tiles per CTA, HW not a multiple of 128, bf16 output."""
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    g = torch.Generator().manual_seed(5)
    B, Cin, Cout, H, W = 3, 96, 288, 200, 168   # HW = 33600 = 262.5 tiles
    x = (torch.rand(B, Cin, H, W, generator=g) * 2 - 1).bfloat16()
    w = (torch.rand(Cout, Cin, generator=g) - 0.5) / Cin ** 0.5
    lw = 1 + 0.2 * (torch.rand(Cin, generator=g) - 0.5)
    lb = 0.1 * (torch.rand(Cin, generator=g) - 0.5)
    ref = F.conv2d(O.layernorm(x.float(), lw, lb), w.view(Cout, Cin, 1, 1))
    packed = ops.pack_conv1x1(w.cuda(), None, L.LN_WITHBIAS, lw.cuda(), lb.cuda())
    y = ops.gemm_tc(x.cuda(), packed, Cout, L.LN_WITHBIAS).float().cpu()
    assert rel_err(y, ref) < 6e-3


def test_gemm_tc_channels_on_lanes_epilogue_matches_default():
    """The opt-in channels-on-lanes epilogue (variant GEMM_RESIDENT_CH, gemm_tc_kernel MODE 4) gives the same bf16 results as
    the default pixels-on-lanes epilogue (identical fp32 accumulators, identical rounding point)."""
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    g = torch.Generator().manual_seed(77)
    for (B, Cin, Cout, H, W, ln, res) in [(2, 96, 510, 16, 24, L.LN_WITHBIAS, False), (1, 255, 96, 16, 16, L.LN_NONE, True),
                                          (3, 48, 144, 8, 40, L.LN_BIASFREE, False), (2, 384, 1152, 8, 8, L.LN_WITHBIAS, False)]:
        x = (torch.rand(B, Cin, H, W, generator=g) * 2 - 1).bfloat16().cuda()
        w = ((torch.rand(Cout, Cin, generator=g) - 0.5) / Cin ** 0.5).cuda()
        bias = (torch.rand(Cout, generator=g) - 0.5).cuda()
        lw, lb = (torch.rand(Cin, generator=g) + 0.5).cuda(), (torch.rand(Cin, generator=g) - 0.5).cuda()
        r = (torch.rand(B, Cout, H, W, generator=g) - 0.5).bfloat16().cuda() if res else None
        packed = ops.pack_conv1x1(w, bias, ln, lw if ln else None, lb if ln == L.LN_WITHBIAS else None)
        y0 = ops.gemm_tc(x, packed, Cout, ln, r, variant=L.GEMM_RESIDENT)
        y1 = ops.gemm_tc(x, packed, Cout, ln, r, variant=L.GEMM_RESIDENT_CH)
        y1f = ops.gemm_tc(x, packed, Cout, ln, r, out_dtype=torch.float32, variant=L.GEMM_RESIDENT_CH)
        torch.cuda.synchronize()
        assert (y0.float() - y1.float()).abs().max() <= 2e-2 * y0.float().abs().max()
        assert ((y1f - y0.float()).norm() / y0.float().norm()) < 4e-3


def test_gemm_stream_matches_resident_weights_kernel():
    """Deep-level shapes go to the streamed-operand kernel (gemm_stream_tc.cu); variant GEMM_RESIDENT forces the
    resident-weight kernel.  Same packed operands, same fp32 accumulation -> results agree to bf16 rounding; ragged pixel
    tiles, several items per CTA, per-sample weights, all LayerNorm kinds, residual, both output dtypes."""
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    g = torch.Generator().manual_seed(99)
    cases = [(3, 384, 2042, 40, 40, L.LN_WITHBIAS, False, False), (2, 1021, 384, 24, 40, L.LN_NONE, True, False),
             (2, 384, 1152, 25, 40, L.LN_BIASFREE, False, False), (3, 384, 384, 32, 32, L.LN_NONE, True, True),
             (1, 192, 1020, 64, 72, L.LN_WITHBIAS, False, False)]
    for (B, Cin, Cout, H, W, ln, res, per_sample) in cases:
        x = (torch.rand(B, Cin, H, W, generator=g) * 2 - 0.5).bfloat16().cuda()
        wshape = (B, Cout, Cin) if per_sample else (Cout, Cin)
        w = ((torch.rand(wshape, generator=g) - 0.5) / Cin ** 0.5).cuda()
        bias = (torch.rand(Cout, generator=g) - 0.5).cuda()
        lw, lb = (torch.rand(Cin, generator=g) + 0.5).cuda(), (torch.rand(Cin, generator=g) - 0.5).cuda()
        r = (torch.rand(B, Cout, H, W, generator=g) - 0.5).bfloat16().cuda() if res else None
        packed = ops.pack_conv1x1(w, bias, ln, lw if ln else None, lb if ln == L.LN_WITHBIAS else None,
                                  per_sample=per_sample)
        y0 = ops.gemm_tc(x, packed, Cout, ln, r, out_dtype=torch.float32, variant=L.GEMM_RESIDENT)
        y1 = ops.gemm_tc(x, packed, Cout, ln, r, out_dtype=torch.float32, variant=L.GEMM_STREAM)
        y1b = ops.gemm_tc(x, packed, Cout, ln, r, variant=L.GEMM_STREAM)
        torch.cuda.synchronize()
        scale = y0.abs().max()
        assert (y0 - y1).abs().max() <= 1e-4 * scale, (Cin, Cout, float((y0 - y1).abs().max()), float(scale))
        assert (y1b.float() - y1).abs().max() <= 1e-2 * scale


@pytest.mark.parametrize("Cout,Cin", [(288, 96), (510, 96), (96, 255), (2042, 384), (384, 1021), (48, 127), (144, 48)])
def test_pack_conv1x1_t_equals_pack_of_transpose(Cout, Cin):
    """b200ir_pack_conv1x1_t (operand of dx = W^T.dy, packed straight from w) == b200ir_pack_conv1x1(w^T), bit for bit,
    including the zero padding."""
    from b200ir import ops
    g = torch.Generator().manual_seed(Cout + Cin)
    w = (torch.rand(Cout, Cin, generator=g) - 0.5).cuda()
    wp0, st0 = ops.pack_conv1x1(w.t().contiguous())
    wp1, st1 = ops.pack_conv1x1_t(w)
    torch.cuda.synchronize()
    assert wp0.shape == wp1.shape
    assert torch.equal(wp0.view(torch.int16), wp1.view(torch.int16))
    assert torch.equal(st0[1], st1[1])   # t = 0 (no bias); s (row sums) is only read by the LayerNorm epilogue


def test_prepacker_fills_caches_with_identical_packs():
    """Prepacker.run(): one launch packs the forward (LayerNorm-folded) and transposed operands of every block conv;
    results are bitwise what pack_conv1x1 / pack_conv1x1_t give, and the forward / backward then find them cached."""
    import b200ir
    from b200ir import ops, autograd_ops as A, modules as M
    L = b200ir._lib
    torch.manual_seed(0)
    blk = b200ir.TransformerBlock(48, 2, 2.66, True, "WithBias").cuda()
    enc = b200ir.EncoderBlock(32, 1, 2.0, False, "BiasFree").cuda()
    net = torch.nn.Sequential(blk)
    for p in list(blk.parameters()) + list(enc.parameters()):
        p.data.add_(0.1 * torch.randn_like(p))
    for model, (n1, attn, n2, ffn) in ((net, (blk.norm1, blk.attn, blk.norm2, blk.ffn)),
                                       (enc, (enc.norms[0], enc.mixer, enc.norms[1], enc.ffn))):
        pre = A.Prepacker(model)
        assert pre.ndesc == 6 and pre.run()
        for cache, key, conv, norm in ((attn._b200ir_cache, ("qkv", id(n1)), attn.qkv, n1),
                                       (ffn._b200ir_cache, ("in", id(n2)), ffn.project_in, n2),
                                       (ffn._b200ir_cache, "out", ffn.project_out, None)):
            cout, cin = conv.weight.shape[:2]
            ln_type, lw, lb = M._ln_args(norm) if norm is not None else (L.LN_NONE, None, None)
            wp0, st0 = ops.pack_conv1x1(conv.weight.detach().reshape(cout, cin), conv.bias.detach() if conv.bias is not None else None,
                                        ln_type, lw.detach() if lw is not None else None, lb.detach() if lb is not None else None)
            wp1, st1 = cache[key][1]
            assert torch.equal(wp0.view(torch.int16), wp1.view(torch.int16)) and torch.equal(st0, st1), key
            wt0, _ = ops.pack_conv1x1_t(conv.weight.detach().reshape(cout, cin))
            assert torch.equal(wt0.view(torch.int16), cache[("T", key)][1][0].view(torch.int16)), key
        # a forward + backward after run() packs nothing itself
        x = (torch.rand(2, attn.qkv.in_channels, 16, 16, device="cuda") - 0.5).bfloat16().requires_grad_(True)
        ops.timing_start(only=["pack_conv1x1"])
        model(x).float().sum().backward()
        torch.cuda.synchronize()
        left = [k for k in ops.timing_stop() if k[1] == 1]     # (name, Bw, Cin, Cout): per-sample MDTA packs (Bw > 1) stay
        assert not left, ("blocks still packed shared weights one by one", left)
