"""tcgen05 3x3 conv (9 tap-shifted GEMMs) + PixelUnshuffle/PixelShuffle store vs the oracle math."""
import pytest
import torch
import torch.nn.functional as F

from oracle import restormer_oracle as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("mode,Cin,Cout,B,H,W,bias", [
    ("unshuffle", 48, 24, 2, 16, 64, False),
    ("unshuffle", 96, 48, 1, 32, 128, False),
    ("unshuffle", 192, 96, 2, 8, 16, False),
    ("shuffle", 384, 768, 2, 8, 8, False),
    ("shuffle", 192, 384, 1, 16, 24, False),
    ("shuffle", 96, 192, 2, 24, 72, False),
    ("none", 48, 48, 2, 10, 40, True),
    ("none", 32, 16, 1, 6, 200, True),
    ("unshuffle", 64, 32, 3, 12, 136, False),
    ("none", 256, 320, 1, 16, 32, True),        # streamed-weights kernel (conv3x3_stream_tc.cu), bias, Cout > 256
    ("shuffle", 384, 768, 1, 32, 32, False),
    ("unshuffle", 200, 96, 2, 20, 48, True),     # Cin not a multiple of 64, ragged tiles
    ("none", 768, 384, 2, 16, 16, False),        # backward of Upsample 384->768: no resident slice fits, stream-only
    ("none", 384, 192, 1, 32, 32, True),
])
def test_conv3x3_tc(mode, Cin, Cout, B, H, W, bias):
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    g = torch.Generator().manual_seed(Cin + Cout + H)
    x = (torch.rand(B, Cin, H, W, generator=g) * 2 - 1).bfloat16()
    w = (torch.rand(Cout, Cin, 3, 3, generator=g) - 0.5) / (Cin * 9) ** 0.5
    b = (torch.rand(Cout, generator=g) - 0.5) if bias else None
    ref = F.conv2d(x.float(), w, b, padding=1)
    if mode == "unshuffle":
        ref = O.pixel_unshuffle2(ref)
    elif mode == "shuffle":
        ref = O.pixel_shuffle2(ref)
    code = {"none": L.SHUFFLE_NONE, "unshuffle": L.UNSHUFFLE2, "shuffle": L.SHUFFLE2}[mode]
    xd = x.cuda()
    assert ops.conv3x3_tc_usable(xd, Cout)
    packed = ops.pack_conv3x3(w.cuda(), b.cuda() if bias else None)
    y = ops.conv3x3_tc(xd, packed, Cout, code, out_dtype=torch.float32)
    torch.cuda.synchronize()
    assert rel_err(y.cpu(), ref) < 4e-3
    yb = ops.conv3x3_tc(xd, packed, Cout, code).float().cpu()
    assert rel_err(yb, ref) < 6e-3
