"""Input pipeline / metrics: the oracle against outputs of the reference's own functions (CPU), the
device kernels against the oracle (GPU)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import pipeline_oracle as PO

P = 16


def _gold():
    return np.load(os.path.join(GOLDEN, "pipeline_aug.npz"))


def test_pipeline_oracle_vs_reference_golden():
    g = _gold()
    img = g["img"]
    for i, (oy, ox) in enumerate(g["offs"]):
        patch = PO.crop_patch(img, int(oy), int(ox), P)
        for mode in range(8):
            aug = PO.data_augmentation(patch, mode)
            assert np.array_equal(aug, g[f"aug_{i}_{mode}"])
            for sigma in (15, 25, 50):
                np.random.seed(1000 + 10 * mode + i)
                noise = np.random.randn(*aug.shape)
                assert np.array_equal(PO.add_gaussian_noise(aug, sigma, noise), g[f"noisy_{i}_{mode}_{sigma}"])


def test_ssim_oracle_properties():
    """No scikit-image here (parity unpinned): the restated SSIM must at least satisfy the definition's invariants."""
    rng = np.random.RandomState(0)
    a = rng.rand(40, 48, 3)
    assert abs(PO.ssim(a, a) - 1.0) < 1e-12
    b = np.clip(a + 0.1 * rng.randn(*a.shape), 0, 1)
    s1, s2 = PO.ssim(a, b), PO.ssim(b, a)
    assert abs(s1 - s2) < 1e-12 and 0.0 < s1 < 1.0
    assert PO.ssim(a, np.clip(a + 0.3 * rng.randn(*a.shape), 0, 1)) < s1
    assert abs(PO.psnr(a, b) - 10 * np.log10(1.0 / np.mean((a - b) ** 2))) < 1e-9


@pytest.mark.gpu
def test_crop_augment_degrade_kernel_bit_exact():
    """Byte work: the device pipeline reproduces the reference's uint8 results exactly for every mode / noise level."""
    import b200ir
    g = _gold()
    img = g["img"]
    offs = g["offs"]
    B = len(offs)
    src = torch.from_numpy(np.stack([img] * B)).cuda()
    for mode in range(8):
        for sigma in (15, 25, 50):
            noises = []
            for i in range(B):
                np.random.seed(1000 + 10 * mode + i)
                noises.append(np.random.randn(P, P, 3))
            noise = torch.from_numpy(np.stack(noises).astype(np.float32)).cuda()
            clean, deg = b200ir.crop_augment_degrade(src, torch.from_numpy(offs).cuda(),
                                                     torch.full((B,), mode, dtype=torch.int32, device="cuda"), P,
                                                     sigma=torch.full((B,), float(sigma), device="cuda"), noise=noise)
            for i in range(B):
                want_c = PO.to_tensor(g[f"aug_{i}_{mode}"])
                assert np.array_equal(clean[i].cpu().numpy(), want_c)
                # the reference adds float64 noise; fp32 noise may land on the other side of an integer boundary in rare
                # pixels (|clean + noise*sigma - n| < 1e-5): allow a handful of off-by-one bytes
                got = np.rint(deg[i].cpu().numpy() * 255).astype(np.int32)
                want = g[f"noisy_{i}_{mode}_{sigma}"].transpose(2, 0, 1).astype(np.int32)
                d = np.abs(got - want)
                assert d.max() <= 1 and (d > 0).sum() <= 2, (mode, sigma, i, int(d.max()), int((d > 0).sum()))


@pytest.mark.gpu
@pytest.mark.parametrize("B,H,W", [(2, 40, 48), (1, 128, 96), (3, 11, 30)])
def test_psnr_ssim_kernel_vs_oracle(B, H, W):
    import b200ir
    rng = np.random.RandomState(H + W)
    clean = rng.rand(B, 3, H, W).astype(np.float32)
    rest = (clean + 0.08 * rng.randn(B, 3, H, W)).astype(np.float32)       # not clamped: the kernel clamps (test.py:104)
    out = b200ir.psnr_ssim(torch.from_numpy(rest).cuda(), torch.from_numpy(clean).cuda()).cpu().numpy()
    for i in range(B):
        c, r = clean[i].transpose(1, 2, 0), np.clip(rest[i], 0, 1).transpose(1, 2, 0)
        assert abs(out[i, 0] - PO.psnr(c, r)) < 1e-3
        assert abs(out[i, 1] - PO.ssim(c, r)) < 2e-5


def test_tile_oracle_identity_and_index_lists():
    """oracle.tile_restore_reference with fn = identity gives the clamped frame back (every pixel is covered, overlapping
    copies average to themselves), and the index lists are the reference's: range(0, h - tile, stride) + [h - tile]."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("b200ir_pipeline_src", os.path.join(os.path.dirname(os.path.dirname(
        os.path.abspath(__file__))), "image-restoration_b200", "pipeline.py"))
    src = open(spec.origin).read()
    ns = {}
    exec(src[src.index("def tile_index_lists"):src.index("@torch.no_grad()")], ns)     # the pure-Python helper only
    assert ns["tile_index_lists"](200, 264, 64, 16) == (64, [0, 48, 96, 136], [0, 48, 96, 144, 192, 200])
    assert ns["tile_index_lists"](96, 40, 128, 0) == (40, [0, 40, 56], [0])
    with pytest.raises(ValueError):
        ns["tile_index_lists"](100, 100, 100, 0)
    g = torch.Generator().manual_seed(3)
    x = torch.rand(2, 3, 72, 104, generator=g) * 1.2 - 0.1
    y = PO.tile_restore_reference(lambda p: p, x, tile=32, tile_overlap=8)
    assert (y - x.clamp(0, 1)).abs().max() <= 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,tile,overlap", [(200, 264, 64, 16), (96, 40, 128, 0), (128, 128, 64, 0)])
def test_tiled_restore_vs_oracle(H, W, tile, overlap):
    """b200ir.tiled_restore (gather kernel, batched tiles through the network, pull-form blend kernel) against the
    reference's tile recipe run tile by tile (oracle/pipeline_oracle.py tile_restore_reference) with the same network."""
    import b200ir
    from oracle import restormer_oracle as O
    net = b200ir.Restormer(dim=16, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1, heads=[1, 2, 2, 4])
    net.load_state_dict(O.synth_state_dict(dict(dim=16, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1, heads=[1, 2, 2, 4]),
                                           seed=4), strict=True)
    net = net.cuda().eval()
    x = O.synth_image(2, H, W, seed=12)
    with torch.no_grad():
        ref = PO.tile_restore_reference(lambda p: net(p.contiguous().cuda()).cpu(), x, tile=tile, tile_overlap=overlap)
        got = b200ir.tiled_restore(net, x.cuda(), tile=tile, tile_overlap=overlap, tile_batch=5).cpu()
        ident = b200ir.tiled_restore(lambda p: p, x.cuda() * 1.2 - 0.1, tile=tile, tile_overlap=overlap).cpu()
    assert got.shape == ref.shape == x.shape
    assert (got - ref).abs().max() <= 2e-5, float((got - ref).abs().max())
    assert (ident - (x * 1.2 - 0.1).clamp(0, 1)).abs().max() <= 1e-6
