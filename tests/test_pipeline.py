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
