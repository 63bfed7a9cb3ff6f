"""CPU oracle of the reference's input pipeline and evaluation metrics.  TEST INFRASTRUCTURE ONLY:
imported by tests/ (and nothing in the product package); numpy restatement, each function citing the reference lines.

Pinned: ``data_augmentation`` / crop / noise / ToTensor against outputs of the UNMODIFIED reference functions
(``oracle/gen_golden_pipeline.py`` imports MoCE-IR-main/src/utils/image_utils.py and data/degradation_utils.py;
tests/golden/pipeline_aug.npz).  **Parity unpinned** for PSNR / SSIM: the arithmetic lives in scikit-image
(``skimage.metrics``; unpinned in MoCE-IR-main/requirements.txt, AdaIR-main/env.yaml pins scikit-image==0.19.3), which is
not installed here, so ``ssim`` restates its published algorithm (Wang et al. 2004 as implemented by
``structural_similarity(gaussian_weights=True)``: sigma 1.5, truncate 3.5 -> 11x11 window, ``use_sample_covariance=True``
-> covariances x NP/(NP-1) with NP = 121, K1 0.01, K2 0.03, mean over the image cropped by 5 pixels, channels averaged)
on ``scipy.ndimage.gaussian_filter`` — the same filter skimage calls.
"""
import numpy as np


def data_augmentation(image: np.ndarray, mode: int) -> np.ndarray:
    """MoCE-IR-main/src/utils/image_utils.py:136-166 (HWC array)."""
    if mode == 0:
        return image
    if mode == 1:
        return np.flipud(image)
    if mode == 2:
        return np.rot90(image)
    if mode == 3:
        return np.flipud(np.rot90(image))
    if mode == 4:
        return np.rot90(image, k=2)
    if mode == 5:
        return np.flipud(np.rot90(image, k=2))
    if mode == 6:
        return np.rot90(image, k=3)
    if mode == 7:
        return np.flipud(np.rot90(image, k=3))
    raise Exception('Invalid choice of image transformation')


def crop_patch(img: np.ndarray, ind_h: int, ind_w: int, patch: int) -> np.ndarray:
    """dataset_utils.py:122-133 (_crop_patch with the random indices given)."""
    return img[ind_h:ind_h + patch, ind_w:ind_w + patch]


def add_gaussian_noise(clean_patch: np.ndarray, sigma: float, noise: np.ndarray) -> np.ndarray:
    """degradation_utils.py:21-24 with the np.random.randn(*shape) samples passed in."""
    return np.clip(clean_patch + noise * sigma, 0, 255).astype(np.uint8)


def to_tensor(img_u8: np.ndarray) -> np.ndarray:
    """torchvision ToTensor on a uint8 HWC array (dataset_utils.py:185-186): CHW float32 / 255."""
    return np.ascontiguousarray(img_u8.transpose(2, 0, 1)).astype(np.float32) / 255.0


def psnr(clean: np.ndarray, restored: np.ndarray) -> float:
    """skimage peak_signal_noise_ratio(clean, restored, data_range=1) (test.py:46-49,108)."""
    mse = np.mean((clean.astype(np.float64) - restored.astype(np.float64)) ** 2)
    return float("inf") if mse == 0 else 10.0 * np.log10(1.0 / mse)


def ssim(clean: np.ndarray, restored: np.ndarray) -> float:
    """structural_similarity(clean, restored, channel_axis=2, gaussian_weights=True, data_range=1.0) (test.py:51-52), HWC."""
    from scipy.ndimage import gaussian_filter
    K1, K2, sigma, truncate, win = 0.01, 0.03, 1.5, 3.5, 11
    C1, C2 = K1 ** 2, K2 ** 2
    cov_norm = win ** 2 / (win ** 2 - 1.0)
    vals = []
    for ch in range(clean.shape[2]):
        X, Y = clean[..., ch].astype(np.float64), restored[..., ch].astype(np.float64)
        f = lambda a: gaussian_filter(a, sigma=sigma, truncate=truncate, mode="reflect")
        ux, uy = f(X), f(Y)
        vx, vy, vxy = cov_norm * (f(X * X) - ux * ux), cov_norm * (f(Y * Y) - uy * uy), cov_norm * (f(X * Y) - ux * uy)
        S = ((2 * ux * uy + C1) * (2 * vxy + C2)) / ((ux ** 2 + uy ** 2 + C1) * (vx + vy + C2))
        pad = (win - 1) // 2
        vals.append(S[pad:-pad, pad:-pad].mean())
    return float(np.mean(vals))


def tile_restore_reference(fn, x, tile=128, tile_overlap=0):
    """Tile helper of the reference (20260104_CG_IR/src/data/dataset_utils.py:513-534; same code in
    MEASNet-main/MEASNet/utils/dataset_utils.py:226-247 and AdaptIR-main/utils/dataset_utils.py:346-367), with the network
    call the reference's copies leave out (``out_patch = in_patch``) supplied as ``fn`` and its last two lines completed the
    way the recipe ends everywhere it is used (``restored = E.div_(W)``, then the reference's own ``torch.clamp(restored, 0,
    1)``).  x: torch fp32 [B, C, H, W] on the CPU; fn: patch -> restored patch.  Parity note: the reference's helper is dead
    code (``restored`` is never assigned), so this pins the index lists and the accumulate order, not an output."""
    import torch
    b, c, h, w = x.shape
    tile = min(tile, h, w)
    assert tile % 8 == 0, "tile size should be multiple of 8"
    stride = tile - tile_overlap
    h_idx_list = list(range(0, h - tile, stride)) + [h - tile]
    w_idx_list = list(range(0, w - tile, stride)) + [w - tile]
    E = W = None
    for h_idx in h_idx_list:
        for w_idx in w_idx_list:
            out_patch = fn(x[..., h_idx:h_idx + tile, w_idx:w_idx + tile])
            if E is None:
                E = torch.zeros(b, out_patch.shape[1], h, w, dtype=x.dtype)
                W = torch.zeros_like(E)
            E[..., h_idx:h_idx + tile, w_idx:w_idx + tile].add_(out_patch)
            W[..., h_idx:h_idx + tile, w_idx:w_idx + tile].add_(torch.ones_like(out_patch))
    return torch.clamp(E.div_(W), 0, 1)
