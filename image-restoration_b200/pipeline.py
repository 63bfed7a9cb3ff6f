"""Input pipeline and evaluation metrics on the device, same semantics as the reference's PIL / numpy /
skimage code: ``crop_augment_degrade`` = ``_crop_patch`` + ``random_augmentation`` + ``Degradation._add_gaussian_noise`` +
``ToTensor`` (MoCE-IR-main/src/data/dataset_utils.py:122-186, utils/image_utils.py:136-185, data/degradation_utils.py:21-24);
``psnr_ssim`` = ``torch.clamp(restored, 0, 1)`` + skimage ``peak_signal_noise_ratio`` / ``structural_similarity(
gaussian_weights=True, data_range=1)`` (MoCE-IR-main/src/test.py:46-52,104-112).  Random choices (crop offsets, the
augmentation mode 1..7, the noise samples) come from the caller so that any generator can drive them."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib as L
from . import ops

lib = L.lib


def crop_augment_degrade(src: torch.Tensor, offsets: torch.Tensor, modes: torch.Tensor, patch: int,
                         sigma: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None):
    """src uint8 CUDA [B, Hs, Ws, 3]; offsets int32 [B, 2] (row, col); modes int32 [B] in 0..7; sigma fp32 [B] on the 0..255
    scale (15 / 25 / 50 in the reference) and noise fp32 [B, patch, patch, 3] ~ N(0, 1), both optional.
    -> (clean, degraded) fp32 [B, 3, patch, patch] in [0, 1] (degraded is None without noise)."""
    if src.dtype != torch.uint8 or not src.is_cuda or src.dim() != 4 or src.shape[-1] != 3 or not src.is_contiguous():
        raise ValueError("crop_augment_degrade: src must be a contiguous uint8 CUDA tensor [B, H, W, 3]")
    B, Hs, Ws, _ = src.shape
    for t, n, dt in ((offsets, 2 * B, torch.int32), (modes, B, torch.int32)):
        if t.dtype != dt or not t.is_cuda or t.numel() != n or not t.is_contiguous():
            raise ValueError("crop_augment_degrade: offsets [B,2] / modes [B] must be contiguous int32 CUDA tensors")
    if (sigma is None) != (noise is None):
        raise ValueError("crop_augment_degrade: sigma and noise go together")
    clean = torch.empty((B, 3, patch, patch), device=src.device, dtype=torch.float32)
    degraded = torch.empty_like(clean) if noise is not None else None
    if noise is not None and (noise.dtype != torch.float32 or tuple(noise.shape) != (B, patch, patch, 3) or not noise.is_contiguous()):
        raise ValueError("crop_augment_degrade: noise must be contiguous fp32 [B, patch, patch, 3]")
    ops.LAUNCHES += 1
    rc = lib.b200ir_crop_augment_degrade(src.data_ptr(), Hs, Ws, offsets.data_ptr(), modes.data_ptr(),
                                         ops._par(sigma, "sigma", B), noise.data_ptr() if noise is not None else None,
                                         clean.data_ptr(), degraded.data_ptr() if degraded is not None else None, B, patch,
                                         ops._stream())
    L.check(rc, "b200ir_crop_augment_degrade")
    return clean, degraded


def psnr_ssim(restored: torch.Tensor, clean: torch.Tensor) -> torch.Tensor:
    """fp32 [B, 2] = (PSNR dB, SSIM) per image of clamp(restored, 0, 1) vs clean (fp32 CUDA NCHW)."""
    if restored.shape != clean.shape or restored.dim() != 4:
        raise ValueError("psnr_ssim: two NCHW tensors of the same shape expected")
    r, c = restored.float().contiguous(), clean.float().contiguous()
    if not r.is_cuda:
        raise ValueError("psnr_ssim: CUDA tensors expected (no CPU fallback)")
    B, C, H, W = r.shape
    nblk = lib.b200ir_psnr_ssim_nblocks(H, W)
    ws = torch.empty(B * (1 + C) * nblk, device=r.device, dtype=torch.float32)
    out = torch.empty((B, 2), device=r.device, dtype=torch.float32)
    ops.LAUNCHES += 3
    L.check(lib.b200ir_psnr_ssim(r.data_ptr(), c.data_ptr(), ws.data_ptr(), out.data_ptr(), B, C, H, W, ops._stream()),
            "b200ir_psnr_ssim")
    return out


def tile_index_lists(h: int, w: int, tile: int, tile_overlap: int):
    """(tile, h_idx_list, w_idx_list) exactly as the reference's tile helper builds them
    (20260104_CG_IR/src/data/dataset_utils.py:515-521): tile = min(tile, h, w), stride = tile - overlap, start offsets
    range(0, h - tile, stride) plus the last one that ends at the border."""
    tile = min(tile, h, w)
    if tile % 8:
        raise ValueError("tile size should be multiple of 8")
    stride = tile - tile_overlap
    if stride <= 0:
        raise ValueError("tile_overlap must be smaller than the tile")
    return tile, list(range(0, h - tile, stride)) + [h - tile], list(range(0, w - tile, stride)) + [w - tile]


@torch.no_grad()
def tiled_restore(net, x: torch.Tensor, tile: int = 128, tile_overlap: int = 0, tile_batch: int = 16,
                  clamp: bool = True) -> torch.Tensor:
    """Full-frame inference for frames too large for one pass: the frame is cut into overlapping ``tile`` x ``tile``
    patches, ``net`` restores them ``tile_batch`` at a time, and the patches are averaged where they overlap
    (``E[.., h:h+t, w:w+t] += patch; W += 1; restored = clamp(E / W, 0, 1)``: the reference's tile helper,
    20260104_CG_IR/src/data/dataset_utils.py:513-534, MEASNet-main/MEASNet/utils/dataset_utils.py:226-247).
    x fp32 CUDA [B, C, H, W]; -> fp32 [B, C_out, H, W].  Gather and blend are libb200ir kernels (csrc/pipeline_ops.cu)."""
    if x.dim() != 4 or not x.is_cuda or x.dtype != torch.float32:
        raise ValueError("tiled_restore: fp32 CUDA NCHW input expected (no CPU fallback)")
    x = x.contiguous()
    B, C, H, W = x.shape
    t, hs, ws = tile_index_lists(H, W, tile, tile_overlap)
    nh, nw = len(hs), len(ws)
    T = nh * nw
    origins = torch.tensor([(a, b) for a in hs for b in ws], dtype=torch.int32).to(x.device)
    hs_d, ws_d = torch.tensor(hs, dtype=torch.int32).to(x.device), torch.tensor(ws, dtype=torch.int32).to(x.device)
    pin = torch.empty((T * B, C, t, t), device=x.device, dtype=torch.float32)
    ops.LAUNCHES += 1
    L.check(lib.b200ir_tile_gather(x.data_ptr(), origins.data_ptr(), pin.data_ptr(), T, B, C, H, W, t, ops._stream()),
            "b200ir_tile_gather")
    pout = None
    for i in range(0, T * B, tile_batch):
        y = net(pin[i:i + tile_batch])
        if pout is None:
            pout = torch.empty((T * B,) + tuple(y.shape[1:]), device=x.device, dtype=torch.float32)
        ops.copy_cast(y, pout[i:i + tile_batch])
    Co = pout.shape[1]
    out = torch.empty((B, Co, H, W), device=x.device, dtype=torch.float32)
    ops.LAUNCHES += 1
    L.check(lib.b200ir_tile_blend(pout.data_ptr(), hs_d.data_ptr(), nh, ws_d.data_ptr(), nw, out.data_ptr(), B, Co, H, W, t,
                                  int(clamp), ops._stream()), "b200ir_tile_blend")
    return out
