"""Mask helpers of the MAE pre-training path (defectGAN/utils/util.py:48-71).

``generate_mask`` / ``generate_shifted_mask`` keep the reference's signatures, RNG calls (host-side ``torch.randint`` for the two
shifts, then ``torch.bernoulli`` on the coarse grid - in that order) and return value (the full-resolution [B,1,H,W] mask).
``draw_shifted_mask`` makes the same draws but returns the COARSE grid and the shifts: what the fused mask-token kernel
(dei2i_mask_token_fwd) consumes - the nearest up-sampling and the crop are evaluated on the fly on the device."""
import torch
import torch.nn.functional as F


def generate_mask(image_size, patch_size, mask_ratio):
    """utils/util.py:48-57: Bernoulli(1 - mask_ratio) per patch, nearest-upsampled to [b, 1, h, w]."""
    gh, gw = image_size[2] // patch_size, image_size[3] // patch_size
    mask = torch.bernoulli(torch.full((image_size[0], 1, gh, gw), (1 - mask_ratio)))
    return F.interpolate(mask, scale_factor=patch_size, mode="nearest")


def draw_shifted_mask(image_size, patch_size, mask_ratio):
    """The draws of generate_shifted_mask (utils/util.py:60-71) -> (coarse [b, gh, gw] fp32 host tensor, h_shift, w_shift)."""
    img_h, img_w = image_size[2], image_size[3]
    h_shift = int(torch.randint(low=0, high=patch_size, size=(1,)))
    w_shift = int(torch.randint(low=0, high=patch_size, size=(1,)))
    gh, gw = (img_h + patch_size) // patch_size, (img_w + patch_size) // patch_size
    coarse = torch.bernoulli(torch.full((image_size[0], 1, gh, gw), (1 - mask_ratio)))
    return coarse[:, 0].contiguous(), h_shift, w_shift


def expand_mask(coarse, h_shift, w_shift, image_size, patch_size):
    """coarse grid + shifts -> the [b, 1, h, w] mask generate_shifted_mask returns."""
    full = F.interpolate(coarse.unsqueeze(1), scale_factor=patch_size, mode="nearest")
    return full[:, :, h_shift: h_shift + image_size[2], w_shift: w_shift + image_size[3]]


def generate_shifted_mask(image_size, patch_size, mask_ratio):
    """utils/util.py:60-71."""
    coarse, hs, ws = draw_shifted_mask(image_size, patch_size, mask_ratio)
    return expand_mask(coarse, hs, ws, image_size, patch_size)
