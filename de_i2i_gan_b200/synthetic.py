"""Seeded synthetic inputs of the hot path - what bench.py, smoke() and the tests feed the trainers.

``synthetic_batch`` is the recipe the survey fixed for every measurement: background and defect images
``randn(B,3,H,H).clamp(-1,1)`` and multi-hot defect labels with column 0 (the 'normal' class, defectgan_model.py:413-415)
forced to 0, all drawn from one ``torch.Generator().manual_seed(seed)`` on the host.
"""
import torch


def synthetic_batch(batch: int, image_size: int, label_nc: int = 6, seed: int = 0):
    """-> (bg [B,3,H,H], labels [B,label_nc], df [B,3,H,H]) FP32 host tensors."""
    gen = torch.Generator().manual_seed(seed)
    bg = torch.randn(batch, 3, image_size, image_size, generator=gen).clamp_(-1, 1)
    df = torch.randn(batch, 3, image_size, image_size, generator=gen).clamp_(-1, 1)
    labels = (torch.rand(batch, label_nc, generator=gen) > 0.5).float()
    labels[:, 0] = 0
    return bg, labels, df


def spatial_labels(batch: int, image_size: int, label_nc: int = 6, seed: int = 0, cells: int = 8):
    """A spatial label map [B,label_nc,H,H] as test_defectgan.py:219-230 builds for mask-guided synthesis: a coarse
    ``cells`` x ``cells`` grid of one-hot classes, nearest-upsampled to the image size."""
    gen = torch.Generator().manual_seed(seed)
    idx = torch.randint(0, label_nc, (batch, cells, cells), generator=gen)
    seg = torch.nn.functional.one_hot(idx, label_nc).permute(0, 3, 1, 2).float()
    return torch.nn.functional.interpolate(seg, size=(image_size, image_size), mode="nearest")
