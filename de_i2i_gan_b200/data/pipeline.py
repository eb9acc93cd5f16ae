"""The reference's training input pipeline (train_defectgan.py:57-79) re-cut for a step that takes 80 ms for 32 images per GPU:

    reference                                           here
    4 DataLoader workers: PIL decode + Resize +         workers only DECODE (uint8 HWC); everything after the decode is ONE device
      RandomResizedCrop + flips + ColorJitter +           kernel sequence on the uint8 batch (dei2i_augment_train): 4x fewer bytes
      ToTensor + Normalize  (FP32 CHW, per image)          over PCIe than FP32 tensors, no per-image Python / PIL arithmetic
    RandomSampler(replacement=False, num_samples)        ShardedRandomSampler: the same permutation stream, strided over the ranks
    pin_memory + .to(device) inside the step             DevicePrefetcher: batch k+1 is copied and augmented on a side stream while
                                                           step k runs; the trainer receives device tensors

Random parameters follow torchvision's distributions (RandomResizedCrop.get_params with scale (0.6, 1), ratio (3/4, 4/3), 10 tries
and the centre-crop fallback; flips with p = 0.5; ColorJitter factors uniform in [0.8, 1.2] applied in a random order) and are
drawn on the host with a ``torch.Generator``.  The arithmetic is compared with torchvision's tensor transforms in
tests/test_gpu_pipeline.py (antialiased bilinear resampling: within 2/255 on natural-image-like inputs; the PIL pipeline itself
rounds to uint8 between stages, which the fused kernel does not).
"""
import math
from typing import Iterable, List, Sequence, Tuple

import torch

from .. import _lib
from ..ops import _p, _stream, check


def draw_train_params(src_hw: Sequence[Tuple[int, int]], size: int, generator=None, jitter=(0.2, 0.2, 0.2), flips=True) -> torch.Tensor:
    """fp32 [N, 12] parameter block of dei2i_augment_train for N source images of sizes ``src_hw``:
    Resize(int(1.5 * size)) (smaller edge) -> RandomResizedCrop((size, size), scale=(0.6, 1.0)) expressed as one crop box in SOURCE
    pixels, horizontal / vertical flip flags, ColorJitter factors (brightness, contrast, saturation) and their order."""
    g = generator
    rnd = lambda lo, hi: float(torch.empty(1).uniform_(lo, hi, generator=g))  # noqa: E731
    out = torch.zeros(len(src_hw), 12)
    for n, (H, W) in enumerate(src_hw):
        r = int(size * 1.5)
        Hr, Wr = (r, int(r * W / H)) if H <= W else (int(r * H / W), r)  # transforms.Resize(int): smaller edge -> r
        area = Hr * Wr
        box = None
        for _ in range(10):  # RandomResizedCrop.get_params
            target = area * rnd(0.6, 1.0)
            ratio = math.exp(rnd(math.log(3 / 4), math.log(4 / 3)))
            w, h = int(round(math.sqrt(target * ratio))), int(round(math.sqrt(target / ratio)))
            if 0 < w <= Wr and 0 < h <= Hr:
                i = int(torch.randint(0, Hr - h + 1, (1,), generator=g))
                j = int(torch.randint(0, Wr - w + 1, (1,), generator=g))
                box = (i, j, h, w)
                break
        if box is None:  # fallback: centre crop at the closest allowed ratio
            in_ratio = Wr / Hr
            if in_ratio < 3 / 4:
                w, h = Wr, int(round(Wr / (3 / 4)))
            elif in_ratio > 4 / 3:
                h, w = Hr, int(round(Hr * (4 / 3)))
            else:
                w, h = Wr, Hr
            box = ((Hr - h) // 2, (Wr - w) // 2, h, w)
        sy, sx = H / Hr, W / Wr  # resized -> source coordinates
        out[n, 0], out[n, 1], out[n, 2], out[n, 3] = box[0] * sy, box[1] * sx, box[2] * sy, box[3] * sx
        if flips:
            out[n, 4] = float(torch.rand(1, generator=g) < 0.5)
            out[n, 5] = float(torch.rand(1, generator=g) < 0.5)
        order = torch.randperm(3, generator=g)
        for k, amp in enumerate(jitter):  # ops: 0 brightness, 1 contrast, 2 saturation
            out[n, 6 + k] = rnd(max(0.0, 1 - amp), 1 + amp) if amp else 1.0
        out[n, 9:12] = order.float()
    return out


class GpuTrainAugment:
    """uint8 HWC images (host, any sizes) -> FP32 NCHW [N, 3, size, size] in [-1, 1] on the device, one fused kernel sequence."""

    def __init__(self, size: int, device, jitter=True):
        self.size, self.device, self.jitter = int(size), torch.device(device), bool(jitter)

    def stage(self, images: List[torch.Tensor]):
        """Concatenate the decoded images into one pinned uint8 buffer (+ offsets / sizes) - the H2D payload of a batch."""
        sizes = [(int(im.shape[0]), int(im.shape[1])) for im in images]
        offs, total = [], 0
        for h, w in sizes:
            offs.append(total)
            total += (h * w * 3 + 15) // 16 * 16
        buf = torch.empty(total, dtype=torch.uint8).pin_memory()
        for im, o, (h, w) in zip(images, offs, sizes):
            assert im.dtype == torch.uint8 and im.dim() == 3 and im.shape[2] == 3, "decoded images are uint8 [H, W, 3]"
            buf[o:o + h * w * 3].copy_(im.reshape(-1))
        return buf, torch.tensor(offs, dtype=torch.int64).pin_memory(), torch.tensor(sizes, dtype=torch.int32).pin_memory(), sizes

    def __call__(self, images: List[torch.Tensor], params: torch.Tensor = None, generator=None):
        buf, offs, hw, sizes = self.stage(images)
        if params is None:
            params = draw_train_params(sizes, self.size, generator)
        n, dev = len(images), self.device
        d_buf, d_offs, d_hw = buf.to(dev, non_blocking=True), offs.to(dev, non_blocking=True), hw.to(dev, non_blocking=True)
        d_par = params.float().pin_memory().to(dev, non_blocking=True)
        lib = _lib.load()
        out = torch.empty(n, 3, self.size, self.size, dtype=torch.float32, device=dev)
        scratch = torch.empty(lib.dei2i_augment_scratch(n, self.size), dtype=torch.float32, device=dev)
        check(lib.dei2i_augment_train(_p(d_buf), _p(d_offs), _p(d_hw), _p(d_par), _p(out), _p(scratch), n, self.size, int(self.jitter),
                                      _stream()), "dei2i_augment_train")
        return out


class ShardedRandomSampler(torch.utils.data.Sampler):
    """RandomSampler(replacement=False, num_samples=k) (train_defectgan.py:70-73) for data parallelism: every rank walks the same
    seeded permutation stream and takes every ``world``-th index, so the ranks' batches are disjoint and their union is what the
    single-process sampler would have drawn."""

    def __init__(self, data_len: int, num_samples: int = None, rank: int = 0, world: int = 1, seed: int = 0):
        self.n, self.k, self.rank, self.world, self.seed, self.epoch = data_len, num_samples or data_len, rank, world, seed, 0

    def set_epoch(self, epoch: int):
        self.epoch = epoch

    def __iter__(self):
        g = torch.Generator().manual_seed(self.seed + self.epoch)
        idx, need = [], self.k
        while need > 0:  # RandomSampler without replacement re-permutes when num_samples exceeds the dataset
            p = torch.randperm(self.n, generator=g).tolist()
            idx += p[:need]
            need -= len(p[:need])
        return iter(idx[self.rank::self.world])

    def __len__(self):
        return (self.k - self.rank + self.world - 1) // self.world


class DevicePrefetcher:
    """Wraps an iterable of host batches ``(list of uint8 HWC images, labels)``: while the trainer consumes batch k, batch k+1 is
    staged, copied and augmented on a side stream.  Yields ``(images fp32 [N,3,S,S] on the device, labels on the device)``."""

    def __init__(self, loader: Iterable, augment: GpuTrainAugment, generator=None):
        self.loader, self.augment, self.generator = loader, augment, generator
        self.stream = torch.cuda.Stream(device=augment.device)

    def _launch(self, batch):
        images, labels = batch[0], batch[1]
        with torch.cuda.stream(self.stream):
            x = self.augment(images, generator=self.generator)
            y = labels.pin_memory().to(self.augment.device, non_blocking=True) if isinstance(labels, torch.Tensor) else labels
        return x, y

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = self._launch(next(it))
        except StopIteration:
            return
        for batch in it:
            cur, nxt = nxt, self._launch(batch)
            torch.cuda.current_stream().wait_stream(self.stream)  # (conservative: also waits for the batch just launched)
            for t in cur:
                if isinstance(t, torch.Tensor):
                    t.record_stream(torch.cuda.current_stream())
            yield cur
        torch.cuda.current_stream().wait_stream(self.stream)
        yield nxt
