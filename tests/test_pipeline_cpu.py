"""Host logic of the device input pipeline (de_i2i_gan_b200/data/pipeline.py; reference: train_defectgan.py:57-79) - no GPU:
the sharded sampler (ranks draw disjoint shards of the single-process permutation stream) and the random-parameter
draw of the augmentation kernel (torchvision's RandomResizedCrop / flips / ColorJitter distributions)."""
import math

import torch

from de_i2i_gan_b200.data import ShardedRandomSampler, draw_train_params


def test_sharded_sampler_partitions_the_single_process_stream():
    for n, k, world in [(10, 25, 2), (7, 7, 4), (100, 64, 8), (5, 3, 2)]:
        full = list(ShardedRandomSampler(n, k, rank=0, world=1, seed=3))
        assert len(full) == k and all(0 <= i < n for i in full)
        for lo in range(0, k, n):  # without replacement inside each pass over the dataset (RandomSampler(replacement=False))
            chunk = full[lo:lo + n]
            assert len(set(chunk)) == len(chunk)
        shards = [list(ShardedRandomSampler(n, k, rank=r, world=world, seed=3)) for r in range(world)]
        for r, s in enumerate(shards):
            assert s == full[r::world]
            assert len(ShardedRandomSampler(n, k, rank=r, world=world, seed=3)) == len(s)
        assert sorted(sum(shards, [])) == sorted(full)


def test_sharded_sampler_epochs_and_seeds():
    s = ShardedRandomSampler(50, 50, rank=1, world=2, seed=11)
    a = list(s)
    assert list(s) == a  # same epoch: same order (every rank must be able to rebuild the stream)
    s.set_epoch(1)
    b = list(s)
    assert b != a and len(b) == len(a)
    assert list(ShardedRandomSampler(50, 50, rank=1, world=2, seed=12)) != a
    assert list(ShardedRandomSampler(50, rank=0, world=1)) and len(ShardedRandomSampler(50)) == 50  # num_samples defaults to the dataset


def test_train_params_follow_the_reference_distributions():
    size = 64
    src = [(200, 300), (300, 200), (128, 128), (97, 411), (500, 90)] * 40
    g = torch.Generator().manual_seed(5)
    p = draw_train_params(src, size, g)
    assert p.shape == (len(src), 12) and p.dtype == torch.float32
    again = draw_train_params(src, size, torch.Generator().manual_seed(5))
    assert torch.equal(p, again)  # a function of the generator alone
    r = int(size * 1.5)
    for (H, W), row in zip(src, p.tolist()):
        y0, x0, h, w = row[:4]
        assert y0 >= 0 and x0 >= 0 and h > 0 and w > 0
        assert y0 + h <= H + 1e-3 and x0 + w <= W + 1e-3  # the crop box lies inside the SOURCE image
        Hr, Wr = (r, int(r * W / H)) if H <= W else (int(r * H / W), r)
        hr, wr = h * Hr / H, w * Wr / W  # the box in the resized image, where RandomResizedCrop draws it
        frac = hr * wr / (Hr * Wr)
        ratio, in_ratio = wr / hr, Wr / Hr
        assert 0.70 < ratio < 1.43, ratio  # ratio in (3/4, 4/3) up to the integer rounding of w and h, fallback included
        assert frac <= 1.0 + 1e-6
        if 0.6 < in_ratio < 1.7:  # (very elongated sources often end in the centre-crop fallback, which ignores the scale range)
            assert frac > 0.55, frac  # scale = (0.6, 1.0)
        assert row[4] in (0.0, 1.0) and row[5] in (0.0, 1.0)
        assert all(0.8 - 1e-6 <= f <= 1.2 + 1e-6 for f in row[6:9])
        assert sorted(row[9:12]) == [0.0, 1.0, 2.0]
    flips = p[:, 4:6].mean(0)
    assert 0.35 < flips[0] < 0.65 and 0.35 < flips[1] < 0.65  # p = 0.5 each
    orders = {tuple(rw) for rw in p[:, 9:12].tolist()}
    assert len(orders) == 6  # every order of the three colour ops occurs


def test_train_params_switches():
    src = [(120, 160)] * 8
    p = draw_train_params(src, 32, torch.Generator().manual_seed(1), jitter=(0.0, 0.0, 0.0), flips=False)
    assert torch.all(p[:, 4:6] == 0) and torch.all(p[:, 6:9] == 1.0)
    # an extreme aspect ratio cannot hold a (3/4 .. 4/3) box of 60 % of the area: the centre-crop fallback at the closest ratio
    p = draw_train_params([(40, 1000)], 32, torch.Generator().manual_seed(2))
    y0, x0, h, w = p[0, :4].tolist()
    r = 48
    Hr, Wr = r, int(r * 1000 / 40)
    hr, wr = h * Hr / 40, w * Wr / 1000
    assert abs(hr - Hr) < 1e-3 and abs(wr - round(Hr * 4 / 3)) < 1e-3
    assert math.isclose(x0 * Wr / 1000, (Wr - round(Hr * 4 / 3)) // 2, abs_tol=1e-3) and y0 == 0
