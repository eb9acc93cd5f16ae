"""Device-side training input pipeline (data/pipeline.py, dei2i_augment_train) vs torchvision's tensor transforms with the same
parameters: Resize(1.5 S) + resized crop (antialiased bilinear) + flips + ColorJitter ops in the drawn order + Normalize."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _smooth_image(h, w, seed):
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(1, 3, h // 8 + 2, w // 8 + 2, generator=g)
    img = torch.nn.functional.interpolate(low, size=(h, w), mode="bicubic", align_corners=False).clamp(0, 1)[0]
    return (img * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()  # HWC uint8, natural-image-like spectrum


def _reference(img_u8, p, S):
    from torchvision.transforms.v2 import functional as TF
    x = img_u8.permute(2, 0, 1).float() / 255
    H, W = x.shape[1:]
    y0, x0, h, w = [float(v) for v in p[:4]]
    # one antialiased bilinear resample of the source box == Resize followed by the resized crop (both are triangle filters)
    out = TF.resized_crop(x, int(round(y0)), int(round(x0)), int(round(h)), int(round(w)), [S, S], antialias=True)
    if p[4] > 0.5:
        out = out.flip(-1)
    if p[5] > 0.5:
        out = out.flip(-2)
    for slot in range(3):
        op = int(p[9 + slot])
        f = float(p[6 + op])
        out = (TF.adjust_brightness, TF.adjust_contrast, TF.adjust_saturation)[op](out, f)
    return (out - 0.5) / 0.5


@pytest.mark.parametrize("S", [64, 128])
def test_augment_train_matches_torchvision(S):
    from de_i2i_gan_b200.data import GpuTrainAugment, draw_train_params
    imgs = [_smooth_image(300, 400, 1), _smooth_image(512, 384, 2), _smooth_image(256, 256, 3), _smooth_image(200, 640, 4)]
    g = torch.Generator().manual_seed(7)
    params = draw_train_params([(im.shape[0], im.shape[1]) for im in imgs], S, g)
    # integer source boxes so that torchvision's integer crop API describes the same region
    params[:, :4] = params[:, :4].round()
    aug = GpuTrainAugment(S, "cuda:0")
    out = aug(imgs, params=params).cpu()
    assert out.shape == (4, 3, S, S) and out.min() >= -1.0001 and out.max() <= 1.0001
    for n, im in enumerate(imgs):
        ref = _reference(im, params[n], S)
        err = (out[n] - ref).abs()
        assert err.mean().item() < 4e-3 and err.max().item() < 4e-2, (n, err.mean().item(), err.max().item())


def test_sharded_sampler_and_prefetcher():
    from de_i2i_gan_b200.data import DevicePrefetcher, GpuTrainAugment, ShardedRandomSampler
    a = list(ShardedRandomSampler(10, 25, rank=0, world=2, seed=3))
    b = list(ShardedRandomSampler(10, 25, rank=1, world=2, seed=3))
    full = list(ShardedRandomSampler(10, 25, rank=0, world=1, seed=3))
    assert len(full) == 25 and sorted(a + b) == sorted(full) and a == full[0::2] and b == full[1::2]
    assert sorted(full[:10]) == list(range(10))  # without replacement inside one permutation
    imgs = [_smooth_image(96, 128, s) for s in range(6)]
    batches = [([imgs[0], imgs[1]], torch.eye(6)[:2]), ([imgs[2], imgs[3]], torch.eye(6)[2:4]), ([imgs[4], imgs[5]], torch.eye(6)[4:])]
    aug = GpuTrainAugment(64, "cuda:0")
    g = torch.Generator().manual_seed(1)
    got = [(x.clone(), y.clone()) for x, y in DevicePrefetcher(batches, aug, generator=g)]
    assert len(got) == 3 and all(x.is_cuda and x.shape == (2, 3, 64, 64) for x, _ in got)
    g2 = torch.Generator().manual_seed(1)
    for (x, y), (ims, lab) in zip(got, batches):  # same generator stream, synchronous path -> same batches
        ref = aug(ims, generator=g2)
        torch.cuda.synchronize()
        assert torch.equal(ref, x) and torch.equal(y.cpu(), lab)
