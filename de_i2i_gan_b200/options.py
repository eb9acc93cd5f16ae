"""Programmatic stand-in for ``TrainOptions().parse()`` (options/defectgan_options.py:77-111 over
options/base_options.py:12-52 and options/train_options.py:9-43): a Namespace carrying the same option names and
defaults, so the mirrored trainer / model / networks read exactly the attributes the reference's do.  The reference's
argparse classes remain the CLI surface (they are host-only code and work unchanged on top of this package); this
helper exists for the benchmark, the smoke test and the parity tests, which have no command line.

Deliberate differences from the parsers' defaults (tests/test_options_cpu.py pins everything else against the reference's own
parsers): ``name`` is 'exp0' (what gather_options turns the default 'exp' into, base_options.py:67-76), ``log_dir`` is None (no
TensorBoard writer unless asked for; the parsers say ./log), and the values the entry scripts add after parsing are present
(``is_train``, ``clf_loss_type``, ``iters_per_epoch``, ``device``)."""
from pathlib import Path
from types import SimpleNamespace

import torch


def train_options(**over) -> SimpleNamespace:
    opt = SimpleNamespace(
        # base_options.py:14-50 with the DefectGAN overrides of defectgan_options.py:14-45
        name="exp0", model="defectgan", gpu_ids="0", ckpt_dir=Path("./ckpt"), phase="train",
        batch_size=4, image_size=128, input_nc=3, output_nc=3, label_nc=6,
        data_dir=Path("./data"), dataset_name="codebrim", dataset_data_type=None, load_from_opt_file=None,
        init_type="normal", init_variance=0.02, use_spectral=False, load_model_name=None, which_epoch="latest",
        ngf=64, num_scales=2, num_res=6, add_noise=False, style_norm_block_type="spade", hidden_nc=128,
        ndf=64, num_layers=5, cycle_gan=False, skip_conn=False,
        embed_nc=768, latent_dim=16, embed_path=None, num_embeds=5, sean_alpha=None, style_distill=False,
        use_running_stats=False,
        # metric options (defectgan_options.py:51-62): carried for namespace parity; FID / IS / LPIPS are not built here
        dims=2048, num_imgs=5_000, npz_path=None, npy_path=None, num_lpips_images=10,
        # train_options.py:9-43 with the overrides of defectgan_options.py:87-109
        num_display_images=8, save_img_freq=4, save_ckpt_freq=10, save_latest_freq=1000,
        optimizer="adam", lr=[2e-4], num_epochs=-1, num_iters=500_000, scheduler="step", lr_decay=5e-3,
        loss_weight=[2, 5, 5, 5, 1], num_critics=5, diff_aug="", continue_training=False, log_dir=None,
        is_train=True,
        # set by the entry script from the dataset (train_defectgan.py:54, codebrim_dataset.py:12) / loader length
        clf_loss_type="bce", iters_per_epoch=1000,
    )
    for k, v in over.items():
        if not hasattr(opt, k):
            raise AttributeError(f"unknown option {k!r}")
        setattr(opt, k, v)
    if not hasattr(opt, "device"):
        # utils/util.py:11-18: cuda:<first gpu id>; under torchrun the local rank picks the device
        import os
        opt.device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", opt.gpu_ids.split(",")[0])))
    return opt


def wgan_options(**over) -> SimpleNamespace:
    """options/wgan_options.py:12-57 defaults (model 'wgan', 64x64, batch 128, noise_dim 100, RMSprop lr 5e-5, clipping 0.03)."""
    opt = SimpleNamespace(
        name="exp0", model="wgan", gpu_ids="0", ckpt_dir=Path("./ckpt"), phase="train", dataset_name="face",
        batch_size=128, image_size=64, input_nc=3, output_nc=3, ngf=64, ndf=64, noise_dim=100,
        data_dir=Path("./data"), dataset_data_type=None, load_from_opt_file=None, use_spectral=False,
        init_type="xavier", init_variance=0.02, load_model_name=None, which_epoch="latest",  # base_options.py:37: WGAN keeps 'xavier'
        num_display_images=64, save_img_freq=1, save_ckpt_freq=10, save_latest_freq=1000,
        optimizer="rmsprop", lr=5e-5, num_epochs=120, num_iters=-1, scheduler="step", lr_decay=1.0,
        num_critics=5, clipping_limit=0.03, continue_training=False, log_dir=None, is_train=True, iters_per_epoch=1000,
    )
    for k, v in over.items():
        if not hasattr(opt, k):
            raise AttributeError(f"unknown option {k!r}")
        setattr(opt, k, v)
    if not hasattr(opt, "device"):
        import os
        opt.device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", opt.gpu_ids.split(",")[0])))
    return opt


def pretrain_options(**over) -> SimpleNamespace:
    """PreTrainOptions (options/defectgan_options.py:143-189): the MAE pre-training namespace - batch 32, AdamW lr 1.5e-4, cosine
    schedule (lr_decay 0.05), loss weights [rec, cls_d, cls_g] = [10, 3, 1], mask ratio 0.75, patch 8, 'position' mask token."""
    kw = dict(batch_size=32, save_latest_freq=300, num_display_images=4, save_img_freq=1, optimizer="adamw", num_epochs=200,
              lr=[1.5e-4], scheduler="cos", lr_decay=0.05, loss_weight=[10, 3, 1], num_critics=1, num_iters=-1)
    extra = dict(mask_ratio=0.75, split_training=False, mask_token_type="position", patch_size=8)
    for k in list(over):
        if k in extra:
            extra[k] = over.pop(k)
    kw.update(over)
    opt = train_options(**kw)
    for k, v in extra.items():
        setattr(opt, k, v)
    return opt


def readme_recipe(**over) -> SimpleNamespace:
    """defectGAN/README.md:21-24: the documented DefectGAN training recipe adds --add_noise --use_spectral."""
    kw = dict(add_noise=True, use_spectral=True)
    kw.update(over)
    return train_options(**kw)
