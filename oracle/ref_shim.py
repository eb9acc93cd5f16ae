"""TEST INFRASTRUCTURE ONLY - import shim for the unmodified reference at /root/reference.

The reference imports a few packages that are absent from this image but are never touched on the hot
path (matplotlib: defectGAN/utils/util.py:8; torchmetrics: defectGAN/trainers/defectgan_trainer.py:15).
They are satisfied with empty ``sys.modules`` stubs.  /root/reference exists only in the build container:
``available()`` is False on the GPU box and every caller must skip.
"""
import os
import sys
import types
from types import SimpleNamespace

REF_ROOT = os.environ.get("DEI2I_REFERENCE_ROOT", "/root/reference")
REF_DEFECTGAN = os.path.join(REF_ROOT, "defectGAN")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_DEFECTGAN, "models", "networks"))


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install_stubs():
    try:
        import matplotlib  # noqa: F401
    except Exception:
        mpl = _stub("matplotlib")
        mpl.pyplot = _stub("matplotlib.pyplot")

    class _NA:
        def __init__(self, *a, **k):
            raise RuntimeError("metric stub: not available in the oracle")

    try:
        import torchmetrics  # noqa: F401
    except Exception:
        tm = _stub("torchmetrics")
        tm.image = _stub("torchmetrics.image", LearnedPerceptualImagePatchSimilarity=_NA)
        tm.image.lpip = _stub("torchmetrics.image.lpip", LearnedPerceptualImagePatchSimilarity=_NA)


_imported = {}


def import_reference():
    """Returns a namespace with the reference's own classes (unmodified code, imported in place)."""
    if not available():
        raise RuntimeError(f"reference not present at {REF_ROOT}")
    if _imported:
        return SimpleNamespace(**_imported)
    install_stubs()
    # The reference uses top-level package names 'models', 'utils', 'options', 'trainers'.
    # Our product package never uses those names at top level, so there is no clash.
    if REF_DEFECTGAN not in sys.path:
        sys.path.insert(0, REF_DEFECTGAN)
    from models.networks import architecture, normalization, generator, discriminator  # type: ignore
    from models import base_model  # type: ignore
    _imported.update(
        architecture=architecture,
        normalization=normalization,
        generator=generator,
        discriminator=discriminator,
        base_model=base_model,
        DefectGanGenerator=generator.DefectGanGenerator,
        DefectGanDiscriminator=discriminator.DefectGanDiscriminator,
        WGanGenerator=generator.WGanGenerator,
        WGanDiscriminator=discriminator.WGanDiscriminator,
    )
    return SimpleNamespace(**_imported)


def default_opt(image_size=128, **over):
    """Network options with the reference defaults (defectGAN/options/defectgan_options.py:25-39)
    plus the README recipe flags --add_noise --use_spectral (defectGAN/README.md:21-24)."""
    opt = SimpleNamespace(
        num_res=6, cycle_gan=False, label_nc=6, skip_conn=False, ngf=64, input_nc=3,
        use_spectral=True, num_scales=2, add_noise=True, style_norm_block_type="spade",
        hidden_nc=128, style_distill=False, embed_nc=768, ndf=64, num_layers=5,
        image_size=image_size, num_epochs=10, sean_alpha=None, use_running_stats=False,
        diff_aug="", clf_loss_type="bce", is_train=True, phase="train",
        init_type="normal", init_variance=0.02,
    )
    for k, v in over.items():
        setattr(opt, k, v)
    return opt


REF_STARGAN = os.path.join(REF_ROOT, "stargan-v2")
_sg = {}


def stargan_available() -> bool:
    return os.path.isdir(os.path.join(REF_STARGAN, "core"))


def import_stargan():
    """The reference's own stargan-v2/core/{model,solver}.py, imported in place.  Absent packages that the hot path never
    touches (munch, skimage, ffmpeg, lpips, imageio, matplotlib) are satisfied with stubs; Munch gets a working stand-in."""
    if _sg:
        return SimpleNamespace(**_sg)
    if not stargan_available():
        raise RuntimeError(f"reference not present at {REF_STARGAN}")

    class Munch(dict):
        __getattr__ = dict.get

        def __setattr__(self, k, v):
            self[k] = v

    try:
        import munch  # noqa: F401
    except Exception:
        _stub("munch", Munch=Munch)
    for name in ("skimage", "skimage.filters", "ffmpeg", "lpips", "imageio", "matplotlib", "matplotlib.pyplot"):
        try:
            __import__(name)
        except Exception:
            _stub(name, gaussian=None)
    if REF_STARGAN not in sys.path:
        sys.path.insert(0, REF_STARGAN)
    from core import model, solver  # type: ignore
    _sg.update(model=model, solver=solver)
    return SimpleNamespace(**_sg)
