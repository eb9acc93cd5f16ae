"""de_i2i_gan_b200/options.py against the reference's own argparse classes (options/base_options.py, train_options.py,
defectgan_options.py, wgan_options.py), imported live through oracle/ref_shim.py: every option the reference's parser defines exists
with the same default, apart from the documented differences.  Skipped where /root/reference is absent (the GPU box)."""
import argparse
import importlib

import pytest

from oracle import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference sources not present")

DOCUMENTED = {"name", "log_dir"}  # 'exp' -> 'exp0' by gather_options; no TensorBoard writer unless asked for
ADDED_AFTER_PARSING = {"is_train", "clf_loss_type", "iters_per_epoch", "device"}  # train_defectgan.py:54, base_options.parse


def _reference_defaults(module, cls, argv):
    ref_shim.import_reference()
    o = getattr(importlib.import_module(module), cls)()
    parser = o.initialize(argparse.ArgumentParser(conflict_handler="resolve"))  # what gather_options builds, without its side effects
    return vars(parser.parse_args(argv))


@pytest.mark.parametrize("module,cls,ours,argv,extra", [
    ("options.defectgan_options", "TrainOptions", "train_options", ["--model", "defectgan", "--dataset_name", "codebrim"], set()),
    ("options.defectgan_options", "PreTrainOptions", "pretrain_options", ["--model", "defectgan", "--dataset_name", "codebrim"], set()),
    ("options.wgan_options", "TrainOptions", "wgan_options", ["--model", "wgan", "--dataset_name", "face"], set()),
])
def test_defaults_match_the_reference_parsers(module, cls, ours, argv, extra):
    from de_i2i_gan_b200 import options
    ref = _reference_defaults(module, cls, argv)
    mine = vars(getattr(options, ours)())
    missing = [k for k in ref if k not in mine]
    assert not missing, missing
    differs = {k: (ref[k], mine[k]) for k in ref if k not in DOCUMENTED and mine[k] != ref[k]}
    assert not differs, differs
    unknown = set(mine) - set(ref) - ADDED_AFTER_PARSING - extra
    assert not unknown, unknown
    assert mine["name"] == "exp0" and mine["log_dir"] is None


def test_unknown_option_is_refused():
    from de_i2i_gan_b200 import options
    with pytest.raises(AttributeError):
        options.train_options(no_such_option=1)
    with pytest.raises(AttributeError):
        options.wgan_options(no_such_option=1)
    assert options.readme_recipe().add_noise and options.readme_recipe().use_spectral  # defectGAN/README.md:21-24
