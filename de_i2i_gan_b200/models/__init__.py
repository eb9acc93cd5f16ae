"""Host-side mirror of the reference's ``models`` package for the hot path (same class names, constructor and
forward signatures, module tree and state-dict keys as /root/reference/defectGAN/models)."""
import importlib


def find_model_using_name(model_name):
    """models/__init__.py:6-29 naming convention: models/<name>_model.py -> <Name>Model."""
    module = importlib.import_module(f"{__name__}.{model_name}_model")
    target = model_name.replace("_", "") + "model"
    for name, cls in module.__dict__.items():
        if name.lower() == target:
            return cls
    raise ValueError(f"no model class for {model_name!r}")


def create_model(opt):
    """models/__init__.py:32-37."""
    return find_model_using_name(opt.model)(opt)
