"""Host-side mirror of the reference's ``trainers`` package for the hot path (trainers/__init__.py:4-25 naming
convention: trainers/<name>_trainer.py -> <Name>Trainer)."""
import importlib


def find_trainer_using_model_name(model_name):
    module = importlib.import_module(f"{__name__}.{model_name}_trainer")
    target = model_name.replace("_", "") + "trainer"
    for name, cls in module.__dict__.items():
        if name.lower() == target and isinstance(cls, type):
            return cls
    raise ValueError(f"In {module.__name__}.py, there should be a trainer class matching {target} in lowercase")
