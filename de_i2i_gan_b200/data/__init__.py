"""Input pipeline of the training step on the device (train_defectgan.py:57-79)."""
from .pipeline import DevicePrefetcher, GpuTrainAugment, ShardedRandomSampler, draw_train_params  # noqa: F401
