"""Host-side helpers mirroring defectGAN/utils (only what the hot path touches: DiffAugment, the MAE mask helpers)."""
