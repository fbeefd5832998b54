"""Batch standardisation helpers (mirror of src/utils/separate.py:56-67)."""


def normalize_batch(batch):
    """Per-item mean / unbiased std over (channel, time) of the mixture, std clamped at
    1e-5; the target (if any) is transformed with the mixture's statistics.
    Returns ((mix, tgt), mean, std)."""
    mix, tgt = batch
    mean = mix.mean(dim=(1, 2), keepdim=True)
    std = mix.std(dim=(1, 2), keepdim=True).clamp(min=1e-5)
    mix = (mix - mean) / std
    if tgt is not None:
        tgt = (tgt - mean) / std
    return (mix, tgt), mean, std


def denormalize_batch(x, mean, std):
    return x * std + mean
