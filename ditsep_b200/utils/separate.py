"""Batch standardisation and source-order helpers (mirror of src/utils/separate.py)."""
import torch


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


def shuffle_sources(x):
    """A different random source order per batch entry (src/utils/separate.py:3-21)."""
    if x.ndim <= 1:
        return x
    order = torch.argsort(torch.rand(x.shape[:2], device=x.device), dim=1)
    return torch.gather(x, 1, order.reshape(order.shape + (1,) * (x.ndim - 2)).expand_as(x))


def power_order_sources(x):
    """Sources ordered by increasing variance over the last axis (src/utils/separate.py:35-53)."""
    if x.ndim <= 1:
        return x
    order = torch.argsort(torch.var(x, dim=-1), dim=1)
    return torch.gather(x, 1, order.reshape(order.shape + (1,) * (x.ndim - 2)).expand_as(x))


def select_elem_at_random(x, dim=-1, batch_dim=0, select=None):
    """One element along `dim`, drawn independently per batch entry, the axis kept with size 1
    (src/utils/separate.py:24-32).  `select`: the (batch,) indices to use instead of a random draw (parity runs)."""
    if select is None:
        select = torch.randint(x.shape[dim], size=(x.shape[batch_dim],), device=x.device)
    x = x.movedim(dim, -1)
    idx = select.to(x.device).reshape((-1,) + (1,) * (x.ndim - 1)).expand(x.shape[:-1] + (1,))
    return torch.gather(x, -1, idx).movedim(-1, dim)
