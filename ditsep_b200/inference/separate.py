"""Separation entry point (mirror of src/inference/separate.py:73-99)."""
from __future__ import annotations

import torch


def scale_output(mix, sep):
    """Per-source least-squares rescale alpha = <mix, sep> / <sep, sep> (separate.py:73-78)."""
    num = (mix * sep).sum(dim=-1, keepdim=True)
    denom = (sep * sep + 1e-10).sum(dim=-1, keepdim=True)
    return num / denom * sep


def separate(mix, model, sampler_kwargs, device):
    """mix: (1, T) waveform -> (1, n_src, T) on the CPU (separate.py:81-99)."""
    mix = mix.to(device)
    mix = mix[None]
    (mix_norm, _), *__ = model.normalize_batch((mix, None))
    sampler = model.get_pc_sampler("reverse_diffusion", "ald2", mix_norm, **sampler_kwargs)
    with torch.no_grad():
        sep, nfe, *_ = sampler()
    sep = scale_output(mix, sep)
    return sep.cpu()
