"""Deterministic synthetic 2-/3-speaker mixtures for benchmarks and parity tests.

There is no network for WSJ0-2mix / LibriMix; SURVEY.md section 8(d) defines the
synthetic stand-in: each source is Gaussian noise coloured by a random 2-pole
resonance (speech-band), multiplied by a slow raised-cosine amplitude envelope
(2-6 Hz), unit variance, with a random relative gain in [-5, 5] dB.  Everything is
drawn from one seeded CPU generator so the same mixtures are regenerated on any box.
"""
from __future__ import annotations

import math

import torch


def synth_sources(batch: int, n_src: int, n_samples: int, fs: int = 8000, seed: int = 1234) -> torch.Tensor:
    """Returns sources (batch, n_src, n_samples) fp32 on CPU."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n_samples, dtype=torch.float64) / fs
    out = torch.empty(batch, n_src, n_samples, dtype=torch.float64)
    for b in range(batch):
        for s in range(n_src):
            white = torch.randn(n_samples + 64, generator=g, dtype=torch.float64)
            f0 = 200.0 + 1800.0 * float(torch.rand((), generator=g))
            r = 0.90 + 0.08 * float(torch.rand((), generator=g))
            a1 = -2.0 * r * math.cos(2 * math.pi * f0 / fs)
            a2 = r * r
            # 2-pole resonator as a truncated impulse response (FIR, length 256) -- exact
            # enough for colouring and trivially reproducible.
            h = torch.zeros(256, dtype=torch.float64)
            h[0] = 1.0
            h[1] = -a1
            for k in range(2, 256):
                h[k] = -a1 * h[k - 1] - a2 * h[k - 2]
            col = torch.nn.functional.conv1d(
                torch.nn.functional.pad(white, (255, 0))[None, None], h.flip(0)[None, None])[0, 0][64:]
            fe = 2.0 + 4.0 * float(torch.rand((), generator=g))
            ph = 2 * math.pi * float(torch.rand((), generator=g))
            env = 0.5 - 0.5 * torch.cos(2 * math.pi * fe * t + ph)
            sig = col * (0.1 + env)
            sig = sig / sig.std()
            gain_db = -5.0 + 10.0 * float(torch.rand((), generator=g))
            out[b, s] = sig * 10 ** (gain_db / 20)
    return out.float()


def synth_mixtures(batch: int, n_src: int = 2, n_samples: int = 32000, fs: int = 8000, seed: int = 1234):
    """Returns (mix (batch,1,T), sources (batch,n_src,T)), not normalised."""
    src = synth_sources(batch, n_src, n_samples, fs, seed)
    return src.sum(dim=1, keepdim=True), src
