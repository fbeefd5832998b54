"""NCSN++ score network (mirror of src/models/diffsep/ncsnpp.py and ncsnpp_utils/).

``NCSNpp`` keeps the reference constructor (ncsnpp.py:48-74), module tree and
``state_dict`` key layout (``all_modules.{i}.{GroupNorm_0,Conv_0,Dense_0,GroupNorm_1,Conv_1,
Conv_2}.{weight,bias}``, ``all_modules.{i}.NIN_k.{W,b}``, ``all_modules.0.W``,
``output_layer.*``) so reference checkpoints load with ``strict=True``.  The sub-modules
are parameter containers only: ``forward`` (ncsnpp.py:320-478) is executed by
``engine.NcsnppEngine`` as a fixed sequence of sm_100a kernel launches on NHWC
activations.  There is no PyTorch-operator or CPU fallback.

Supported configuration family = the one every DiffSep config uses: biggan res-blocks,
FIR resampling, ``progressive="output_skip"``, ``progressive_input="input_skip"``,
``progressive_combine="sum"``, fourier embedding, ``skip_rescale=True``, swish.
"""
from __future__ import annotations

import math
import os
from typing import List

import numpy as np
import torch
import torch.nn as nn

from .engine import NcsnppEngine, layout_of


def _variance_scaling_(t: torch.Tensor, scale: float, in_axis=1, out_axis=0):
    """default_init = variance_scaling(scale, "fan_avg", "uniform") (layers.py:48-102)."""
    scale = 1e-10 if scale == 0 else scale
    shape = t.shape
    rf = float(np.prod(shape)) / shape[in_axis] / shape[out_axis]
    fan_in, fan_out = shape[in_axis] * rf, shape[out_axis] * rf
    variance = scale / ((fan_in + fan_out) / 2)
    with torch.no_grad():
        t.uniform_(-1.0, 1.0).mul_(math.sqrt(3 * variance))
    return t


def _conv(cin, cout, k, init_scale=1.0):
    """ddpm_conv3x3 / ddpm_conv1x1 (layers.py:112-156): DDPM init, zero bias."""
    m = nn.Conv2d(cin, cout, k, stride=1, padding=k // 2)
    _variance_scaling_(m.weight.data, init_scale)
    nn.init.zeros_(m.bias)
    return m


def _gn(c):
    return nn.GroupNorm(num_groups=min(c // 4, 32), num_channels=c, eps=1e-6)


class GaussianFourierProjection(nn.Module):
    """layerspp.py:37-46: fixed random frequencies W ~ N(0, scale^2)."""

    def __init__(self, embedding_size=256, scale=1.0):
        super().__init__()
        self.W = nn.Parameter(torch.randn(embedding_size) * scale, requires_grad=False)


class NIN(nn.Module):
    """layers.py:678-689: 1x1 channel mix y = x W + b with W of shape (in, out)."""

    def __init__(self, in_dim, num_units, init_scale=0.1):
        super().__init__()
        self.W = nn.Parameter(_variance_scaling_(torch.empty(in_dim, num_units), init_scale, in_axis=0, out_axis=1))
        self.b = nn.Parameter(torch.zeros(num_units))


class AttnBlockpp(nn.Module):
    """layerspp.py:67-97 parameter container."""

    def __init__(self, channels, skip_rescale=True, init_scale=0.0):
        super().__init__()
        self.GroupNorm_0 = _gn(channels)
        self.NIN_0 = NIN(channels, channels)
        self.NIN_1 = NIN(channels, channels)
        self.NIN_2 = NIN(channels, channels)
        self.NIN_3 = NIN(channels, channels, init_scale=init_scale)
        self.skip_rescale = skip_rescale


class Combine(nn.Module):
    """layerspp.py:49-64 (method "sum") parameter container."""

    def __init__(self, dim1, dim2, method="sum"):
        super().__init__()
        self.Conv_0 = _conv(dim1, dim2, 1)
        self.method = method


class ResnetBlockBigGANpp(nn.Module):
    """layerspp.py:251-328 parameter container."""

    def __init__(self, in_ch, out_ch, temb_dim, up=False, down=False, init_scale=0.0):
        super().__init__()
        self.GroupNorm_0 = _gn(in_ch)
        self.Conv_0 = _conv(in_ch, out_ch, 3)
        self.Dense_0 = nn.Linear(temb_dim, out_ch)
        _variance_scaling_(self.Dense_0.weight.data, 1.0)
        nn.init.zeros_(self.Dense_0.bias)
        self.GroupNorm_1 = _gn(out_ch)
        self.Dropout_0 = nn.Dropout(0.0)
        self.Conv_1 = _conv(out_ch, out_ch, 3, init_scale=init_scale)
        if in_ch != out_ch or up or down:
            self.Conv_2 = _conv(in_ch, out_ch, 1)
        self.up, self.down, self.in_ch, self.out_ch = up, down, in_ch, out_ch


class NCSNpp(nn.Module):
    """NCSN++ (ncsnpp.py:40-478).  ``precision`` ("bf16" | "fp32", default from
    ``$DITSEP_PRECISION`` or "bf16") is this repo's only extra argument: bf16 runs the wide
    convolutions on tcgen05 tensor cores, fp32 keeps every operand in fp32 (parity mode)."""

    @staticmethod
    def add_argparse_args(parser):
        return parser

    def __init__(self, scale_by_sigma=True, nonlinearity="swish", nf=128, ch_mult=(1, 2, 1, 1, 1), num_res_blocks=2,
                 attn_resolutions=(4, 8, 16), resamp_with_conv=True, conditional=True, fir=True,
                 fir_kernel=[1, 3, 3, 1], skip_rescale=True, resblock_type="biggan", progressive="output_skip",
                 progressive_input="input_skip", progressive_combine="sum", init_scale=0.0, fourier_scale=16,
                 image_size=64, num_channels_in=4, num_channels_out=4, embedding_type="fourier", dropout=0.0,
                 centered=False, precision=None, **unused_kwargs):
        super().__init__()
        unsupported = {
            "nonlinearity": (nonlinearity.lower(), ("swish", "silu")), "conditional": (conditional, (True,)),
            "fir": (fir, (True,)), "fir_kernel": (tuple(fir_kernel), ((1, 3, 3, 1),)),
            "skip_rescale": (skip_rescale, (True,)), "resblock_type": (resblock_type.lower(), ("biggan",)),
            "progressive": (progressive.lower(), ("output_skip",)),
            "progressive_input": (progressive_input.lower(), ("input_skip",)),
            "progressive_combine": (progressive_combine.lower(), ("sum",)),
            "embedding_type": (embedding_type.lower(), ("fourier",)), "dropout": (float(dropout), (0.0,)),
        }
        for k, (v, ok) in unsupported.items():
            if v not in ok:
                raise NotImplementedError(f"NCSNpp({k}={v!r}) has no sm_100a kernel path; supported: {ok}")
        self.nf, self.ch_mult, self.num_res_blocks = nf, tuple(ch_mult), num_res_blocks
        self.attn_resolutions = tuple(attn_resolutions)
        self.num_resolutions = len(ch_mult)
        self.all_resolutions = [image_size // (2 ** i) for i in range(self.num_resolutions)]
        self.conditional, self.centered, self.scale_by_sigma = conditional, centered, scale_by_sigma
        self.skip_rescale, self.resblock_type = skip_rescale, resblock_type.lower()
        self.progressive, self.progressive_input = progressive.lower(), progressive_input.lower()
        self.embedding_type = embedding_type.lower()
        self.image_size = image_size
        self.num_channels_in, self.num_channels_out = num_channels_in, num_channels_out
        self.precision = (precision or os.environ.get("DITSEP_PRECISION", "bf16")).lower()
        if self.precision not in ("bf16", "fp32"):
            raise ValueError(f"precision must be 'bf16' or 'fp32', got {self.precision!r}")

        self.output_layer = nn.Conv2d(num_channels_in, num_channels_out, 1)
        self.layout = layout_of(self)
        modules: List[nn.Module] = []
        for spec in self.layout:
            if spec.kind == "gfp":
                modules.append(GaussianFourierProjection(embedding_size=nf, scale=fourier_scale))
            elif spec.kind == "linear":
                lin = nn.Linear(spec.in_ch, spec.out_ch)
                _variance_scaling_(lin.weight.data, 1.0)
                nn.init.zeros_(lin.bias)
                modules.append(lin)
            elif spec.kind == "conv3x3":
                modules.append(_conv(spec.in_ch, spec.out_ch, 3, init_scale=1.0 if len(modules) == 3 else init_scale))
            elif spec.kind == "resblock":
                modules.append(ResnetBlockBigGANpp(spec.in_ch, spec.out_ch, nf * 4, up=spec.up, down=spec.down,
                                                   init_scale=init_scale))
            elif spec.kind == "attn":
                modules.append(AttnBlockpp(spec.in_ch, skip_rescale=True, init_scale=init_scale))
            elif spec.kind == "combine":
                modules.append(Combine(spec.in_ch, spec.out_ch, "sum"))
            elif spec.kind == "gn":
                modules.append(_gn(spec.in_ch))
            else:
                raise AssertionError(spec.kind)
        self.all_modules = nn.ModuleList(modules)
        self._engine = None

    def engine(self) -> NcsnppEngine:
        if self._engine is None:
            self._engine = NcsnppEngine(self)
        return self._engine

    def forward(self, x, time_cond):
        """x: (B, Cin, H, W) fp32 NCHW like the reference; returns (B, Cout, H, W) fp32.
        (``ScoreModelNCSNpp`` bypasses this NCHW entry and feeds NHWC features directly.)"""
        feat = x.permute(0, 2, 3, 1).contiguous().float()
        out = self.engine().forward_nhwc(feat, time_cond, affine_done=False)
        return out.permute(0, 3, 1, 2).contiguous()
