"""Oracle: NCSN++ score network + STFT wrapper, functional fp32 on CPU.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates, for the
configuration family the DiffSep configs use (biggan res-blocks, FIR
resampling, input_skip / output_skip / sum, fourier embedding, skip_rescale):

* ``NCSNpp.__init__`` / ``forward``      -- src/models/diffsep/ncsnpp.py:48-309, 320-478
* ``ResnetBlockBigGANpp.forward``        -- src/models/diffsep/ncsnpp_utils/layerspp.py:296-328
* ``AttnBlockpp.forward``                -- layerspp.py:81-97 ; ``NIN`` layers.py:678-689
* ``Combine``                            -- layerspp.py:49-64
* ``GaussianFourierProjection``          -- layerspp.py:37-46
* ``upsample_2d`` / ``downsample_2d``    -- up_or_down_sampling.py:206-273,
  through ``upfirdn2d_native``           -- op/upfirdn2d.py:159-200
* ``ScoreModelNCSNpp``                   -- src/models/diffsep/score_models.py:41-138
* ``LatentScoreModelNCSNpp.forward``     -- score_models.py:173-186

All functions take a ``state_dict``-style mapping with the reference's key
layout (``backbone.all_modules.{i}.<name>.{weight,bias}``) instead of modules.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------- config
@dataclass
class NCSNppConfig:
    """Constructor arguments of the reference ``NCSNpp`` (ncsnpp.py:48-74) that the
    DiffSep configs vary; defaults are the hot-path config
    (src/config/diffsep/model/default.yaml:25-30)."""

    nf: int = 64
    ch_mult: Sequence[int] = (1, 1, 2, 2, 2, 2, 2)
    num_res_blocks: int = 2
    attn_resolutions: Sequence[int] = (16,)
    image_size: int = 256
    num_channels_in: int = 6
    num_channels_out: int = 4
    fourier_scale: float = 16.0
    centered: bool = False
    scale_by_sigma: bool = True
    fir_kernel: Sequence[int] = (1, 3, 3, 1)

    @property
    def num_resolutions(self) -> int:
        return len(self.ch_mult)

    @property
    def all_resolutions(self) -> List[int]:
        return [self.image_size // (2 ** i) for i in range(self.num_resolutions)]


@dataclass
class ModuleSpec:
    """One entry of ``NCSNpp.all_modules`` (ncsnpp.py:107-309)."""

    kind: str  # gfp | linear | conv3x3 | resblock | attn | combine | gn
    in_ch: int = 0
    out_ch: int = 0
    up: bool = False
    down: bool = False


def module_layout(cfg: NCSNppConfig) -> List[ModuleSpec]:
    """Replays the module-list construction of ncsnpp.py:107-309 (biggan /
    input_skip / output_skip / sum branch only) and returns the spec of each
    ``all_modules[i]``."""
    nf = cfg.nf
    mods: List[ModuleSpec] = [
        ModuleSpec("gfp", out_ch=nf),
        ModuleSpec("linear", 2 * nf, 4 * nf),
        ModuleSpec("linear", 4 * nf, 4 * nf),
    ]
    channels = cfg.num_channels_in
    mods.append(ModuleSpec("conv3x3", channels, nf))
    hs_c = [nf]
    in_ch = nf
    R = cfg.num_resolutions
    res = cfg.all_resolutions
    for i_level in range(R):
        for _ in range(cfg.num_res_blocks):
            out_ch = nf * cfg.ch_mult[i_level]
            mods.append(ModuleSpec("resblock", in_ch, out_ch))
            in_ch = out_ch
            if res[i_level] in cfg.attn_resolutions:
                mods.append(ModuleSpec("attn", in_ch, in_ch))
            hs_c.append(in_ch)
        if i_level != R - 1:
            mods.append(ModuleSpec("resblock", in_ch, in_ch, down=True))
            mods.append(ModuleSpec("combine", channels, in_ch))
            hs_c.append(in_ch)
    in_ch = hs_c[-1]
    mods.append(ModuleSpec("resblock", in_ch, in_ch))
    mods.append(ModuleSpec("attn", in_ch, in_ch))
    mods.append(ModuleSpec("resblock", in_ch, in_ch))
    for i_level in reversed(range(R)):
        for _ in range(cfg.num_res_blocks + 1):
            out_ch = nf * cfg.ch_mult[i_level]
            mods.append(ModuleSpec("resblock", in_ch + hs_c.pop(), out_ch))
            in_ch = out_ch
        if res[i_level] in cfg.attn_resolutions:
            mods.append(ModuleSpec("attn", in_ch, in_ch))
        mods.append(ModuleSpec("gn", in_ch, in_ch))
        mods.append(ModuleSpec("conv3x3", in_ch, channels))
        if i_level != 0:
            mods.append(ModuleSpec("resblock", in_ch, in_ch, up=True))
    assert not hs_c
    return mods


# --------------------------------------------------------------------------- blocks
def _setup_kernel(k: Sequence[int]) -> np.ndarray:
    """up_or_down_sampling.py:192-199."""
    k = np.asarray(k, dtype=np.float32)
    if k.ndim == 1:
        k = np.outer(k, k)
    k /= np.sum(k)
    return k


def upfirdn2d_ref(x: Tensor, kernel: Tensor, up: int, down: int, pad: Tuple[int, int]) -> Tensor:
    """``upfirdn2d_native`` (op/upfirdn2d.py:159-200): zero-insert by ``up``,
    zero-pad by ``pad``, correlate with the flipped kernel, decimate by ``down``.
    x: (N, C, H, W)."""
    n, c, h, w = x.shape
    kh, kw = kernel.shape
    p0, p1 = pad
    z = x.reshape(n * c, 1, h, w)
    if up > 1:
        up_buf = z.new_zeros(n * c, 1, h * up, w * up)
        up_buf[:, :, ::up, ::up] = z
        z = up_buf
    z = F.pad(z, [p0, p1, p0, p1])
    wgt = torch.flip(kernel, [0, 1]).view(1, 1, kh, kw)
    z = F.conv2d(z, wgt)
    z = z[:, :, ::down, ::down]
    return z.reshape(n, c, z.shape[-2], z.shape[-1])


def upsample_2d_ref(x: Tensor, k: Sequence[int] = (1, 3, 3, 1), factor: int = 2) -> Tensor:
    """up_or_down_sampling.py:206-239 (gain 1 -> kernel scaled by factor**2,
    pad ((p+1)//2 + factor - 1, p//2) with p = len(k) - factor)."""
    kk = _setup_kernel(k) * (factor ** 2)
    p = kk.shape[0] - factor
    return upfirdn2d_ref(x, torch.tensor(kk, device=x.device, dtype=x.dtype), up=factor, down=1,
                         pad=((p + 1) // 2 + factor - 1, p // 2))


def downsample_2d_ref(x: Tensor, k: Sequence[int] = (1, 3, 3, 1), factor: int = 2) -> Tensor:
    """up_or_down_sampling.py:242-273 (pad ((p+1)//2, p//2))."""
    kk = _setup_kernel(k)
    p = kk.shape[0] - factor
    return upfirdn2d_ref(x, torch.tensor(kk, device=x.device, dtype=x.dtype), up=1, down=factor,
                         pad=((p + 1) // 2, p // 2))


def _gn(x: Tensor, sd, key: str) -> Tensor:
    c = x.shape[1]
    return F.group_norm(x, min(c // 4, 32), sd[key + ".weight"], sd[key + ".bias"], eps=1e-6)


def _conv(x: Tensor, sd, key: str, pad: int) -> Tensor:
    return F.conv2d(x, sd[key + ".weight"], sd[key + ".bias"], padding=pad)


def _nin(x: Tensor, sd, key: str) -> Tensor:
    """layers.py:678-689: y[b,o,h,w] = sum_c x[b,c,h,w] W[c,o] + b[o]."""
    return torch.einsum("bchw,co->bohw", x, sd[key + ".W"]) + sd[key + ".b"][None, :, None, None]


def resblock_ref(x: Tensor, temb: Tensor, sd, key: str, spec: ModuleSpec,
                 fir_kernel=(1, 3, 3, 1), trace: Optional[dict] = None) -> Tensor:
    """``ResnetBlockBigGANpp.forward`` layerspp.py:296-328 (skip_rescale=True,
    dropout 0)."""
    h = F.silu(_gn(x, sd, key + ".GroupNorm_0"))
    if spec.up:
        h = upsample_2d_ref(h, fir_kernel)
        x = upsample_2d_ref(x, fir_kernel)
    elif spec.down:
        h = downsample_2d_ref(h, fir_kernel)
        x = downsample_2d_ref(x, fir_kernel)
    h = _conv(h, sd, key + ".Conv_0", 1)
    h = h + F.linear(F.silu(temb), sd[key + ".Dense_0.weight"], sd[key + ".Dense_0.bias"])[:, :, None, None]
    if trace is not None:
        trace[key + ".h1"] = h
    h = F.silu(_gn(h, sd, key + ".GroupNorm_1"))
    h = _conv(h, sd, key + ".Conv_1", 1)
    if spec.in_ch != spec.out_ch or spec.up or spec.down:
        x = _conv(x, sd, key + ".Conv_2", 0)
    return (x + h) / np.sqrt(2.0)


def attn_ref(x: Tensor, sd, key: str) -> Tensor:
    """``AttnBlockpp.forward`` layerspp.py:81-97 (skip_rescale=True)."""
    b, c, hh, ww = x.shape
    h = _gn(x, sd, key + ".GroupNorm_0")
    q = _nin(h, sd, key + ".NIN_0").reshape(b, c, hh * ww)
    k = _nin(h, sd, key + ".NIN_1").reshape(b, c, hh * ww)
    v = _nin(h, sd, key + ".NIN_2").reshape(b, c, hh * ww)
    w = torch.einsum("bcq,bck->bqk", q, k) * (int(c) ** (-0.5))
    w = F.softmax(w, dim=-1)
    h = torch.einsum("bqk,bck->bcq", w, v).reshape(b, c, hh, ww)
    h = _nin(h, sd, key + ".NIN_3")
    return (x + h) / np.sqrt(2.0)


def temb_ref(sd, prefix: str, time_cond: Tensor) -> Tensor:
    """ncsnpp.py:325-346 + GaussianFourierProjection layerspp.py:44-46: fourier
    features of log(t), Linear, SiLU, Linear."""
    am = prefix + "all_modules."
    x_proj = torch.log(time_cond)[:, None] * sd[am + "0.W"][None, :] * 2 * np.pi
    temb = torch.cat([torch.sin(x_proj), torch.cos(x_proj)], dim=-1)
    temb = F.linear(temb, sd[am + "1.weight"], sd[am + "1.bias"])
    temb = F.linear(F.silu(temb), sd[am + "2.weight"], sd[am + "2.bias"])
    return temb


def ncsnpp_forward_ref(sd: Dict[str, Tensor], cfg: NCSNppConfig, x: Tensor, time_cond: Tensor,
                       prefix: str = "backbone.", trace: Optional[dict] = None) -> Tensor:
    """``NCSNpp.forward`` ncsnpp.py:320-478.  x: (B, Cin, H, W) fp32, time_cond: (B,)."""
    mods = module_layout(cfg)
    am = prefix + "all_modules."
    R = cfg.num_resolutions

    def key(i):
        return am + str(i)

    def rec(name, val):
        if trace is not None:
            trace[name] = val
        return val

    temb = rec("temb", temb_ref(sd, prefix, time_cond))
    m = 3
    if not cfg.centered:
        x = 2 * x - 1.0  # ncsnpp.py:348-350
    input_pyramid = x
    hs = [rec(key(m), _conv(x, sd, key(m), 1))]
    m += 1
    for i_level in range(R):
        for _ in range(cfg.num_res_blocks):
            h = rec(key(m), resblock_ref(hs[-1], temb, sd, key(m), mods[m], cfg.fir_kernel, trace))
            m += 1
            if h.shape[-2] in cfg.attn_resolutions:
                h = rec(key(m), attn_ref(h, sd, key(m)))
                m += 1
            hs.append(h)
        if i_level != R - 1:
            h = rec(key(m), resblock_ref(hs[-1], temb, sd, key(m), mods[m], cfg.fir_kernel, trace))
            m += 1
            input_pyramid = downsample_2d_ref(input_pyramid, cfg.fir_kernel)
            # Combine(method="sum") layerspp.py:57-62
            h = rec(key(m), _conv(input_pyramid, sd, key(m) + ".Conv_0", 0) + h)
            m += 1
            hs.append(h)
    h = hs[-1]
    h = rec(key(m), resblock_ref(h, temb, sd, key(m), mods[m], cfg.fir_kernel, trace)); m += 1
    h = rec(key(m), attn_ref(h, sd, key(m))); m += 1
    h = rec(key(m), resblock_ref(h, temb, sd, key(m), mods[m], cfg.fir_kernel, trace)); m += 1
    pyramid = None
    for i_level in reversed(range(R)):
        for _ in range(cfg.num_res_blocks + 1):
            h = rec(key(m), resblock_ref(torch.cat([h, hs.pop()], dim=1), temb, sd, key(m), mods[m],
                                         cfg.fir_kernel, trace))
            m += 1
        if h.shape[-2] in cfg.attn_resolutions:
            h = rec(key(m), attn_ref(h, sd, key(m)))
            m += 1
        pyramid_h = F.silu(_gn(h, sd, key(m))); m += 1
        pyramid_h = _conv(pyramid_h, sd, key(m), 1)
        if i_level == R - 1:
            pyramid = pyramid_h
        else:
            pyramid = upsample_2d_ref(pyramid, cfg.fir_kernel) + pyramid_h
        rec(key(m), pyramid)
        m += 1
        if i_level != 0:
            h = rec(key(m), resblock_ref(h, temb, sd, key(m), mods[m], cfg.fir_kernel, trace))
            m += 1
    assert not hs and m == len(mods)
    h = pyramid
    if cfg.scale_by_sigma:
        h = h / time_cond.reshape(-1, 1, 1, 1)  # ncsnpp.py:473-475 (divides by t itself)
    return F.conv2d(h, sd[prefix + "output_layer.weight"], sd[prefix + "output_layer.bias"])


# --------------------------------------------------------------------------- STFT wrapper
@dataclass
class ScoreConfig:
    """``ScoreModelNCSNpp`` arguments (score_models.py:11-39; values from
    src/config/diffsep/model/default.yaml:15-35)."""

    num_sources: int = 2
    n_fft: int = 510
    hop_length: int = 128
    spec_abs_exponent: float = 0.5
    spec_factor: float = 0.15
    backbone: NCSNppConfig = field(default_factory=NCSNppConfig)

    def __post_init__(self):
        # score_models.py:24-26
        self.backbone.num_channels_in = 2 * self.num_sources + 2
        self.backbone.num_channels_out = 2 * self.num_sources


def hann_window(n_fft: int) -> Tensor:
    """torchaudio Spectrogram default window = torch.hann_window(win_length), periodic."""
    n = torch.arange(n_fft, dtype=torch.float64)
    return (0.5 - 0.5 * torch.cos(2 * math.pi * n / n_fft)).to(torch.float32)


def stft_ref(x: Tensor, n_fft: int, hop: int, window: Tensor) -> Tensor:
    """torchaudio ``Spectrogram(power=None, center=True, pad_mode="constant")``
    (score_models.py:29,112): centre zero-padding of n_fft//2, onesided,
    unnormalised.  x: (..., T) -> (..., n_fft//2+1, frames) complex64."""
    shp = x.shape
    z = x.reshape(-1, shp[-1])
    z = F.pad(z, (n_fft // 2, n_fft // 2))
    spec = torch.stft(z, n_fft, hop_length=hop, win_length=n_fft, window=window, center=False,
                      normalized=False, onesided=True, return_complex=True)
    return spec.reshape(*shp[:-1], spec.shape[-2], spec.shape[-1])


def istft_ref(spec: Tensor, n_fft: int, hop: int, window: Tensor) -> Tensor:
    """torchaudio ``InverseSpectrogram`` (score_models.py:30,122) = torch.istft with
    center=True: windowed overlap-add divided by the squared-window envelope, n_fft//2
    trimmed on both ends, length hop*(frames-1)."""
    shp = spec.shape
    z = spec.reshape(-1, shp[-2], shp[-1])
    y = torch.istft(z, n_fft, hop_length=hop, win_length=n_fft, window=window, center=True,
                    normalized=False, onesided=True, length=None)
    return y.reshape(*shp[:-2], y.shape[-1])


def transform_forward_ref(spec: Tensor, e: float, factor: float) -> Tensor:
    """score_models.py:41-57, transform="exponent"."""
    if e != 1:
        spec = spec.abs() ** abs(e) * torch.exp(1j * spec.angle())
    return spec * factor


def transform_backward_ref(spec: Tensor, e: float, factor: float) -> Tensor:
    """score_models.py:59-70, transform="exponent"."""
    spec = spec / abs(factor)
    if e != 1:
        spec = spec.abs() ** (1 / abs(e)) * torch.exp(1j * spec.angle())
    return spec


def pre_process_ref(x: Tensor, cfg: ScoreConfig, window: Tensor) -> Tuple[Tensor, int, int]:
    """score_models.py:107-116: right-pad n_fft-hop, STFT, compress, stack
    [Re c0.., Im c0..] (complex_to_real :72-76), zero-pad frames to a multiple
    of 64 (:83-91)."""
    n_samples = x.shape[-1]
    x = F.pad(x, (0, cfg.n_fft - cfg.hop_length))
    spec = stft_ref(x, cfg.n_fft, cfg.hop_length, window)
    spec = transform_forward_ref(spec, cfg.spec_abs_exponent, cfg.spec_factor)
    feat = torch.cat((spec.real, spec.imag), dim=1)
    rem = feat.shape[-1] % 64
    n_pad = 0 if rem == 0 else 64 - rem
    if n_pad:
        feat = F.pad(feat, (0, n_pad))
    return feat, n_samples, n_pad


def post_process_ref(x: Tensor, n_samples: int, n_pad: int, cfg: ScoreConfig, window: Tensor) -> Tensor:
    """score_models.py:118-124: unpad, real_to_complex (:78-81: channel k pairs with
    k + C/2), expand, iSTFT, trim / zero-pad to n_samples (:99-105)."""
    if n_pad:
        x = x[..., :-n_pad]
    half = x.shape[1] // 2
    spec = torch.complex(x[:, :half].contiguous(), x[:, half:].contiguous())
    spec = transform_backward_ref(spec, cfg.spec_abs_exponent, cfg.spec_factor)
    y = istft_ref(spec, cfg.n_fft, cfg.hop_length, window)
    if y.shape[-1] < n_samples:
        y = F.pad(y, (0, n_samples - y.shape[-1]))
    elif y.shape[-1] > n_samples:
        y = y[..., :n_samples]
    return y


def score_forward_ref(sd: Dict[str, Tensor], cfg: ScoreConfig, xt: Tensor, time_cond: Tensor,
                      mix: Tensor, trace: Optional[dict] = None) -> Tensor:
    """``ScoreModelNCSNpp.forward`` score_models.py:126-138."""
    window = sd.get("stft.window", None)
    if window is None:
        window = hann_window(cfg.n_fft)
    x = torch.cat((xt, mix), dim=1)
    feat, n_samples, n_pad = pre_process_ref(x, cfg, window)
    if trace is not None:
        trace["features"] = feat
    out = ncsnpp_forward_ref(sd, cfg.backbone, feat, time_cond, "backbone.", trace)
    if trace is not None:
        trace["backbone_out"] = out
    return post_process_ref(out, n_samples, n_pad, cfg, window)


def latent_score_forward_ref(sd: Dict[str, Tensor], bcfg: NCSNppConfig, xt: Tensor, time_cond: Tensor,
                             mix: Tensor, max_latent_length: int = 16) -> Tensor:
    """``LatentScoreModelNCSNpp.forward`` score_models.py:173-186: cat, pad the last
    axis to a multiple of max_latent_length, backbone, unpad."""
    x = torch.cat((xt, mix), dim=1)
    rem = x.shape[-1] % max_latent_length
    n_pad = 0 if rem == 0 else max_latent_length - rem
    if n_pad:
        x = F.pad(x, (0, n_pad))
    out = ncsnpp_forward_ref(sd, bcfg, x, time_cond, "backbone.")
    return out[..., :-n_pad] if n_pad else out


# --------------------------------------------------------------------------- weights
def _fan_avg_uniform(shape, scale, gen, in_axis=1, out_axis=0) -> Tensor:
    """``default_init`` = variance_scaling(scale, "fan_avg", "uniform") layers.py:48-102."""
    scale = 1e-10 if scale == 0 else scale
    rf = float(np.prod(shape)) / shape[in_axis] / shape[out_axis]
    fan_in, fan_out = shape[in_axis] * rf, shape[out_axis] * rf
    variance = scale / ((fan_in + fan_out) / 2)
    return (torch.rand(*shape, generator=gen) * 2.0 - 1.0) * math.sqrt(3 * variance)


def make_state_dict(cfg: NCSNppConfig, seed: int = 0, prefix: str = "backbone.",
                    init_scale: float = 1.0, with_window: Optional[int] = 510) -> Dict[str, Tensor]:
    """Deterministic random weights in the reference's key layout.

    Distribution follows the reference initialisers (default_init fan-avg uniform,
    layers.py:99-102; NIN scale 0.1, layers.py:679-684; GFP randn*scale,
    layerspp.py:42) EXCEPT that the ``init_scale=0`` tensors (Conv_1, NIN_3, pyramid
    convs) are drawn at ``init_scale`` (default 1.0) so residual branches are
    numerically visible, and biases / GroupNorm affine
    parameters are perturbed so that they are exercised.  Drawn from one seeded
    CPU generator in module order, so the same dict is regenerated anywhere."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    am = prefix + "all_modules."

    def conv(key, cin, cout, k, scale=1.0):
        sd[key + ".weight"] = _fan_avg_uniform((cout, cin, k, k), scale, g)
        sd[key + ".bias"] = 0.05 * torch.randn(cout, generator=g)

    def gn(key, c):
        sd[key + ".weight"] = 1.0 + 0.1 * torch.randn(c, generator=g)
        sd[key + ".bias"] = 0.1 * torch.randn(c, generator=g)

    def linear(key, cin, cout):
        sd[key + ".weight"] = _fan_avg_uniform((cout, cin), 1.0, g)
        sd[key + ".bias"] = 0.05 * torch.randn(cout, generator=g)

    def nin(key, cin, cout, scale):
        sd[key + ".W"] = _fan_avg_uniform((cin, cout), scale, g)
        sd[key + ".b"] = 0.05 * torch.randn(cout, generator=g)

    for i, spec in enumerate(module_layout(cfg)):
        key = am + str(i)
        if spec.kind == "gfp":
            sd[key + ".W"] = torch.randn(spec.out_ch, generator=g) * cfg.fourier_scale
        elif spec.kind == "linear":
            linear(key, spec.in_ch, spec.out_ch)
        elif spec.kind == "conv3x3":
            conv(key, spec.in_ch, spec.out_ch, 3, 1.0 if i == 3 else init_scale)
        elif spec.kind == "gn":
            gn(key, spec.in_ch)
        elif spec.kind == "combine":
            conv(key + ".Conv_0", spec.in_ch, spec.out_ch, 1)
        elif spec.kind == "attn":
            gn(key + ".GroupNorm_0", spec.in_ch)
            for j in range(3):
                nin(f"{key}.NIN_{j}", spec.in_ch, spec.in_ch, 0.1)
            nin(f"{key}.NIN_3", spec.in_ch, spec.in_ch, init_scale)
        elif spec.kind == "resblock":
            gn(key + ".GroupNorm_0", spec.in_ch)
            conv(key + ".Conv_0", spec.in_ch, spec.out_ch, 3)
            linear(key + ".Dense_0", 4 * cfg.nf, spec.out_ch)
            gn(key + ".GroupNorm_1", spec.out_ch)
            conv(key + ".Conv_1", spec.out_ch, spec.out_ch, 3, init_scale)
            if spec.in_ch != spec.out_ch or spec.up or spec.down:
                conv(key + ".Conv_2", spec.in_ch, spec.out_ch, 1)
        else:
            raise AssertionError(spec.kind)
    conv(prefix + "output_layer", cfg.num_channels_in, cfg.num_channels_out, 1)
    if with_window:
        sd["stft.window"] = hann_window(with_window)
        sd["stft_inv.window"] = hann_window(with_window)
    return sd
