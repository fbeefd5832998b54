"""TEST INFRASTRUCTURE ONLY (tests/ and the pin script import this; the product never does).

Functional CPU restatement (torch, fp32) of the Oobleck VAE the latent DiffSep model wraps around its sampler
(SURVEY.md 8 f3): LatentDiffSep.encode / decode (src/diffsep_latent.py:107-128) over
stable_audio_tools.models.autoencoders.OobleckEncoder / OobleckDecoder (:281-357) with the `oobleck_finetune.json`
configuration: ELU activations (use_snake absent), weight-normalised 1-D convolutions, residual units with dilations
1 / 3 / 9, strided convolutions down and transposed convolutions up, VAE bottleneck (bottleneck.py:57-86), final tanh.
Works from a plain state_dict with the reference's key layout (weight_g / weight_v of torch.nn.utils.weight_norm).
Pinned against the unmodified reference classes by oracle/pin_oobleck.py (tests/golden/oobleck.npz)."""
import math
from typing import Dict, List, Sequence

import torch
import torch.nn.functional as F
from torch import Tensor


def wn_weight(sd: Dict[str, Tensor], key: str) -> Tensor:
    """torch.nn.utils.weight_norm (dim 0): w = g * v / ||v|| per slice along dim 0 (for ConvTranspose1d that is the
    INPUT channel axis, as torch does it)."""
    if key + ".weight" in sd:
        return sd[key + ".weight"]
    v, g = sd[key + ".weight_v"], sd[key + ".weight_g"]
    return v * (g / v.flatten(1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1))))


def _conv(sd, key, x, stride=1, padding=0, dilation=1):
    return F.conv1d(x, wn_weight(sd, key), sd.get(key + ".bias"), stride=stride, padding=padding, dilation=dilation)


def residual_unit(sd, key, x, dilation):
    """autoencoders.py:59-85: x + conv1(ELU(conv7_dilated(ELU(x))))."""
    h = _conv(sd, key + ".layers.1", F.elu(x), padding=3 * dilation, dilation=dilation)
    return x + _conv(sd, key + ".layers.3", F.elu(h))


def encoder_forward(sd, x: Tensor, strides: Sequence[int], prefix: str = "encoder.") -> Tensor:
    """OobleckEncoder.forward (:281-311): conv7, EncoderBlocks (:228-244: three residual units, ELU, strided conv of kernel
    2 s / padding ceil(s / 2)), ELU, conv3."""
    p = prefix + "layers."
    h = _conv(sd, p + "0", x, padding=3)
    for i, s in enumerate(strides):
        b = f"{p}{i + 1}.layers."
        for j, d in enumerate((1, 3, 9)):
            h = residual_unit(sd, f"{b}{j}", h, d)
        h = _conv(sd, b + "4", F.elu(h), stride=s, padding=math.ceil(s / 2))
    return _conv(sd, f"{p}{len(strides) + 2}", F.elu(h), padding=1)


def vae_sample(latents: Tensor, noise: Tensor) -> Tensor:
    """bottleneck.py:57-65 (VAEBottleneck.encode): mean, scale = chunk(2); z = noise * (softplus(scale) + 1e-4) + mean."""
    mean, scale = latents.chunk(2, dim=1)
    return noise * (F.softplus(scale) + 1e-4) + mean


def decoder_forward(sd, z: Tensor, strides: Sequence[int], prefix: str = "decoder.", final_tanh: bool = True) -> Tensor:
    """OobleckDecoder.forward (:314-357): conv7, DecoderBlocks in reverse stride order (:246-279: ELU, transposed conv of
    kernel 2 s / stride s / padding ceil(s / 2), three residual units), ELU, conv7 without bias, tanh."""
    p = prefix + "layers."
    h = _conv(sd, p + "0", z, padding=3)
    for i, s in enumerate(reversed(list(strides))):
        b = f"{p}{i + 1}.layers."
        h = F.conv_transpose1d(F.elu(h), wn_weight(sd, b + "1"), sd.get(b + "1.bias"), stride=s, padding=math.ceil(s / 2))
        for j, d in enumerate((1, 3, 9)):
            h = residual_unit(sd, f"{b}{j + 2}", h, d)
    h = _conv(sd, f"{p}{len(strides) + 2}", F.elu(h), padding=3)
    return torch.tanh(h) if final_tanh else h


def pad_to_hop(x: Tensor, hop: int) -> Tensor:
    """src/utils/torch_utils.py:12-18: ALWAYS pads (a full hop when the length is already a multiple)."""
    return F.pad(x, (0, hop - x.shape[-1] % hop))


def latent_encode(sd, mix: Tensor, strides, noise: Tensor) -> Tensor:
    """LatentDiffSep.encode for the mixture (diffsep_latent.py:107-111): (B, 1, T) -> (B, 1, latent, T')."""
    hop = int(torch.tensor(list(strides)).prod())
    return vae_sample(encoder_forward(sd, pad_to_hop(mix, hop), strides), noise).unsqueeze(1)


def latent_decode(sd, est: Tensor, strides, target_dim=None) -> Tensor:
    """LatentDiffSep.decode (diffsep_latent.py:120-128): (B, n_src, latent, T') -> (B, n_src, T)."""
    B, n, C, L = est.shape
    out = decoder_forward(sd, est.reshape(B * n, C, L), strides).reshape(B, n, -1)
    return out[..., :target_dim] if target_dim is not None else out


def make_state_dict(channels: int, c_mults: Sequence[int], strides: Sequence[int], enc_latent: int, seed: int = 0):
    """Seeded weights in the reference's state_dict layout (encoder.layers.N... / decoder.layers.N..., weight_g / weight_v
    of torch.nn.utils.weight_norm), conditioned so that every layer stays audible in the output: weight-norm gains ~1 (a
    convolution preserves variance), x0.3 on the closing 1x1 convolution of every residual unit (15 units in a row do not
    blow the signal up), x0.4 on each model's last convolution (the decoder's tanh is not saturated).  The pin script
    loads it into the reference classes with strict=True, which is what checks the key layout and the shapes."""
    g = torch.Generator().manual_seed(seed)
    cm = [1] + list(c_mults)
    sd: Dict[str, Tensor] = {}

    def conv(key, cout, cin, k, bias=True, gain=1.0, transpose=False):
        shape = (cin, cout, k) if transpose else (cout, cin, k)   # ConvTranspose1d stores (in, out, k); weight_norm dim 0
        sd[key + ".weight_v"] = torch.randn(shape, generator=g)
        sd[key + ".weight_g"] = gain * (0.8 + 0.4 * torch.rand((shape[0], 1, 1), generator=g))
        if bias:
            sd[key + ".bias"] = 0.05 * torch.randn(cout, generator=g)

    def runit(key, c):
        conv(key + ".layers.1", c, c, 7)
        conv(key + ".layers.3", c, c, 1, gain=0.3)

    n = len(strides)
    conv("encoder.layers.0", cm[0] * channels, 1, 7)
    for i, s in enumerate(strides):
        cin, cout = cm[i] * channels, cm[i + 1] * channels
        for j in range(3):
            runit(f"encoder.layers.{i + 1}.layers.{j}", cin)
        conv(f"encoder.layers.{i + 1}.layers.4", cout, cin, 2 * s)
    conv(f"encoder.layers.{n + 2}", enc_latent, cm[-1] * channels, 3, gain=0.4)
    conv("decoder.layers.0", cm[-1] * channels, enc_latent // 2, 7)
    for idx, i in enumerate(range(n, 0, -1)):
        cin, cout, s = cm[i] * channels, cm[i - 1] * channels, strides[i - 1]
        conv(f"decoder.layers.{idx + 1}.layers.1", cout, cin, 2 * s, transpose=True)
        for j in range(3):
            runit(f"decoder.layers.{idx + 1}.layers.{j + 2}", cout)
    conv(f"decoder.layers.{n + 2}", 1, cm[0] * channels, 7, bias=False, gain=0.4)
    return sd
