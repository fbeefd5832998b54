"""Waveform <-> compressed-spectrogram wrappers around the NCSN++ backbone
(mirror of src/models/diffsep/score_models.py).

``ScoreModelNCSNpp`` (:10-138) and ``LatentScoreModelNCSNpp`` (:140-186) keep the reference
constructor arguments, attributes (``backbone``, ``stft``, ``stft_inv``, ``stft_args``),
helper methods and ``state_dict`` keys (``stft.window``, ``stft_inv.window``,
``backbone.*``).  One ``forward`` is: fused STFT+compress kernel -> NCSN++ launch plan ->
fused expand+iSTFT kernel; the spectrogram never exists as a complex tensor and is written
straight into the backbone's NHWC input buffer.
"""
from __future__ import annotations

import importlib

import torch

from ... import ops

try:  # hydra is optional: the reference builds the backbone with hydra.utils.instantiate
    from hydra.utils import instantiate as _hydra_instantiate  # type: ignore
except Exception:  # pragma: no cover - hydra is not installed in the build image
    _hydra_instantiate = None

# reference dotted paths -> this package (configs keep their ``_target_`` strings)
_TARGET_ALIASES = {
    "models.diffsep.ncsnpp.NCSNpp": "ditsep_b200.models.diffsep.ncsnpp.NCSNpp",
    "src.models.diffsep.ncsnpp.NCSNpp": "ditsep_b200.models.diffsep.ncsnpp.NCSNpp",
    "models.diffsep.score_models.ScoreModelNCSNpp": "ditsep_b200.models.diffsep.score_models.ScoreModelNCSNpp",
    "models.diffsep.score_models.LatentScoreModelNCSNpp":
        "ditsep_b200.models.diffsep.score_models.LatentScoreModelNCSNpp",
    "sdes.MixSDE": "ditsep_b200.sdes.MixSDE", "sdes.sdes.MixSDE": "ditsep_b200.sdes.MixSDE",
    "sdes.OUVESDE": "ditsep_b200.sdes.OUVESDE", "sdes.sdes.OUVESDE": "ditsep_b200.sdes.OUVESDE",
    "sdes.PriorMixSDE": "ditsep_b200.sdes.PriorMixSDE", "sdes.sdes.PriorMixSDE": "ditsep_b200.sdes.PriorMixSDE",
    "sdes.SBVESDE": "ditsep_b200.sdes.SBVESDE", "sdes.sdes.SBVESDE": "ditsep_b200.sdes.SBVESDE",
}


def instantiate(cfg, *args, **kwargs):
    """``hydra.utils.instantiate`` for the ``_target_`` dicts of the DiffSep configs
    (score_models.py:27, diffsep.py:39,44).  Reference dotted paths are mapped onto this
    package; nested ``_target_`` dicts are instantiated unless ``_recursive_=False``."""
    cfg = dict(cfg)
    recursive = kwargs.pop("_recursive_", cfg.pop("_recursive_", True))
    target = cfg.pop("_target_")
    target = _TARGET_ALIASES.get(target, target)
    mod, name = target.rsplit(".", 1)
    cls = getattr(importlib.import_module(mod), name)
    if recursive:
        cfg = {k: (instantiate(v) if isinstance(v, dict) and "_target_" in v else v) for k, v in cfg.items()}
    return cls(*args, **cfg, **kwargs)


class Spectrogram(torch.nn.Module):
    """Stand-in for ``torchaudio.transforms.Spectrogram(power=None, **stft_args)``
    (score_models.py:29): owns the ``window`` buffer and the STFT geometry; the transform through the
    fused kernels happens inside ``ScoreModelNCSNpp``."""

    def __init__(self, n_fft=400, hop_length=None, win_length=None, center=True, pad_mode="reflect", **unused):
        super().__init__()
        self.n_fft = n_fft
        self.win_length = win_length or n_fft
        self.hop_length = hop_length or self.win_length // 2
        self.center, self.pad_mode = center, pad_mode
        self.register_buffer("window", torch.hann_window(self.win_length), persistent=True)

    def check(self):
        if (self.n_fft, self.hop_length, self.win_length, self.center, self.pad_mode) != (510, 128, 510, True, "constant"):
            raise NotImplementedError("the fused sm_100a STFT kernels implement n_fft=510, hop_length=128, center=True, "
                                      f"pad_mode='constant' (every DiffSep config); got n_fft={self.n_fft}, "
                                      f"hop={self.hop_length}, win={self.win_length}, center={self.center}, "
                                      f"pad_mode={self.pad_mode!r}")

    def forward(self, x):
        raise NotImplementedError("stand-alone Spectrogram.forward is not part of the sampler path; "
                                  "use ScoreModelNCSNpp.pre_process")


class InverseSpectrogram(Spectrogram):
    """Stand-in for ``torchaudio.transforms.InverseSpectrogram(**stft_args)`` (score_models.py:30)."""

    def forward(self, x):
        raise NotImplementedError("stand-alone InverseSpectrogram.forward is not part of the sampler path; "
                                  "use ScoreModelNCSNpp.post_process")


def _frames(n_samples: int) -> int:
    return 1 + (n_samples + 510 - 128) // 128


class ScoreModelNCSNpp(torch.nn.Module):
    def __init__(self, num_sources, stft_args, backbone_args, transform="exponent", spec_abs_exponent=0.5,
                 spec_factor=3.0, spec_trans_learnable=False):
        super().__init__()
        # infer input / output channels of the backbone from the number of sources; like the
        # reference this MUTATES backbone_args (score_models.py:24-26)
        backbone_args.update(num_channels_in=2 * num_sources + 2, num_channels_out=2 * num_sources)
        self.backbone = (_hydra_instantiate or instantiate)(
            {**backbone_args, "_target_": _TARGET_ALIASES.get(backbone_args["_target_"], backbone_args["_target_"])})
        self.num_sources = num_sources
        self.stft_args = stft_args
        self.stft = Spectrogram(**stft_args)
        self.stft_inv = InverseSpectrogram(**stft_args)
        if transform not in ("exponent", "log", "none"):
            raise ValueError("transform must be one of 'exponent'|'log'|'none'")
        self.transform = transform
        self.spec_abs_exponent = spec_abs_exponent
        self.spec_factor = spec_factor
        if spec_trans_learnable:
            self.spec_abs_exponent = torch.nn.Parameter(torch.tensor(self.spec_abs_exponent))
            self.spec_factor = torch.nn.Parameter(torch.tensor(spec_factor))

    # ------------------------------------------------------------------ compression constants
    def _exp_factor(self):
        if self.transform == "exponent":
            return abs(float(self.spec_abs_exponent)), float(self.spec_factor)
        if self.transform == "none":
            return 1.0, 1.0
        if self.transform == "log":  # log1p(|X|) e^{j arg X} * |factor| (score_models.py:49-51,66-68): exponent 0 in the C ABI
            return 0.0, abs(float(self.spec_factor))
        raise ValueError("transform must be one of 'exponent'|'log'|'none'")

    # ------------------------------------------------------------------ reference helper API
    def pad(self, x):
        rem = x.shape[-1] % 64
        if rem == 0:
            return x, 0
        return torch.nn.functional.pad(x, (0, 64 - rem)), 64 - rem

    def unpad(self, x, pad):
        return x if pad == 0 else x[..., :-pad]

    def adjust_length(self, x, n_samples):
        if x.shape[-1] < n_samples:
            return torch.nn.functional.pad(x, (0, n_samples - x.shape[-1]))
        return x[..., :n_samples]

    def pre_process(self, x):
        """x: (B, C, T) -> ((B, 2C, 256, frames padded to x64) fp32, n_samples, n_pad)
        (score_models.py:107-116), channels [Re c0.., Im c0..]."""
        self.stft.check()
        e, f = self._exp_factor()
        B, C, T = x.shape
        x = x.float()
        frames = _frames(T)
        Wp = (frames + 63) // 64 * 64
        feat = torch.empty(B, 256, Wp, 2 * C, device=x.device, dtype=torch.float32)
        ops.stft_compress(x[:, :C - 1].contiguous(), x[:, C - 1:].contiguous(), self.stft.window, feat, e, f, False)
        return feat.permute(0, 3, 1, 2).contiguous(), T, Wp - frames

    def post_process(self, x, n_samples, n_pad):
        """x: (B, 2n, 256, W) fp32 -> (B, n, n_samples) (score_models.py:118-124).  The frame
        count is derived from n_samples; n_pad is accepted for signature parity."""
        self.stft_inv.check()
        e, f = self._exp_factor()
        B, C2, F, Wp = x.shape
        spec = x.float().permute(0, 2, 3, 1).contiguous()
        out = torch.empty(B, C2 // 2, n_samples, device=x.device, dtype=torch.float32)
        ops.expand_istft(spec, self.stft_inv.window, out, e, f)
        return out

    # ------------------------------------------------------------------ the hot path
    def note_constant_mixture(self, mix):
        """Called by the samplers of this package with the conditioning tensor they will pass, unchanged, to every score
        evaluation of the run (None at the end of the run).  Only a tensor announced here is ever served from the
        mixture cache: `_version` does not see writes made by this library's own kernels through raw pointers (e.g. the
        rescaled mixture of the Schroedinger-bridge sampler, rewritten every step), so identity alone would not do."""
        self._mix_const = mix
        if mix is None:
            self._mixc = None

    def _mixture_cache(self, mix, bw, consts):
        """(buffer, mode) for ops.stft_compress: the compressed STFT of the mixture channel is the same for every score
        evaluation of a sampler run (the reference re-transforms cat(xt, mix) each time, score_models.py:126-131), so it
        is kept while the SAME tensor object, unmodified (`_version`), comes back.  The cached tensor is referenced from
        here, so its storage cannot be handed to another tensor meanwhile.  DITSEP_STFT_MIX_CACHE=0 turns it off."""
        import os
        if os.environ.get("DITSEP_STFT_MIX_CACHE", "1") == "0" or not mix.is_cuda or mix is not getattr(self, "_mix_const", None):
            return None, 0
        c = getattr(self, "_mixc", None)
        if c is not None and c["ref"] is mix and c["ver"] == mix._version and c["bw"] == bw and c["consts"] == consts:
            return c["buf"], 2
        buf = c["buf"] if c is not None and c["bw"] == bw and c["buf"].device == mix.device else \
            torch.empty(bw[0], bw[1], 256, 2, device=mix.device, dtype=torch.float32)
        self._mixc = {"ref": mix, "ver": mix._version, "bw": bw, "consts": consts, "buf": buf}
        return buf, 1

    def forward(self, xt, time_cond, mix):
        """xt: (B, n, T), time_cond: (B,), mix: (B, 1, T) -> score (B, n, T)
        (score_models.py:126-138)."""
        self.stft.check()
        e, f = self._exp_factor()
        B, n, T = xt.shape
        Wp = (_frames(T) + 63) // 64 * 64
        plan = self.backbone.engine().plan(B, 256, Wp, xt.device)
        mixc = mix.float().contiguous()
        cache, mode = self._mixture_cache(mixc, (B, Wp), (e, f, not self.backbone.centered))
        ops.stft_compress(xt.float().contiguous(), mixc, self.stft.window, plan.feat, e, f, not self.backbone.centered,
                          mix_cache=cache, mix_mode=mode)
        plan.t.copy_(time_cond)
        plan.run()
        out = torch.empty(B, n, T, device=xt.device, dtype=torch.float32)
        ops.expand_istft(plan.out, self.stft_inv.window, out, e, f)
        return out


class LatentScoreModelNCSNpp(torch.nn.Module):
    """score_models.py:140-186: cat(xt, mix) on channels, pad the last axis to a multiple of
    ``max_latent_length``, backbone, unpad.  No STFT."""

    def __init__(self, num_sources, backbone_args, max_latent_length=16):
        super().__init__()
        backbone_args.update(num_channels_in=num_sources + 1, num_channels_out=num_sources)
        self.backbone = (_hydra_instantiate or instantiate)(
            {**backbone_args, "_target_": _TARGET_ALIASES.get(backbone_args["_target_"], backbone_args["_target_"])})
        self.max_latent_length = max_latent_length

    def pad(self, x):
        rem = x.shape[-1] % self.max_latent_length
        if rem == 0:
            return x, 0
        pad = self.max_latent_length - rem
        return torch.nn.functional.pad(x, (0, pad)), pad

    def unpad(self, x, pad):
        return x if pad == 0 else x[..., :-pad]

    def forward(self, xt, time_cond, mix):
        x = torch.cat((xt, mix), dim=1)
        x, n_pad = self.pad(x)
        x = self.backbone(x, time_cond)
        return self.unpad(x, n_pad)
