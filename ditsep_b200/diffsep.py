"""``DiffSepModel``: the API surface of the separation sampler (mirror of src/diffsep.py).

Only the inference side of the reference Lightning module is on the hot path and is kept:
``__init__`` (diffsep.py:28-80: build score model + SDE from the Hydra config),
``forward`` (:341-343), ``separate`` (:82-98), ``get_pc_sampler`` (:623-695),
``normalize_batch`` / ``denormalize_batch``, ``eval(no_ema=...)``.  Training losses, EMA
and optimiser plumbing are out of scope.  pytorch_lightning,
hydra and omegaconf are optional: plain dicts / namespaces work as configs.
"""
from __future__ import annotations

import math
from typing import Any

import torch

from . import sdes, utils
from .models.diffsep.score_models import instantiate

try:  # pragma: no cover - not installed in the build image
    import pytorch_lightning as pl

    _Base = pl.LightningModule
except Exception:
    _Base = torch.nn.Module


class _Cfg:
    """Attribute + item access over nested dicts (the subset of DictConfig the path uses)."""

    def __init__(self, d):
        object.__setattr__(self, "_d", d)

    def __getattr__(self, k):
        try:
            v = self._d[k]
        except KeyError:
            raise AttributeError(k) from None
        return _Cfg(v) if isinstance(v, dict) else v

    def __getitem__(self, k):
        return getattr(self, k)

    def __contains__(self, k):
        return k in self._d

    def items(self):
        return self._d.items()

    def copy(self):
        return dict(self._d)

    def to_dict(self):
        return self._d


def _to_plain(cfg) -> Any:
    """DictConfig / _Cfg / dict -> plain nested dict."""
    if isinstance(cfg, _Cfg):
        return cfg.to_dict()
    try:  # pragma: no cover
        from omegaconf import OmegaConf

        if OmegaConf.is_config(cfg):
            return OmegaConf.to_container(cfg, resolve=True)
    except Exception:
        pass
    return cfg


class DiffSepModel(_Base):
    def __init__(self, config):
        super().__init__()
        self.config = _Cfg(config) if isinstance(config, dict) else config
        model_cfg = self.config.model
        self.score_model = instantiate(_to_plain(model_cfg.score_model), _recursive_=False)
        self.sde = instantiate(_to_plain(model_cfg.sde))
        self.t_eps = model_cfg.t_eps
        self.t_max = self.sde.T
        self.normalize_batch = utils.normalize_batch
        self.denormalize_batch = utils.denormalize_batch

    def forward(self, xt, time, mix):
        return self.score_model(xt, time, mix)

    def eval(self, no_ema=False):  # the reference swaps EMA weights in unless no_ema (diffsep.py:586-606)
        return super().train(False)

    def separate(self, mix, **kwargs):
        """diffsep.py:82-98, including quirk Q1: the sampler runs twice and the second,
        un-denormalised result tuple is what is returned."""
        (mix, _), *stats = self.normalize_batch((mix, None))
        sampler_kwargs = dict(_to_plain(self.config.model.sampler))
        sampler_kwargs.update(kwargs)
        sampler = self.get_pc_sampler("reverse_diffusion", "ald2", mix, **sampler_kwargs)
        est, *others = sampler()
        est = self.denormalize_batch(est, *stats)
        return sampler()

    def get_pc_sampler(self, predictor_name, corrector_name, y, N=None, minibatch=None, schedule=None, **kwargs):
        """diffsep.py:623-695."""
        N = self.sde.N if N is None else N
        sde = self.sde.copy()
        sde.N = N
        kwargs = {"eps": self.t_eps, **kwargs}
        factory = sdes.get_pc_sampler if schedule is None else sdes.get_pc_scheduled_sampler
        if schedule is not None:
            kwargs["schedule"] = schedule
        if minibatch is None:
            return factory(predictor_name, corrector_name, sde=sde, score_fn=self, y=y, **kwargs)
        M = y.shape[0]

        def batched_sampling_fn():
            samples, ns, intmet = [], [], []
            for i in range(int(math.ceil(M / minibatch))):
                y_mini = y[i * minibatch:(i + 1) * minibatch]
                kw_i = dict(kwargs)
                if kw_i.get("seed") is not None:  # an explicit seed must not give every minibatch the same stream
                    kw_i["seed"] = int(kw_i["seed"]) + 7919 * i
                if kw_i.get("lengths") is not None:
                    kw_i["lengths"] = kw_i["lengths"][i * minibatch:(i + 1) * minibatch]
                sampler = factory(predictor_name, corrector_name, sde=sde, score_fn=self, y=y_mini, **kw_i)
                sample, n, *other = sampler()
                samples.append(sample)
                ns.append(n)
                if len(other) > 0:
                    intmet.append(other[0])
            samples = torch.cat(samples, dim=0)
            if len(intmet) > 0:
                return samples, ns, intmet
            return samples, ns

        return batched_sampling_fn


def default_config(nf=64, n_src=2, N=30, precision=None, fs=8000):
    """src/config/diffsep/model/default.yaml:15-48 as a plain dict (the hot-path config)."""
    backbone = {"_target_": "models.diffsep.ncsnpp.NCSNpp", "nf": nf, "ch_mult": [1, 1, 2, 2, 2, 2, 2],
                "num_res_blocks": 2, "attn_resolutions": [16], "resamp_with_conv": True, "image_size": 256}
    if precision is not None:
        backbone["precision"] = precision
    return {"model": {
        "fs": fs, "n_speakers": n_src, "t_eps": 0.03, "t_rev_init": 0.03,
        "score_model": {"_target_": "models.diffsep.score_models.ScoreModelNCSNpp", "num_sources": n_src,
                        "stft_args": {"n_fft": 510, "hop_length": 128, "center": True, "pad_mode": "constant"},
                        "backbone_args": backbone, "transform": "exponent", "spec_abs_exponent": 0.5,
                        "spec_factor": 0.15, "spec_trans_learnable": False},
        "sde": {"_target_": "sdes.MixSDE", "ndim": n_src, "d_lambda": 2.0, "sigma_min": 0.05, "sigma_max": 0.5, "N": N},
        "sampler": {"N": N, "snr": 0.5, "corrector_steps": 1},
    }}
