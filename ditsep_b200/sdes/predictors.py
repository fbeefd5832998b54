"""Predictors of the PC sampler (mirror of src/sdes/predictors.py).

``reverse_diffusion`` (:55-66) runs as ONE fused sm_100a kernel per update
(``csrc/pc_step.cu``): drift, diffusion, the score term, the noise draw and the state
update in a single coalesced pass.  ``none`` (:69-76) is the identity.
``euler_maruyama`` (:39-52) is not on the hot path and not fused yet.
"""
from __future__ import annotations

import abc

import torch

from .. import ops
from ..utils.registry import Registry
from .sdes import NoiseSource

PredictorRegistry = Registry("Predictor")


class Predictor(abc.ABC):
    """Abstract predictor (predictors.py:10-36).  ``noise`` is this repo's extension: the
    shared ``NoiseSource`` of the sampler (tape or Philox); default = fresh Philox source."""

    def __init__(self, sde, score_fn, probability_flow=False, noise: NoiseSource = None):
        super().__init__()
        self.sde = sde
        self.rsde = sde.reverse(score_fn)
        self.score_fn = score_fn
        self.probability_flow = probability_flow
        self.noise = noise

    def _noise(self):
        if self.noise is None:
            self.noise = NoiseSource()
        return self.noise

    @abc.abstractmethod
    def update_fn(self, x, t, *args):
        pass

    def debug_update_fn(self, x, t, *args):
        raise NotImplementedError(f"Debug update function not implemented for predictor {self}.")


@PredictorRegistry.register("reverse_diffusion")
class ReverseDiffusionPredictor(Predictor):
    def __init__(self, sde, score_fn, probability_flow=False, noise: NoiseSource = None):
        super().__init__(sde, score_fn, probability_flow=probability_flow, noise=noise)
        if probability_flow:
            raise NotImplementedError("probability_flow=True has no fused kernel (not on the PC sampler path)")

    def update_fn(self, x, t, *args, _inplace=False, _want_mean=True, **kwargs):
        """x_mean = x - (f - G^2 score), x = x_mean + G z (predictors.py:60-66 with
        sdes.py:94-108,165-173).  Returns (x, x_mean).  The private keywords let the sampler
        update its own state in place and skip the x_mean store when it is not consumed."""
        y = args[0] if args else None
        score = self.score_fn(x, t, *args)
        p = self.sde.fused_params(n_src=x.shape[1], y=y)
        x_out = x if _inplace else x.clone()
        x_mean = torch.empty_like(x_out) if _want_mean else None
        src = self._noise()
        z, seed, draw = src.next(x_out)
        ops.pc_predictor(p, x_out, score.contiguous(), t.contiguous(), None if y is None else y.contiguous(), noise=z,
                         seed=seed, draw=draw, x_mean=x_mean, lengths=src.lengths)
        return x_out, x_mean


@PredictorRegistry.register("euler_maruyama")
class EulerMaruyamaPredictor(ReverseDiffusionPredictor):
    """predictors.py:42-57: x_mean = x - (drift - g^2 score) dt, x = x_mean + g sqrt(dt) z over the continuous reverse
    SDE (sdes.py:131-160).  With the reference's `discretize` (drift dt, g sqrt(dt): sdes.py:94-108) and its dt, which is
    1/N for both (the `getattr(kwargs, "dt", ...)` on a dict, quirk Q2), this is term by term the reverse-diffusion
    update, so it runs the same fused kernel row; pinned against the reference's own euler_maruyama run
    (tests/golden/variants.npz, `em_*`)."""


@PredictorRegistry.register("none")
class NonePredictor(Predictor):
    """An empty predictor that does nothing (predictors.py:69-76)."""

    def __init__(self, *args, **kwargs):
        pass

    def update_fn(self, x, t, *args, **kwargs):
        return x, x
