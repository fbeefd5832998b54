"""Correctors of the PC sampler (mirror of src/sdes/correctors.py).

``ald2`` (:87-121, MixSDE) and ``ald`` (:58-84, scalar-std SDEs) run as ONE fused sm_100a
kernel per Langevin step (``csrc/pc_step.cu``); the snr-scaled step sizes and the noise
draw never leave registers.  ``none`` (:124-133) is the identity.  ``langevin`` (:35-55)
couples the batch through mean norms and is not on the hot path.
"""
from __future__ import annotations

import abc

import torch

from .. import ops
from ..utils.registry import Registry
from . import sdes
from .sdes import NoiseSource

CorrectorRegistry = Registry("Corrector")


class Corrector(abc.ABC):
    """Abstract corrector (correctors.py:11-32); ``noise`` as in ``Predictor``."""

    def __init__(self, sde, score_fn, snr, n_steps, noise: NoiseSource = None):
        super().__init__()
        self.rsde = sde.reverse(score_fn)
        self.score_fn = score_fn
        self.snr = snr
        self.n_steps = n_steps
        self.noise = noise

    def _noise(self):
        if self.noise is None:
            self.noise = NoiseSource()
        return self.noise

    @abc.abstractmethod
    def update_fn(self, x, t, *args):
        pass


class _FusedLangevin(Corrector):
    def __init__(self, sde, score_fn, snr, n_steps, noise: NoiseSource = None):
        super().__init__(sde, score_fn, snr, n_steps, noise=noise)
        self.sde = sde

    def update_fn(self, x, t, *args, _inplace=False, _want_mean=True, **kwargs):
        """n_steps Langevin steps; returns (x, x_mean) (x_mean = x when n_steps == 0)."""
        x_out = x if _inplace else x.clone()
        x_mean = x_out
        p = self.sde.fused_params(snr=self.snr, n_src=x.shape[1])
        for _ in range(self.n_steps):
            score = self.score_fn(x_out, t, *args)
            x_mean = torch.empty_like(x_out) if _want_mean else None
            src = self._noise()
            z, seed, draw = src.next(x_out)
            ops.pc_corrector(p, x_out, score.contiguous(), t.contiguous(), noise=z, seed=seed, draw=draw, x_mean=x_mean,
                             lengths=src.lengths)
        return x_out, x_mean


@CorrectorRegistry.register(name="ald")
class AnnealedLangevinDynamics(_FusedLangevin):
    """eps = 2 (snr std(t))^2; x_mean = x + eps score; x = x_mean + sqrt(2 eps) z (correctors.py:70-84)."""

    def __init__(self, sde, score_fn, snr, n_steps, noise: NoiseSource = None):
        if isinstance(sde, (sdes.MixSDE,)):
            raise NotImplementedError("'ald' needs a scalar-std SDE (OUVESDE); use 'ald2' with MixSDE")
        super().__init__(sde, score_fn, snr, n_steps, noise=noise)


@CorrectorRegistry.register(name="ald2")
class AnnealedLangevinDynamics2(_FusedLangevin):
    """x_mean = x + 2 snr^2 L L score; x = x_mean + 2 snr L z (correctors.py:102-121)."""

    def __init__(self, sde, score_fn, snr, n_steps, noise: NoiseSource = None):
        if not isinstance(sde, (sdes.MixSDE,)):
            raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")
        super().__init__(sde, score_fn, snr, n_steps, noise=noise)


@CorrectorRegistry.register(name="langevin")
class LangevinCorrector(Corrector):
    def update_fn(self, x, t, *args, **kwargs):
        raise NotImplementedError("'langevin' is outside the fused PC-sampler hot path (SURVEY.md 8f item 2)")


@CorrectorRegistry.register(name="none")
class NoneCorrector(Corrector):
    """An empty corrector that does nothing (correctors.py:124-133)."""

    def __init__(self, *args, **kwargs):
        self.snr = 0
        self.n_steps = 0

    def update_fn(self, x, t, *args, **kwargs):
        return x, x
