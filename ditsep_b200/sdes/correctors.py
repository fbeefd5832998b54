"""Correctors of the PC sampler (mirror of src/sdes/correctors.py).

``ald2`` (:87-121, MixSDE) and ``ald`` (:58-84, scalar-std SDEs) run as ONE fused sm_100a
kernel per Langevin step (``csrc/pc_step.cu``); the snr-scaled step sizes and the noise
draw never leave registers.  ``none`` (:124-133) is the identity.  ``langevin`` (:35-55)
couples the batch through mean norms: two reductions + the same fused kernel with a device-resident step size.
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
        p = self.sde.fused_params(snr=self.snr, n_src=x.shape[1], y=args[0] if args else None)
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
        if not isinstance(sde, (sdes.MixSDE,)):  # PriorMixSDE is a MixSDE here (correctors.py:92 names both)
            raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")
        super().__init__(sde, score_fn, snr, n_steps, noise=noise)


@CorrectorRegistry.register(name="langevin")
class LangevinCorrector(Corrector):
    """correctors.py:35-55: ONE step size for the whole batch, eps = 2 (snr mean_b|z_b| / mean_b|score_b|)^2, then
    x_mean = x + eps score, x = x_mean + sqrt(2 eps) z.  The step depends on the norm of the noise draw, so the draw is
    materialised (tape, or the Philox stream written out), two device reductions form eps, and the update is one
    launch of the fused kernel reading eps from device memory (no host synchronisation).  The batch means couple the
    mixtures, so a ragged batch cannot be made equivalent to batch-1 runs: `lengths` is refused."""

    def update_fn(self, x, t, *args, _inplace=False, _want_mean=True, **kwargs):
        x_out = x if _inplace else x.clone()
        x_mean = x_out
        src = self._noise()
        if src.lengths is not None:
            raise NotImplementedError("'langevin' couples the batch through mean norms: no ragged batches")
        B = x.shape[0]
        for _ in range(self.n_steps):
            score = self.score_fn(x_out, t, *args).contiguous()
            z, seed, draw = src.next(x_out)
            if z is None:
                z = torch.empty_like(x_out)
                ops.philox_normal(z, seed, draw)
            gn = torch.linalg.vector_norm(score.reshape(B, -1), dim=-1).mean()
            zn = torch.linalg.vector_norm(z.reshape(B, -1), dim=-1).mean()
            step = ((self.snr * zn / gn) ** 2 * 2).reshape(1).float()
            x_mean = torch.empty_like(x_out) if _want_mean else None
            ops.pc_langevin(x.shape[1], x_out, score, z.contiguous(), step, x_mean=x_mean)
        return x_out, x_mean


@CorrectorRegistry.register(name="none")
class NoneCorrector(Corrector):
    """An empty corrector that does nothing (correctors.py:124-133)."""

    def __init__(self, *args, **kwargs):
        self.snr = 0
        self.n_steps = 0

    def update_fn(self, x, t, *args, **kwargs):
        return x, x
