"""Sampler factories (mirror of src/sdes/__init__.py).

``get_pc_sampler`` (:133-193) keeps the reference signature and return convention
``sampler() -> (x, nfe[, intermediates])``.  Two keyword extensions (absorbed by the
reference's ``**kwargs``, so existing call sites stay valid):

* ``noise_tape``: callable ``(i, like) -> tensor`` supplying draw i (parity runs);
* ``seed``: Philox seed of the in-kernel noise (default: taken from torch's generator).

The loop body is two fused-kernel launches plus two score-network evaluations per step;
the state ``xt`` is updated in place and ``xt_mean`` is only materialised when it is
consumed (last step with ``denoise=True``, or ``intermediate=True``).
"""
from __future__ import annotations

import torch

from .correctors import Corrector, CorrectorRegistry
from .predictors import Predictor, PredictorRegistry, ReverseDiffusionPredictor
from .sdes import SDE, MixSDE, NoiseSource, OUVESDE, PriorMixSDE, SDERegistry

__all__ = ["PredictorRegistry", "CorrectorRegistry", "SDERegistry", "Predictor", "Corrector", "SDE", "MixSDE",
           "PriorMixSDE", "OUVESDE", "NoiseSource", "ReverseDiffusionPredictor", "get_pc_sampler", "get_pc_scheduled_sampler",
           "get_ode_sampler", "get_sb_sampler"]


def get_pc_sampler(predictor_name, corrector_name, sde, score_fn, y, true_mean=None, denoise=True, eps=3e-2, snr=0.1,
                   corrector_steps=1, probability_flow: bool = False, intermediate=False, n_spkrs=2, noise_tape=None,
                   seed=None, lengths=None, **kwargs):
    """Create a Predictor-Corrector sampler (sdes/__init__.py:133-193).  ``lengths``: optional (B,) valid sample
    counts when ``y`` is a batch of zero-padded mixtures of different lengths (utils/shard.py)."""
    predictor_cls = PredictorRegistry.get_by_name(predictor_name)
    corrector_cls = CorrectorRegistry.get_by_name(corrector_name)
    if lengths is not None:
        lengths = torch.as_tensor(lengths, dtype=torch.int32).to(y.device).contiguous()
    noise = NoiseSource(tape=noise_tape, seed=seed, lengths=lengths)
    predictor = predictor_cls(sde, score_fn, probability_flow=probability_flow, noise=noise)
    corrector = corrector_cls(sde, score_fn, snr=snr, n_steps=corrector_steps, noise=noise)

    def pc_sampler():
        im = [] if intermediate else None
        noise.begin_run()
        with torch.no_grad():
            shape = torch.Size((y.shape[0], n_spkrs, *y.shape[2:]))
            cond = true_mean if true_mean is not None else y
            xt = sde.prior_sampling(shape, cond, noise=noise).to(cond.device)
            # all N time vectors in one go (the reference builds ones(B) * t every step)
            timesteps = torch.linspace(sde.T, eps, sde.N, device=y.device)
            vec_ts = timesteps[:, None].expand(sde.N, y.shape[0]).contiguous()
            xt_mean = xt
            for i in range(sde.N):
                vec_t = vec_ts[i]
                last = i == sde.N - 1
                xt, xt_mean = corrector.update_fn(xt, vec_t, y, _inplace=not intermediate, _want_mean=intermediate)
                if intermediate:
                    im.append((xt, xt_mean))
                xt, xt_mean = predictor.update_fn(xt, vec_t, y, _inplace=not intermediate,
                                                  _want_mean=intermediate or (last and denoise))
            x_result = xt_mean if denoise else xt
            ns = sde.N * (corrector.n_steps + 1)
            if intermediate:
                return x_result, ns, im
            return x_result, ns

    return pc_sampler


def _schedule_timesteps(schedule, T, eps, N, device):
    """The N + 1 point time grids of the scheduled sampler (sdes/__init__.py:96-117)."""
    import math
    base = 10
    if schedule == "linear":
        return torch.linspace(T, eps, N + 1, device=device)
    if schedule == "log":
        return torch.logspace(math.log(T) / math.log(base), math.log(eps) / math.log(base), N + 1, base=base, device=device)
    if schedule == "revlog":
        return torch.logspace(math.log(eps) / math.log(base), math.log(T) / math.log(base), N + 1, base=base,
                              device=device).flip(dims=(0,))
    raise NotImplementedError(f"Schedule '{schedule}' does not exist")


def get_pc_scheduled_sampler(predictor_name, corrector_name, sde, score_fn, y, denoise=True, true_mean=None, eps=3e-2,
                             snr=0.1, corrector_steps=1, probability_flow: bool = False, intermediate=False,
                             schedule="linear", noise_tape=None, seed=None, lengths=None, **kwargs):
    """PC sampler on a scheduled time grid (sdes/__init__.py:49-130): "linear", "log" or "revlog" spacing of N + 1
    points between T and eps, of which the first N are visited.  Same fused update kernels as get_pc_sampler.  The
    reference hands the grid spacing to update_fn as ``dt=``, but it never reaches ``SDE.discretize`` (quirk Q2:
    ``getattr`` on the kwargs dict), so the step size stays 1/N -- reproduced.  Like the reference, the prior is drawn
    with the CONDITIONING tensor's shape (y.shape, not (B, n_spkrs, ...))."""
    predictor_cls = PredictorRegistry.get_by_name(predictor_name)
    corrector_cls = CorrectorRegistry.get_by_name(corrector_name)
    if lengths is not None:
        lengths = torch.as_tensor(lengths, dtype=torch.int32).to(y.device).contiguous()
    noise = NoiseSource(tape=noise_tape, seed=seed, lengths=lengths)
    predictor = predictor_cls(sde, score_fn, probability_flow=probability_flow, noise=noise)
    corrector = corrector_cls(sde, score_fn, snr=snr, n_steps=corrector_steps, noise=noise)

    def pc_sampler():
        im = [] if intermediate else None
        noise.begin_run()
        with torch.no_grad():
            cond = true_mean if true_mean is not None else y
            xt = sde.prior_sampling(cond.shape, cond, noise=noise).to(cond.device)
            timesteps = _schedule_timesteps(schedule, sde.T, eps, sde.N, y.device)
            vec_ts = timesteps[:sde.N, None].expand(sde.N, y.shape[0]).contiguous()
            xt_mean = xt
            for i in range(sde.N):
                vec_t = vec_ts[i]
                last = i == sde.N - 1
                xt, xt_mean = corrector.update_fn(xt, vec_t, y, _inplace=not intermediate, _want_mean=intermediate)
                if intermediate:
                    im.append((xt, xt_mean))
                xt, xt_mean = predictor.update_fn(xt, vec_t, y, _inplace=not intermediate,
                                                  _want_mean=intermediate or (last and denoise))
            x_result = xt_mean if denoise else xt
            ns = sde.N * (corrector.n_steps + 1)
            if intermediate:
                return x_result, ns, im
            return x_result, ns

    return pc_sampler


def get_ode_sampler(*args, **kwargs):
    raise NotImplementedError("PF-ODE sampler (sdes/__init__.py:196-281) is SURVEY.md 8f item 2, not built yet")


def get_sb_sampler(*args, **kwargs):
    raise NotImplementedError("SB sampler (sdes/__init__.py:284-389) is SURVEY.md 8f item 2, not built yet")
