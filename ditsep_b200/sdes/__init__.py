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
from .sdes import SDE, MixSDE, NoiseSource, OUVESDE, PriorMixSDE, SBVESDE, SDERegistry

__all__ = ["PredictorRegistry", "CorrectorRegistry", "SDERegistry", "Predictor", "Corrector", "SDE", "MixSDE",
           "PriorMixSDE", "OUVESDE", "SBVESDE", "NoiseSource", "ReverseDiffusionPredictor", "get_pc_sampler", "get_pc_scheduled_sampler",
           "get_ode_sampler", "get_sb_sampler"]


def _note_constant_mixture(score_fn, y):
    """Tell a score model of this package that `y` conditions every evaluation of the run unchanged (it may then keep
    y's transform, models/diffsep/score_models.py::note_constant_mixture); any other callable is left alone."""
    hook = getattr(score_fn, "note_constant_mixture", None)
    if hook is not None:
        hook(y)


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
        _note_constant_mixture(score_fn, y)
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
            _note_constant_mixture(score_fn, None)
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
        _note_constant_mixture(score_fn, y)
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
            _note_constant_mixture(score_fn, None)
            if intermediate:
                return x_result, ns, im
            return x_result, ns

    return pc_sampler


def get_ode_sampler(*args, **kwargs):
    raise NotImplementedError("PF-ODE sampler (sdes/__init__.py:196-281) is SURVEY.md 8f item 2, not built yet")


def get_sb_sampler(sde, model, y: torch.Tensor, eps=1e-4, n_steps=50, sampler_type="ode", pad_dim=None, noise_tape=None,
                   seed=None, lengths=None, **kwargs):
    """Schroedinger-bridge samplers (sdes/__init__.py:284-389): first-order "sde" (:288-333) or "ode" (:335-382)
    discretisation on ``linspace(T, eps, N + 1)``, state initialised with the mixture repeated twice (the reference
    hard-codes two sources), returning ``(x, n_steps)`` -- the ARGUMENT n_steps (default 50), like the reference.

    Each step is one network evaluation plus ONE ``ops.sb_step`` launch.  If ``model`` exposes ``edm_table(times)``
    and ``score_model`` (``DiffSepOU`` with an ``SBVESDE``), the EDM preconditioning of ``DiffSepOU.forward``
    (diffsep.py:813-828) is folded into that launch as well: it also emits the c_in-scaled inputs of the next
    evaluation, so the raw network is called directly.  Any other callable ``model(xt, time, y)`` is used as is.
    ``noise_tape`` / ``seed`` / ``lengths`` as in ``get_pc_sampler`` ("sde" draws one tensor per step: draw i at
    step i)."""
    from .. import ops
    if sampler_type not in ("sde", "ode"):
        raise ValueError("Invalid type. Choose 'ode' or 'sde'.")
    if lengths is not None:
        lengths = torch.as_tensor(lengths, dtype=torch.int32).to(y.device).contiguous()
    noise = NoiseSource(tape=noise_tape, seed=seed, lengths=lengths)
    use_noise = sampler_type == "sde"

    def sampler():
        noise.begin_run()
        with torch.no_grad():
            B, N = y.shape[0], sde.N
            yc = y.to(torch.float32).contiguous()
            time_steps = torch.linspace(sde.T, eps, N + 1)
            # coefficient rows [w_prev, w_est, w_y, w_z, c_skip, c_out, f_div, c_in_next]; row 0 is the start-up call
            # that only forms the first evaluation's scaled inputs, row i + 1 belongs to step i
            coef = torch.zeros(N + 1, 8)
            coef[0, 0] = 1.0
            coef[:, 5:] = 1.0
            coef[1:, :4] = sde.step_table(time_steps, sampler_type)
            edm = getattr(model, "edm_table", None)
            edm = edm(time_steps[1:]) if edm is not None else None  # (N, 4): c_in, c_skip, c_out, f_div at each evaluation
            if edm is not None:
                coef[1:, 4:7] = edm[:, 1:4]
                coef[:N, 7] = edm[:, 0]
            coef = coef[:, None, :].expand(N + 1, B, 8).contiguous().to(y.device)
            vec_ts = time_steps[1:, None].expand(N, B).contiguous().to(y.device)
            xt = yc.repeat(1, 2, 1)
            if edm is not None:
                x_in, y_in = torch.empty_like(xt), torch.empty_like(yc)
                ops.sb_step(xt, None, yc, coef[0], x_in=x_in, y_in=y_in, lengths=lengths)
            for i in range(N):
                z, sd, draw = noise.next(xt) if use_noise else (None, 0, 0)
                if edm is not None:
                    f = model.score_model(x_in, vec_ts[i], y_in).contiguous()
                    ops.sb_step(xt, f, yc, coef[i + 1], x_out=xt, x_in=x_in, y_in=y_in, noise=z, use_noise=use_noise,
                                seed=sd, draw=draw, lengths=lengths)
                else:
                    f = model(xt, vec_ts[i], y).to(torch.float32).contiguous()
                    ops.sb_step(xt, f, yc, coef[i + 1], x_out=xt, noise=z, use_noise=use_noise, seed=sd, draw=draw,
                                lengths=lengths)
            return xt, n_steps

    return sampler
