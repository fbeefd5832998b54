"""Oracle: mixing-SDE math, predictor / corrector updates and the PC sampler loop.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates

* ``MixSDE``                 -- src/sdes/sdes.py:182-351
* ``OUVESDE``                -- src/sdes/sdes.py:595-698
* ``SDE.discretize`` / ``RSDE.discretize``  -- src/sdes/sdes.py:94-108, 165-173
* ``ReverseDiffusionPredictor.update_fn``   -- src/sdes/predictors.py:60-66
* ``AnnealedLangevinDynamics2.update_fn``   -- src/sdes/correctors.py:102-121
* ``AnnealedLangevinDynamics.update_fn``    -- src/sdes/correctors.py:70-84
* ``get_pc_sampler``         -- src/sdes/__init__.py:133-193
* ``normalize_batch``        -- src/utils/separate.py:56-67
* ``sample_prior``, ``compute_score_loss``, ``compute_score_loss_init_hack_pit`` (forward values)
                             -- src/diffsep.py:113-181, 305-358
* ``SBVESDE``, ``get_sb_sampler``, ``DiffSepOU.forward`` (EDM preconditioning)
                             -- src/sdes/sdes.py:701-779, src/sdes/__init__.py:284-389, src/diffsep.py:813-828,1165-1202

with the matrix algebra written out the way the reference does it (2x2 / nxn
``std @ x`` products), NOT in the closed form the CUDA kernel uses, so that the
closed form is checked against an independent formulation.  Noise comes from a
"tape": a callable ``noise(i, like)`` returning the i-th draw, so the oracle,
the reference and the CUDA path consume identical noise.
"""
from __future__ import annotations

import math
from typing import Callable, List, Optional, Tuple

import numpy as np
import torch

Tensor = torch.Tensor
NoiseFn = Callable[[int, Tensor], Tensor]


class NoiseTape:
    """Deterministic noise source: draw i is ``randn(shape, Generator(seed + i))`` on
    CPU (fp32), so any draw can be regenerated independently and in any order."""

    def __init__(self, seed: int = 4321):
        self.seed = seed
        self.count = 0

    def draw(self, i: int, shape) -> Tensor:
        g = torch.Generator().manual_seed(self.seed + 1000003 * i)
        return torch.empty(tuple(shape), dtype=torch.float32).normal_(generator=g)

    def __call__(self, i: int, like: Tensor) -> Tensor:
        self.count = max(self.count, i + 1)
        return self.draw(i, like.shape).to(like.device)


# --------------------------------------------------------------------------- MixSDE
class MixSDERef:
    """sdes.py:182-351."""

    def __init__(self, ndim=2, d_lambda=2.0, sigma_min=0.05, sigma_max=0.5, N=30):
        self.ndim, self.d_lambda = ndim, d_lambda
        self.sigma_min, self.sigma_max = sigma_min, sigma_max
        self.ratiosig = sigma_max / sigma_min
        self.logsig = math.log(self.ratiosig)
        self.N = N
        ones = torch.ones((ndim, 1))
        self.A = (ones @ ones.T / ndim)[None]          # :244-250 averaging matrix
        self.Pn = (torch.eye(ndim) - self.A[0])[None]  # projector on its null space
        self.T = 1.0

    def sde(self, x, t):
        """:277-286."""
        drift = -self.d_lambda * self.Pn @ x
        sigma = self.sigma_min * self.ratiosig ** t
        return drift, sigma * np.sqrt(2 * self.logsig)

    def cov_eigval(self, t):
        """:298-311."""
        mult = self.sigma_min ** 2
        srp = self.ratiosig ** (2 * t)
        ev1 = mult * (srp - 1)
        ev2 = mult * (srp - torch.exp(-2.0 * self.d_lambda * t)) / (1.0 + self.d_lambda / self.logsig)
        return ev1, ev2

    def std(self, t):
        """:317-322."""
        ev1, ev2 = self.cov_eigval(t)
        return ev1[:, None, None].sqrt() * self.A + ev2[:, None, None].sqrt() * self.Pn

    def mean(self, x0, t):
        """:288-296."""
        decay = torch.exp(-t[:, None, None] * self.d_lambda)
        return (self.A + decay * self.Pn) @ x0

    def prior_sampling(self, y, noise: Tensor):
        """:336-348 (hard-codes 2 sources, mean 0.5*y)."""
        t = torch.ones((y.shape[0],)) * self.T
        mean = torch.broadcast_to(0.5 * y, (y.shape[0], 2, y.shape[2]))
        return mean + self.std(t) @ noise


class PriorMixSDERef(MixSDERef):
    """sdes.py:354-592: MixSDE whose diffusion is scaled, per mixture and per sample, by an energy envelope of the
    mixture: sigma_mix = 0.5 sqrt(max(box_avg_len(mix^2), 1e-4)) (:480-492: avg_pool1d, stride 1, padding avg_len // 2,
    zeros counted; the extra output sample of an even window is dropped)."""

    def __init__(self, ndim=2, d_lambda=2.0, sigma_min=0.05, sigma_max=0.5, N=30, avg_len=510):
        super().__init__(ndim, d_lambda, sigma_min, sigma_max, N)
        self.avg_len = avg_len

    def sigma_mix(self, mix):
        k = self.avg_len
        m = torch.nn.functional.avg_pool1d(mix ** 2, kernel_size=k, stride=1, padding=k // 2)
        m = m.clamp(min=1e-4).sqrt()
        if k % 2 == 0:
            m = m[..., :-1]
        return 0.5 * m

    def sde_env(self, x, t, mix):
        """:453-472: drift as MixSDE, diffusion (B, ndim, L) = g(t) * sigma_mix."""
        drift, g = self.sde(x, t)
        return drift, _pad(g, x) * torch.broadcast_to(self.sigma_mix(mix), x.shape)

    def std_env(self, t, mix):
        """:519-531: L(t)[b] (ndim x ndim) times sigma_mix[b, 0, l] -> (B, ndim, ndim, L)."""
        return self.std(t)[..., None] * self.sigma_mix(mix)[:, None, :, :]

    def prior_sampling(self, y, noise: Tensor):
        """:566-589 with a one-channel mixture: 0.5 y + L(T) sigma_mix z."""
        t = torch.ones((y.shape[0],)) * self.T
        mean = torch.broadcast_to(0.5 * y, (y.shape[0], self.ndim, y.shape[2]))
        return mean + torch.einsum("bcdt,bdt->bct", self.std_env(t, y), noise)


class OUVESDERef:
    """sdes.py:595-698."""

    def __init__(self, theta=1.5, sigma_min=0.05, sigma_max=0.5, N=30):
        self.theta, self.sigma_min, self.sigma_max = theta, sigma_min, sigma_max
        self.logsig = np.log(sigma_max / sigma_min)
        self.N = N
        self.T = 1

    def sde(self, x, t, y):
        """:652-660."""
        drift = self.theta * (y - x)
        sigma = self.sigma_min * (self.sigma_max / self.sigma_min) ** t
        return drift, sigma * np.sqrt(2 * self.logsig)

    def std(self, t):
        """:670-682."""
        s, th, ls = self.sigma_min, self.theta, self.logsig
        return torch.sqrt(s ** 2 * torch.exp(-2 * th * t) * (torch.exp(2 * (th + ls) * t) - 1) * ls / (th + ls))

    def mean(self, x0, t, y):
        """:662-668: e^{-theta t} x0 + (1 - e^{-theta t}) y."""
        e = _pad(torch.exp(-self.theta * t), x0)
        return e * x0 + (1 - e) * y

    def prior_sampling(self, y, noise: Tensor):
        """:687-695: y + randn(shape) * std(1)."""
        std = self.std(torch.ones((y.shape[0],)))
        pad = (...,) + (None,) * (y.ndim - std.ndim)
        return y + noise * std[pad]


# --------------------------------------------------------------------------- update rules
def _pad(v: Tensor, like: Tensor) -> Tensor:
    return v[(...,) + (None,) * (like.ndim - v.ndim)]


def corrector_ald2_ref(sde: MixSDERef, x, t, score, noise, snr):
    """correctors.py:102-121 for one step: grad = L @ (L @ score);
    x_mean = x + 2 snr^2 grad; x = x_mean + (2 snr L) @ noise."""
    L = sde.std(t)
    grad = L @ (L @ score)
    x_mean = x + 2 * snr ** 2 * grad
    return x_mean + (2 * snr * L) @ noise, x_mean


def corrector_ald2_env_ref(sde: "PriorMixSDERef", x, t, y, score, noise, snr):
    """correctors.py:102-121 with PriorMixSDE's 4-D L (mult_std = einsum 'bcdt,bdt->bct', sdes.py:532-534)."""
    L = sde.std_env(t, y)
    mul = lambda a, v: torch.einsum("bcdt,bdt->bct", a, v)
    grad = mul(L, mul(L, score))
    x_mean = x + 2 * snr ** 2 * grad
    return x_mean + mul(2 * snr * L, noise), x_mean


def corrector_langevin_ref(x, score, noise, snr):
    """correctors.py:35-55 for one step: ONE step size for the whole batch from the batch means of the per-mixture
    norms, eps = 2 (snr mean|z| / mean|score|)^2; x_mean = x + eps score; x = x_mean + sqrt(2 eps) z."""
    gn = torch.norm(score.reshape(score.shape[0], -1), dim=-1).mean()
    zn = torch.norm(noise.reshape(noise.shape[0], -1), dim=-1).mean()
    step = (snr * zn / gn) ** 2 * 2
    x_mean = x + step * score
    return x_mean + noise * torch.sqrt(step * 2), x_mean


def corrector_ald_ref(sde, x, t, score, noise, snr):
    """correctors.py:70-84 for one step with a scalar std (OUVESDE)."""
    std = sde.std(t)
    step = (snr * std) ** 2 * 2
    x_mean = x + _pad(step, x) * score
    return x_mean + noise * _pad(torch.sqrt(step * 2), x), x_mean


def predictor_rd_ref(sde, x, t, y, score, noise):
    """predictors.py:60-66 + sdes.py:94-108,165-173.  dt is always 1/N (the
    ``getattr(kwargs, "dt", ...)`` on a dict never finds the key -- quirk Q2)."""
    dt = 1 / sde.N
    if isinstance(sde, PriorMixSDERef):
        drift, diffusion = sde.sde_env(x, t, y)
    else:
        drift, diffusion = sde.sde(x, t) if isinstance(sde, MixSDERef) else sde.sde(x, t, y)
    f = drift * dt
    G = diffusion * torch.sqrt(torch.tensor(dt))
    rev_f = f - _pad(G, f) ** 2 * score
    x_mean = x - rev_f
    return x_mean + _pad(G, x) * noise, x_mean


def predictor_em_ref(sde, x, t, y, score, noise):
    """predictors.py:42-57 (euler_maruyama) over the reverse SDE of sdes.py:131-160: x_mean = x - (drift - g^2 score) dt,
    x = x_mean + g sqrt(dt) z, dt = 1/N (the same dict-getattr quirk).  Algebraically the reverse-diffusion update."""
    dt = 1.0 / sde.N
    if isinstance(sde, PriorMixSDERef):
        drift, g = sde.sde_env(x, t, y)
    else:
        drift, g = sde.sde(x, t) if isinstance(sde, MixSDERef) else sde.sde(x, t, y)
    total = drift - _pad(g, x) ** 2 * score
    x_mean = x + total * (-dt)
    return x_mean + _pad(g, x) * np.sqrt(dt) * noise, x_mean


def schedule_timesteps_ref(schedule: Optional[str], T: float, eps: float, N: int) -> Tensor:
    """Time grids: ``linspace(T, eps, N)`` of get_pc_sampler (src/sdes/__init__.py:175), or the N + 1 point grids of
    get_pc_scheduled_sampler (:96-117)."""
    import math
    if schedule is None:
        return torch.linspace(T, eps, N)
    if schedule == "linear":
        return torch.linspace(T, eps, N + 1)
    if schedule == "log":
        return torch.logspace(math.log(T) / math.log(10), math.log(eps) / math.log(10), N + 1, base=10)
    if schedule == "revlog":
        return torch.logspace(math.log(eps) / math.log(10), math.log(T) / math.log(10), N + 1, base=10).flip(dims=(0,))
    raise NotImplementedError(f"Schedule '{schedule}' does not exist")


def pc_sampler_ref(score_fn, sde, y: Tensor, noise: NoiseFn, corrector: str = "ald2",
                   eps: float = 3e-2, snr: float = 0.5, corrector_steps: int = 1, denoise: bool = True,
                   n_spkrs: int = 2, on_score: Optional[Callable] = None,
                   schedule: Optional[str] = None, predictor: str = "reverse_diffusion") -> Tuple[Tensor, int]:
    """``get_pc_sampler`` closure (src/sdes/__init__.py:168-191) with the
    ``reverse_diffusion`` predictor.  Noise draw order (quirk Q3): #0 prior, then per
    step i: corrector draws 1+2i.., predictor draw; 1 + N*(corrector_steps+1) draws.
    ``on_score(k, x, t, score)`` is called after each of the N*(corrector_steps+1)
    score evaluations.
    ``schedule`` ("linear" | "log" | "revlog"): the time grid of ``get_pc_scheduled_sampler``
    (src/sdes/__init__.py:49-130) instead: N + 1 points of which the first N are visited; the step size the
    reference computes from the grid never reaches ``discretize`` (quirk Q2), so dt stays 1/N."""
    with torch.no_grad():
        shape = (y.shape[0], n_spkrs, *y.shape[2:])
        draw = 0
        z = noise(draw, torch.empty(shape)); draw += 1
        xt = sde.prior_sampling(y, z)
        timesteps = schedule_timesteps_ref(schedule, sde.T, eps, sde.N)
        k = 0
        xt_mean = xt
        for i in range(sde.N):
            vec_t = torch.ones(y.shape[0]) * timesteps[i]
            for _ in range(corrector_steps):
                s = score_fn(xt, vec_t, y)
                if on_score:
                    on_score(k, xt, vec_t, s)
                k += 1
                z = noise(draw, xt); draw += 1
                if corrector == "ald2" and isinstance(sde, PriorMixSDERef):
                    xt, xt_mean = corrector_ald2_env_ref(sde, xt, vec_t, y, s, z, snr)
                elif corrector == "ald2":
                    xt, xt_mean = corrector_ald2_ref(sde, xt, vec_t, s, z, snr)
                elif corrector == "langevin":
                    xt, xt_mean = corrector_langevin_ref(xt, s, z, snr)
                elif corrector == "ald":
                    xt, xt_mean = corrector_ald_ref(sde, xt, vec_t, s, z, snr)
                else:
                    raise ValueError(corrector)
            s = score_fn(xt, vec_t, y)
            if on_score:
                on_score(k, xt, vec_t, s)
            k += 1
            z = noise(draw, xt); draw += 1
            step = predictor_em_ref if predictor == "euler_maruyama" else predictor_rd_ref
            xt, xt_mean = step(sde, xt, vec_t, y, s, z)
        return (xt_mean if denoise else xt), sde.N * (corrector_steps + 1)


# --------------------------------------------------------------------------- Schroedinger bridge
class SBVESDERef:
    """``SBVESDE`` (src/sdes/sdes.py:701-779): variance-exploding Schroedinger bridge, g(t) = sqrt(c) k^t, alpha = 1."""

    def __init__(self, k: float = 2.6, c: float = 0.4, N: int = 50, eps: float = 1e-8):
        self.k, self.c, self.N, self.eps = k, c, N, eps
        self.T = 1

    def sigmas_alphas(self, t: Tensor):
        """sdes.py:741-752; everything in fp32 tensors like the reference."""
        alpha_t = torch.ones_like(t)
        alpha_T = torch.ones_like(t)
        log_k = torch.log(torch.tensor(self.k))
        sigma_t = torch.sqrt((self.c * (self.k ** (2 * t) - 1.0)) / (2 * log_k))
        sigma_T = torch.sqrt((self.c * (self.k ** (2 * self.T) - 1.0)) / (2 * log_k))
        alpha_bart = alpha_t / (alpha_T + self.eps)
        sigma_bart = torch.sqrt(sigma_T ** 2 - sigma_t ** 2 + self.eps)
        return sigma_t, sigma_T, sigma_bart, alpha_t, alpha_T, alpha_bart

    def std(self, t: Tensor) -> Tensor:
        """sdes.py:763-767."""
        sigma_t, sigma_T, sigma_bart, alpha_t, _, _ = self.sigmas_alphas(t)
        return (alpha_t * sigma_bart * sigma_t) / (sigma_T + self.eps)

    def mean(self, x0: Tensor, y: Tensor, t: Tensor) -> Tensor:
        """sdes.py:754-761."""
        sigma_t, sigma_T, sigma_bart, alpha_t, _, alpha_bart = self.sigmas_alphas(t)
        w_x = alpha_t * sigma_bart ** 2 / (sigma_T ** 2 + self.eps)
        w_y = alpha_bart * sigma_t ** 2 / (sigma_T ** 2 + self.eps)
        return _pad(w_x, x0) * x0 + _pad(w_y, x0) * y

    def marginal_prob(self, x0, y, t):
        return self.mean(x0, y, t), self.std(t)


def edm_forward_ref(net, sde: SBVESDERef, xt: Tensor, time: Tensor, mix: Tensor, sigma_data: float = 0.1,
                    network_scaling: str = "1/sigma") -> Tensor:
    """``DiffSepOU.forward`` with an ``SBVESDE`` (src/diffsep.py:813-828) and the coefficient branches it actually
    takes (``pad_dim`` is always a non-empty tuple there, :1165-1202): c_in = sd^2 / (s^2 + sd^2) (NOT the
    1/sqrt form of the other branch), c_out = s sd / sqrt(sd^2 + s^2), c_skip = s^2 / (s^2 + sd^2)."""
    s = sde.std(time)
    c_in = _pad(sigma_data ** 2 / (s ** 2 + sigma_data ** 2), xt)
    c_out = _pad((s * sigma_data) / torch.sqrt(sigma_data ** 2 + s ** 2), xt)
    c_skip = _pad(s ** 2 / (s ** 2 + sigma_data ** 2), xt)
    F = net(c_in * xt, time, c_in * mix)
    if network_scaling == "1/sigma":
        F = F / _pad(s, xt)
    elif network_scaling == "1/t":
        F = F / _pad(time, xt)
    return c_skip * xt + c_out * F


def sb_sampler_ref(sde: SBVESDERef, model, y: Tensor, noise: Optional[NoiseFn] = None, eps: float = 1e-4,
                   n_steps: int = 50, sampler_type: str = "ode") -> Tuple[Tensor, int]:
    """``get_sb_sampler`` closures (src/sdes/__init__.py:284-389): first-order SB-SDE / SB-ODE discretisation on
    ``linspace(T, eps, N + 1)``, state initialised with the mixture repeated twice; the returned count is the
    ``n_steps`` ARGUMENT (default 50), not the number of network evaluations (N) -- reproduced.  The SDE variant
    draws one noise tensor per step (draw i at step i), unused at the last step where its weight is 0."""
    with torch.no_grad():
        xt = y.repeat(1, 2, 1)
        ts = torch.linspace(sde.T, eps, sde.N + 1)
        B = xt.shape[0]
        sig_p, sig_T, sigb_p, al_p, al_T, _ = sde.sigmas_alphas(ts[0] * torch.ones(B))
        for i, t in enumerate(ts[1:]):
            time = t * torch.ones(B)
            sig, sig_T, sigb, al, al_T, _ = sde.sigmas_alphas(time)
            est = model(xt, time, y)
            if sampler_type == "sde":
                w_prev = al * sig ** 2 / (al_p * sig_p ** 2 + sde.eps)
                tmp = 1 - sig ** 2 / (sig_p ** 2 + sde.eps)
                w_est = al * tmp
                w_z = al * sig * torch.sqrt(tmp)
                z = noise(i, xt)
                if i == sde.N - 1:
                    xt = _pad(w_prev, xt) * xt + _pad(w_est, xt) * est + 0.0 * z
                else:
                    xt = _pad(w_prev, xt) * xt + _pad(w_est, xt) * est + _pad(w_z, xt) * z
            elif sampler_type == "ode":
                w_prev = al * sig * sigb / (al_p * sig_p * sigb_p + sde.eps)
                w_est = al / (sig_T ** 2 + sde.eps) * (sigb ** 2 - sigb_p * sig * sigb / (sig_p + sde.eps))
                w_mix = al / (al_T * sig_T ** 2 + sde.eps) * (sig ** 2 - sig_p * sig * sigb / (sigb_p + sde.eps))
                xt = _pad(w_prev, xt) * xt + _pad(w_est, xt) * est + _pad(w_mix, xt) * y
            else:
                raise ValueError("Invalid type. Choose 'ode' or 'sde'.")
            sig_p, sigb_p, al_p = sig, sigb, al
        return xt, n_steps


# --------------------------------------------------------------------------- score-matching loss, forward
def sample_prior_ref(sde: MixSDERef, mix: Tensor, target: Tensor, time: Tensor, z: Tensor, init_hack=0,
                     t_rev_init: float = 0.03, select4: Optional[Tensor] = None):
    """``DiffSepModel.sample_prior`` (src/diffsep.py:113-181) with the time draw, the normal draw and (init_hack 4) the
    Bernoulli selection passed in: returns (x_t, time, L, z)."""
    true_mix = torch.broadcast_to(mix, target.shape) / target.shape[1]
    if init_hack == 4:
        time = torch.where(select4, torch.ones_like(time) * sde.T, time)
    mean, L = sde.mean(target, time), sde.std(time)
    if init_hack in (1, 4):
        keep = (time < sde.T - t_rev_init) if init_hack == 1 else ~select4
        keep = torch.broadcast_to(keep[:, None, None], z.shape)
        z = torch.where(keep, z, z + torch.linalg.solve(L, true_mix - mean))
        return mean + L @ z, time, L, z
    if init_hack in (2, 3):
        Tm = sde.T - t_rev_init
        beta = torch.clamp((time - Tm) / (sde.T - Tm), min=0.0, max=1.0)[:, None, None]
        x_t = true_mix * beta + mean * (1.0 - beta) + L @ z
        if init_hack == 3:
            z = torch.linalg.solve(L, x_t - mean)
        return x_t, time, L, z
    return mean + L @ z, time, L, z


def score_loss_ref(model, sde: MixSDERef, mix: Tensor, target: Tensor, time: Tensor, z: Tensor, **hack) -> Tensor:
    """``compute_score_loss`` (src/diffsep.py:344-358) with MSELoss(reduction="none"): (B,) per-item means."""
    x_t, time, L, z = sample_prior_ref(sde, mix, target, time, z, **hack)
    pred = model(x_t, time, mix)
    return ((L @ pred) - (-z)).square().mean(dim=(-2, -1))


def score_loss_init_hack_pit_ref(model, sde: MixSDERef, mix: Tensor, target: Tensor, z0: Tensor) -> Tensor:
    """``compute_score_loss_init_hack_pit`` (src/diffsep.py:305-338)."""
    import itertools
    time = torch.ones(mix.shape[0]) * sde.T
    true_mix = torch.broadcast_to(mix, target.shape) / target.shape[1]
    losses = []
    for perm in itertools.permutations(range(target.shape[1])):
        mean, L = sde.mean(target[:, perm, :], time), sde.std(time)
        z = z0 + torch.linalg.solve(L, true_mix - mean)
        pred = model(mean + L @ z, time, mix)
        losses.append(((L @ pred) + z).square().mean(dim=(-2, -1)))
    return torch.stack(losses, dim=1).min(dim=1).values


def score_loss_with_pit_ref(model, sde: MixSDERef, mix: Tensor, target: Tensor, time: Tensor, z: Tensor, select: Tensor,
                            mmnr_thresh_pit: float = -10.0) -> Tensor:
    """``compute_score_loss_with_pit`` (src/diffsep.py:183-258) with the time draw, the normal draw and the random
    order selection (``select``: (B,) indices into the permutation list) passed in.  Written per item, NOT with the
    reference's batched broadcast / flatten bookkeeping; PIT items first, like the reference's ``torch.cat``."""
    import itertools
    perms = list(itertools.permutations(range(target.shape[1])))
    L = sde.std(time)
    means = torch.stack([sde.mean(target[:, p, :], time) for p in perms], dim=1)
    Lz = L @ z
    pit, reg = [], []
    for b in range(target.shape[0]):
        m_sel = means[b, int(select[b])]
        x_t = m_sel + Lz[b]
        err = means[b] - m_sel
        err_pow = err.square().sum() / ((len(perms) - 1) * err.shape[1] * err.shape[2])
        mmnr = 10.0 * torch.log10(err_pow / Lz[b].square().mean().clamp(min=1e-5))
        pred = model(x_t[None], time[b:b + 1], mix[b:b + 1])[0]
        if mmnr < mmnr_thresh_pit:
            pit.append(torch.stack([((L[b] @ pred) + z[b] + torch.linalg.solve(L[b], e)).square().mean() for e in err]).min())
        else:
            reg.append(((L[b] @ pred) + z[b]).square().mean())
    return torch.stack(pit + reg)


def score_loss_with_pit_allthetime_ref(model, sde: MixSDERef, mix: Tensor, target: Tensor, time: Tensor, z0: Tensor) -> Tensor:
    """``compute_score_loss_with_pit_allthetime`` (src/diffsep.py:260-303) without the source shuffle."""
    import itertools
    mean_0, L = sde.mean(target, time), sde.std(time)
    pred = model(mean_0 + L @ z0, time, mix)
    losses = []
    for p in itertools.permutations(range(target.shape[1])):
        z_p = z0 + torch.linalg.solve(L, mean_0 - sde.mean(target[:, p, :], time))
        losses.append(((L @ pred) + z_p).square().mean(dim=(-2, -1)))
    return torch.stack(losses, dim=0).min(dim=0).values


def ou_sample_prior_ref(sde, mix: Tensor, target: Tensor, time: Tensor, z: Tensor):
    """``DiffSepOU.sample_prior`` (src/diffsep.py:800-811) for an ``OUVESDERef`` or ``SBVESDERef``: (x_t, sigma)."""
    if isinstance(sde, SBVESDERef):
        mean, std = sde.marginal_prob(target, mix, time)
    else:
        mean, std = sde.mean(target, time, mix), sde.std(time)
    sigma = _pad(std, mean)
    return mean + sigma * z, sigma


def ou_score_loss_ref(model, sde, mix: Tensor, target: Tensor, time: Tensor, z: Tensor) -> Tensor:
    """``DiffSepOU.compute_score_loss`` (src/diffsep.py:836-841) with MSELoss(reduction="none"): (B, n_src)."""
    x_t, sigma = ou_sample_prior_ref(sde, mix, target, time, z)
    loss = (model(x_t, time, mix) * sigma + z).square()
    return loss.mean(dim=tuple(range(2 - loss.ndim, 0)))


def ou_score_loss_init_hack_pit_ref(model, sde: OUVESDERef, mix: Tensor, target: Tensor, z0: Tensor) -> Tensor:
    """``DiffSepOU.compute_score_loss_init_hack_pit`` (src/diffsep.py:843-869), OU branch."""
    import itertools
    time = torch.ones(mix.shape[0]) * sde.T
    losses = []
    for perm in itertools.permutations(range(target.shape[1])):
        mean, sigma = sde.mean(target[:, perm, ...], time, mix), _pad(sde.std(time), mix)
        z = z0 + (mix - mean) / sigma
        loss = (model(mix + sigma * z0, time, mix) * sigma + z).square()
        losses.append(loss.mean(dim=tuple(range(2 - loss.ndim, 0))))
    return torch.stack(losses, dim=1).min(dim=1).values


def normalize_batch_ref(mix: Tensor):
    """src/utils/separate.py:56-63: per-item mean / unbiased std over (chan, time), clamp 1e-5."""
    mean = mix.mean(dim=(1, 2), keepdim=True)
    std = mix.std(dim=(1, 2), keepdim=True).clamp(min=1e-5)
    return (mix - mean) / std, mean, std


# --------------------------------------------------------------------------- metric
def si_sdr_pit_ref(ref: Tensor, est: Tensor) -> Tuple[Tensor, List[Tuple[int, ...]]]:
    """Scale-invariant SDR with the best source permutation.

    Restates ``fast_bss_eval.si_bss_eval_sources(ref, est, zero_mean=False,
    compute_permutation=True)`` as called at src/evaluate_mp.py:173-175 (third-party
    package fast-bss-eval==0.1.4, env/environment.yaml:294, not installed here):
    alpha = <est,ref>/<ref,ref>, SI-SDR = 10 log10(|alpha ref|^2 / |est - alpha ref|^2),
    permutation maximising the mean.  Parity unpinned against the package; only
    DIFFERENCES of two SI-SDRs computed by this same function are asserted.
    ref, est: (B, n_src, T) -> (B, n_src) dB, per-item permutation."""
    import itertools

    b, n, _ = ref.shape
    ref = ref.double()
    est = est.double()
    out = torch.empty(b, n, dtype=torch.float64)
    perms = []
    for bi in range(b):
        best, best_p, best_v = -float("inf"), None, None
        for p in itertools.permutations(range(n)):
            vals = []
            for r in range(n):
                e = est[bi, p[r]]
                rr = ref[bi, r]
                alpha = (e @ rr) / (rr @ rr).clamp(min=1e-30)
                tgt = alpha * rr
                vals.append(10 * torch.log10((tgt @ tgt) / ((e - tgt) @ (e - tgt)).clamp(min=1e-30)))
            v = torch.stack(vals)
            if float(v.mean()) > best:
                best, best_p, best_v = float(v.mean()), p, v
        out[bi] = best_v
        perms.append(best_p)
    return out.float(), perms
