"""Mixing / OU-VE SDE classes of the DiffSep sampler (mirror of src/sdes/sdes.py).

API parity with the reference: ``SDE`` (:44-179), ``MixSDE`` (:182-351), ``OUVESDE``
(:595-698), ``SDERegistry`` (:14) -- same constructor arguments, attributes, method
names, warnings and errors.  The small closed-form helpers (``sde``, ``marginal_prob``,
``_std`` ...) are plain tensor algebra on (B,) / (B,n,n) operands and stay in PyTorch;
the sampler's hot path (prior draw, corrector and predictor updates) does not go through
them: it calls the fused sm_100a kernel of ``csrc/pc_step.cu`` via ``fused_params()``.
"""
from __future__ import annotations

import abc
import math
import warnings

import numpy as np
import torch

from .. import ops
from .._lib import SdeParams
from ..utils.registry import Registry

SDERegistry = Registry("SDE")

SDE_MIX, SDE_OUVE = 0, 1


def batch_broadcast(a, x):
    """Broadcast a (B,) tensor over all non-batch dimensions of x (sdes.py:26-41)."""
    if len(a.shape) != 1:
        a = a.squeeze()
        if len(a.shape) != 1:
            raise ValueError(
                f"Don't know how to batch-broadcast tensor `a` with more than one effective dimension (shape {a.shape})")
    if a.shape[0] != x.shape[0] and a.shape[0] != 1:
        raise ValueError(
            f"Don't know how to batch-broadcast shape {a.shape} over {x.shape} as the batch dimension is not matching")
    return a.view((x.shape[0], *(1 for _ in range(len(x.shape) - 1))))


class NoiseSource:
    """Where the sampler's Gaussian draws come from.

    * ``tape`` given (callable ``tape(i, like) -> tensor``): draw i is read from the tape
      (parity runs: the reference, the oracle and this path consume identical noise).
    * otherwise: in-register Philox4x32-10 inside the fused kernel, keyed by (seed, i).
      ``seed=None`` takes 63 bits from torch's global CPU generator, so
      ``torch.manual_seed`` makes a run reproducible like it does for the reference.
    The reference draws 1 + N*(corrector_steps+1) tensors per run (prior, then corrector
    and predictor alternately); ``count`` numbers them in that order.
    """

    def __init__(self, tape=None, seed=None, lengths=None):
        self.tape = tape
        if seed is None and tape is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        self.seed = 0 if seed is None else int(seed)
        self.count = 0
        # ragged batch: (B,) int32 valid sample counts; the fused updates keep the state at 0 past them
        self.lengths = lengths

    def begin_run(self):
        """Start of one sampler invocation.  A tape is replayed from draw 0 (parity runs number the draws per run);
        the Philox stream is NOT rewound: like the reference's global generator, a second call of the same sampler
        closure continues the stream and returns a fresh sample."""
        if self.tape is not None:
            self.count = 0

    reset = begin_run

    def next(self, like: torch.Tensor):
        """Returns (noise tensor or None, seed, draw index)."""
        i = self.count
        self.count += 1
        if self.tape is None:
            return None, self.seed, i
        z = self.tape(i, like)
        if z.shape != like.shape:
            raise ValueError(f"noise tape draw {i} has shape {tuple(z.shape)}, expected {tuple(like.shape)}")
        return z.to(device=like.device, dtype=torch.float32).contiguous(), self.seed, i


class SDE(abc.ABC):
    """SDE abstract class (sdes.py:44-179)."""

    def __init__(self, N):
        super().__init__()
        self.N = N

    @property
    @abc.abstractmethod
    def T(self):
        pass

    @abc.abstractmethod
    def sde(self, x, t, *args):
        pass

    @abc.abstractmethod
    def marginal_prob(self, x, t, *args):
        pass

    @abc.abstractmethod
    def prior_sampling(self, shape, *args):
        pass

    @abc.abstractmethod
    def prior_logp(self, z):
        pass

    @abc.abstractmethod
    def copy(self):
        pass

    def fused_params(self, snr: float = 0.0, n_src: int = 2, y=None) -> SdeParams:
        """Scalars of this SDE for the fused update kernel (ditsep_sde_params); y: the conditioning mixture (only
        PriorMixSDE's coefficients depend on it)."""
        raise NotImplementedError(f"SDE class {self.__class__.__name__} has no fused sm_100a update kernel")

    def discretize(self, x, t, *args, **kwargs):
        """x_{i+1} = x_i + f_i(x_i) + G_i z_i, Euler-Maruyama (sdes.py:94-108).

        ``getattr(kwargs, "dt", ...)`` on a dict never finds the key, so dt is ALWAYS 1/N in
        the reference; reproduced."""
        dt = 1 / self.N
        drift, diffusion = self.sde(x, t, *args)
        f = drift * dt
        G = diffusion * math.sqrt(dt)
        return f, G

    def reverse(oself, score_model, probability_flow=False):
        """Reverse-time SDE/ODE (sdes.py:110-175)."""
        N = oself.N
        T = oself.T
        sde_fn = oself.sde
        discretize_fn = oself.discretize
        fused_fn = oself.fused_params

        class RSDE(oself.__class__):
            def __init__(self):
                self.N = N
                self.probability_flow = probability_flow

            @property
            def T(self):
                return T

            def fused_params(self, snr: float = 0.0, n_src: int = 2, y=None):
                return fused_fn(snr, n_src, y=y)

            def sde(self, x, t, *args):
                parts = self.rsde_parts(x, t, *args)
                return parts["total_drift"], parts["diffusion"]

            def rsde_parts(self, x, t, *args):
                sde_drift, sde_diffusion = sde_fn(x, t, *args)
                pad_dim = (...,) + (None,) * (x.ndim - sde_diffusion.ndim)
                score = score_model(x, t, *args)
                score_drift = -sde_diffusion[pad_dim] ** 2 * score * (0.5 if self.probability_flow else 1.0)
                diffusion = torch.zeros_like(sde_diffusion) if self.probability_flow else sde_diffusion
                return {"total_drift": sde_drift + score_drift, "diffusion": diffusion, "sde_drift": sde_drift,
                        "sde_diffusion": sde_diffusion, "score_drift": score_drift, "score": score}

            def discretize(self, x, t, *args, **kwargs):
                f, G = discretize_fn(x, t, *args, **kwargs)
                pad_dim = (...,) + (None,) * (f.ndim - G.ndim)
                rev_f = f - G[pad_dim] ** 2 * score_model(x, t, *args) * (0.5 if self.probability_flow else 1.0)
                rev_G = torch.zeros_like(G) if self.probability_flow else G
                return rev_f, rev_G

        return RSDE()


@SDERegistry.register("mix")
class MixSDE(SDE):
    """dx = -lambda Pn x dt + g(t) dw, the DiffSep mixing SDE (sdes.py:182-351)."""

    def __init__(self, ndim, d_lambda, sigma_min, sigma_max, N=1000):
        super().__init__(N)
        self.ndim = ndim
        self.d_lambda = d_lambda
        self.sigma_min = sigma_min
        self.sigma_max = sigma_max
        self.ratiosig = self.sigma_max / self.sigma_min
        self.logsig = math.log(self.ratiosig)
        self.N = N
        self.A, self.Pn = self.get_mix_mat()

    def send_to(self, device):
        if self.A.device != device:
            self.A = self.A.to(device)
        if self.Pn.device != device:
            self.Pn = self.Pn.to(device)

    def get_mix_mat(self):
        ones = torch.ones((self.ndim, 1))
        avg_mat = ones @ ones.T / self.ndim
        return avg_mat[None, ...], (torch.eye(self.ndim) - avg_mat)[None, ...]

    def copy(self):
        return MixSDE(self.ndim, self.d_lambda, self.sigma_min, self.sigma_max, N=self.N)

    @property
    def T(self):
        return 1.0

    def fused_params(self, snr: float = 0.0, n_src: int = None, y=None) -> SdeParams:
        return SdeParams(SDE_MIX, self.ndim if n_src is None else n_src, float(self.d_lambda), float(self.sigma_min),
                         float(self.sigma_max), int(self.N), float(snr))

    def sde(self, x, t, mix=None):
        self.send_to(x.device)
        drift = -self.d_lambda * self.Pn @ x
        sigma = self.sigma_min * self.ratiosig ** t
        return drift, sigma * np.sqrt(2 * self.logsig)

    def _mean_mix_mat(self, t):
        decay = torch.exp(-t[:, None, None] * self.d_lambda)
        return self.A + decay * self.Pn

    def _mean(self, x0, t):
        return self._mean_mix_mat(t) @ x0

    def _cov_eigval(self, t):
        mult = self.sigma_min ** 2
        srp = self.ratiosig ** (2 * t)
        ev1 = mult * (srp - 1)
        ev2 = mult * (srp - torch.exp(-2.0 * self.d_lambda * t)) / (1.0 + self.d_lambda / self.logsig)
        return ev1, ev2

    def _var(self, t):
        ev1, ev2 = self._cov_eigval(t)
        return 0.5 * (ev1 + ev2)

    def _std(self, t):
        ev1, ev2 = self._cov_eigval(t)
        return ev1[:, None, None].sqrt() * self.A + ev2[:, None, None].sqrt() * self.Pn

    def marginal_prob(self, x0, t, *args):
        self.send_to(x0.device)
        return self._mean(x0, t), self._std(t)

    def sample_time_varprop(self, n, t_eps=0.0, device=None):
        """Times with density proportional to the noise standard deviation sqrt(_var(t)), by rejection under the
        constant envelope sqrt(_var(T)) (sdes.py:261-275)."""
        top = float(self._var(torch.tensor([float(self.T)], device=device)).sqrt())
        kept, have = [], 0
        while have < n:
            m = 3 * (n - have)
            t = torch.empty(m, device=device).uniform_(t_eps, self.T)
            u = torch.empty(m, device=device).uniform_(0, top)
            t = t[u < self._var(t).sqrt()]
            kept.append(t)
            have += t.shape[0]
        return torch.cat(kept)[:n]

    @staticmethod
    def mult_std(std, x):
        return std @ x

    @staticmethod
    def mult_std_inv(std, x):
        return torch.linalg.solve(std, x)

    def prior_sampling(self, shape, y, noise: NoiseSource = None):
        """x_T = 0.5 y + L(T) z, two sources hard-coded like the reference (sdes.py:336-348)."""
        if shape != y.shape:
            warnings.warn(f"Target shape {shape} does not match shape of y {y.shape}! Ignoring target shape.")
        noise = NoiseSource() if noise is None else noise
        x = torch.empty((y.shape[0], 2, y.shape[2]), device=y.device, dtype=torch.float32)
        z, seed, draw = noise.next(x)
        ops.pc_prior(self.fused_params(n_src=2), x, y.contiguous(), noise=z, seed=seed, draw=draw, lengths=noise.lengths)
        return x

    def prior_logp(self, z):
        raise NotImplementedError("prior_logp for OU SDE not yet implemented!")


@SDERegistry.register("priormix")
class PriorMixSDE(MixSDE):
    """MixSDE with a mixture-dependent diffusion (sdes.py:354-592): g(t, b, l) = g_mix(t) m[b, l], where
    m = 0.5 sqrt(max(box_avg_len(mix^2), 1e-4)) is a running energy envelope of the conditioning mixture (:480-492).
    Nothing else changes, so every closed form of MixSDE carries over with the factor m: terms linear in the diffusion
    (the noise of the prior, the corrector and the predictor; the marginal std L) scale by m, terms quadratic in it (the
    score terms) by m^2.  The fused kernel takes m as `ditsep_sde_params.elem_scale`; m is computed once per mixture
    tensor (`ditsep_energy_envelope`) and cached."""

    def __init__(self, ndim, d_lambda, sigma_min, sigma_max, N=1000, avg_len=510):
        super().__init__(ndim, d_lambda, sigma_min, sigma_max, N=N)
        self.avg_len = avg_len
        self._env_key, self._env = None, None

    def copy(self):
        return PriorMixSDE(self.ndim, self.d_lambda, self.sigma_min, self.sigma_max, N=self.N, avg_len=self.avg_len)

    def _std_sigma_mix(self, mix):
        """(B, 1, L) envelope m of `mix` (B, 1, L); one kernel launch per distinct mixture tensor."""
        key = (mix.data_ptr(), tuple(mix.shape), mix._version, str(mix.device))
        if key != self._env_key:
            if mix.dim() != 3 or mix.shape[1] != 1:
                raise ValueError(f"PriorMixSDE needs a one-channel mixture (B, 1, L), got {tuple(mix.shape)}")
            env = torch.empty(mix.shape, device=mix.device, dtype=torch.float32)
            ops.energy_envelope(mix.float().contiguous(), env, self.avg_len)
            self._env_key, self._env = key, env
        return self._env

    def fused_params(self, snr: float = 0.0, n_src: int = None, y=None) -> SdeParams:
        if y is None:
            raise ValueError("PriorMixSDE's coefficients depend on the mixture: pass y")
        p = super().fused_params(snr, n_src)
        env = self._std_sigma_mix(y)
        p.elem_scale = env.data_ptr()
        p._keep = env  # the struct holds a raw device pointer
        return p

    def sde(self, x, t, mix):
        drift, g = super().sde(x, t)
        g = g[(...,) + (None,) * (x.ndim - g.ndim)]
        return drift, g * torch.broadcast_to(self._std_sigma_mix(mix), x.shape)

    def _std(self, t, mix):
        """(B, n, n, L): the matrix square root of the covariance, per sample."""
        return super()._std(t)[..., None] * self._std_sigma_mix(mix)[:, None, :, :]

    def marginal_prob(self, x0, t, mix):
        self.send_to(x0.device)
        return self._mean(x0, t), self._std(t, mix)

    @staticmethod
    def mult_std(std, x):
        return torch.einsum("bcdt,bdt->bct", std, x)

    @staticmethod
    def mult_std_inv(std, x):
        return torch.linalg.solve(std.permute(0, 3, 1, 2), x.permute(0, 2, 1)[..., None])[..., 0].permute(0, 2, 1).contiguous()

    def prior_sampling(self, shape, mix, noise: NoiseSource = None):
        """x_T = 0.5 mix + m L(T) z for a one-channel mixture (sdes.py:566-589)."""
        if shape != mix.shape:
            warnings.warn(f"Target shape {shape} does not match shape of mix {mix.shape}! Ignoring target shape.")
        if mix.shape[1] != 1:
            raise NotImplementedError("PriorMixSDE.prior_sampling: only the one-channel mixture form has a fused kernel")
        noise = NoiseSource() if noise is None else noise
        x = torch.empty((mix.shape[0], self.ndim, mix.shape[2]), device=mix.device, dtype=torch.float32)
        z, seed, draw = noise.next(x)
        ops.pc_prior(self.fused_params(n_src=self.ndim, y=mix), x, mix.contiguous(), noise=z, seed=seed, draw=draw,
                     lengths=noise.lengths)
        return x


@SDERegistry.register("ouve")
class OUVESDE(SDE):
    """dx = theta (y - x) dt + g(t) dw (sdes.py:595-698)."""

    def __init__(self, theta, sigma_min, sigma_max, N=1000, **ignored_kwargs):
        super().__init__(N)
        self.theta = theta
        self.sigma_min = sigma_min
        self.sigma_max = sigma_max
        self.logsig = np.log(self.sigma_max / self.sigma_min)
        self.N = N

    def copy(self):
        return OUVESDE(self.theta, self.sigma_min, self.sigma_max, N=self.N)

    @property
    def T(self):
        return 1

    def fused_params(self, snr: float = 0.0, n_src: int = 2, y=None) -> SdeParams:
        return SdeParams(SDE_OUVE, int(n_src), float(self.theta), float(self.sigma_min), float(self.sigma_max),
                         int(self.N), float(snr))

    def sde(self, x, t, y):
        drift = self.theta * (y - x)
        sigma = self.sigma_min * (self.sigma_max / self.sigma_min) ** t
        return drift, sigma * np.sqrt(2 * self.logsig)

    def _mean(self, x0, t, y):
        e = torch.exp(-self.theta * t)
        e = e[(...,) + (None,) * (x0.ndim - e.ndim)]
        return e * x0 + (1 - e) * y

    def _std(self, t):
        s, th, ls = self.sigma_min, self.theta, self.logsig
        return torch.sqrt((s ** 2 * torch.exp(-2 * th * t) * (torch.exp(2 * (th + ls) * t) - 1) * ls) / (th + ls))

    def marginal_prob(self, x0, t, y):
        return self._mean(x0, t, y), self._std(t)

    def prior_sampling(self, shape, y, noise: NoiseSource = None):
        """x_T = y + std(1) z with z of `shape` (sdes.py:687-695)."""
        if shape != y.shape:
            warnings.warn(f"Target shape {shape} does not match shape of y {y.shape}! Brodcasting mix shape.")
        noise = NoiseSource() if noise is None else noise
        x = torch.empty(tuple(shape), device=y.device, dtype=torch.float32)
        z, seed, draw = noise.next(x)
        ops.pc_prior(self.fused_params(n_src=shape[1]), x, y.contiguous(), noise=z, seed=seed, draw=draw,
                     lengths=noise.lengths)
        return x

    def prior_logp(self, z):
        raise NotImplementedError("prior_logp for OU SDE not yet implemented!")


@SDERegistry.register("sbve")
class SBVESDE(SDE):
    """Schroedinger bridge with a variance-exploding diffusion g(t) = sqrt(c) k^t, alpha = 1 (sdes.py:701-779; Jukic et
    al., "Schroedinger Bridge for Generative Speech Enhancement", 2024).  The closed forms below are per-batch scalars
    (a handful of flops on (B,) tensors) and stay in torch; the per-sample work of the samplers built on them runs in
    ``ops.sb_step``, which takes the coefficient tables ``step_table`` prepares."""

    def __init__(self, k, c, N=50, eps=1e-8, sampler_type="ode", **ignored_kwargs):
        super().__init__(N)
        self.k = k
        self.c = c
        self.N = N
        self.eps = eps
        self.sampler_type = sampler_type

    def copy(self):
        return SBVESDE(self.k, self.c, N=self.N)  # like the reference (:730), eps and sampler_type go back to defaults

    @property
    def T(self):
        return 1

    def sde(self, x, y, t):
        return 0.0, torch.sqrt(torch.tensor(self.c)) * self.k ** t

    def _sigmas_alphas(self, t):
        """sdes.py:741-752, fp32 like the reference."""
        alpha_t = torch.ones_like(t)
        alpha_T = torch.ones_like(t)
        log_k = torch.log(torch.tensor(self.k))
        sigma_t = torch.sqrt((self.c * (self.k ** (2 * t) - 1.0)) / (2 * log_k))
        sigma_T = torch.sqrt((self.c * (self.k ** (2 * self.T) - 1.0)) / (2 * log_k))
        alpha_bart = alpha_t / (alpha_T + self.eps)
        sigma_bart = torch.sqrt(sigma_T ** 2 - sigma_t ** 2 + self.eps)
        return sigma_t, sigma_T, sigma_bart, alpha_t, alpha_T, alpha_bart

    def _mean(self, x0, y, t):
        sigma_t, sigma_T, sigma_bart, alpha_t, alpha_T, alpha_bart = self._sigmas_alphas(t)
        w_xt = alpha_t * sigma_bart ** 2 / (sigma_T ** 2 + self.eps)
        w_yt = alpha_bart * sigma_t ** 2 / (sigma_T ** 2 + self.eps)
        pad_dim = (...,) + (None,) * (x0.ndim - t.ndim)
        return w_xt[pad_dim] * x0 + w_yt[pad_dim] * y

    def _std(self, t):
        sigma_t, sigma_T, sigma_bart, alpha_t, alpha_T, alpha_bart = self._sigmas_alphas(t)
        return (alpha_t * sigma_bart * sigma_t) / (sigma_T + self.eps)

    def marginal_prob(self, x0, y, t):
        return self._mean(x0, y, t), self._std(t)

    def prior_sampling(self, shape, y, noise: NoiseSource = None):
        if shape != y.shape:
            warnings.warn(f"Target shape {shape} does not match shape of y {y.shape}! Ignoring target shape.")
        return y

    def prior_logp(self, z):
        raise NotImplementedError("prior_logp for SBVE SDE not yet implemented!")

    def step_table(self, time_steps: torch.Tensor, sampler_type: str) -> torch.Tensor:
        """(N, 4) f32 rows [w_prev, w_est, w_y, w_z] of the first-order SB discretisation on the grid ``time_steps``
        (N + 1 points, CPU), formed with the reference's own expressions (sdes/__init__.py:305-309 "sde", :353-364
        "ode"); the last "sde" step has w_z = 0 (:325-326)."""
        ts = time_steps.detach().to("cpu", torch.float32)
        sig, sig_T, sigb, al, al_T, _ = self._sigmas_alphas(ts[1:])
        sig_p, _, sigb_p, al_p, _, _ = self._sigmas_alphas(ts[:-1])
        zero = torch.zeros_like(sig)
        if sampler_type == "sde":
            w_prev = al * sig ** 2 / (al_p * sig_p ** 2 + self.eps)
            tmp = 1 - sig ** 2 / (sig_p ** 2 + self.eps)
            w_est = al * tmp
            w_z = al * sig * torch.sqrt(tmp)
            w_z[-1] = 0.0
            return torch.stack([w_prev, w_est, zero, w_z], dim=1)
        if sampler_type == "ode":
            w_prev = al * sig * sigb / (al_p * sig_p * sigb_p + self.eps)
            w_est = al / (sig_T ** 2 + self.eps) * (sigb ** 2 - sigb_p * sig * sigb / (sig_p + self.eps))
            w_mix = al / (al_T * sig_T ** 2 + self.eps) * (sig ** 2 - sig_p * sig * sigb / (sigb_p + self.eps))
            return torch.stack([w_prev, w_est, w_mix, zero], dim=1)
        raise ValueError("Invalid type. Choose 'ode' or 'sde'.")
