"""``DiffSepModel``: the API surface of the separation sampler (mirror of src/diffsep.py).

Only the inference side of the reference Lightning module is on the hot path and is kept:
``__init__`` (diffsep.py:28-80: build score model + SDE from the Hydra config),
``forward`` (:341-343), ``separate`` (:82-98), ``get_pc_sampler`` (:623-695),
``normalize_batch`` / ``denormalize_batch``, ``eval(no_ema=...)``; and of ``DiffSepOU`` (:729-1217) its EDM-preconditioned
``forward``, ``separate`` and ``get_sb_sampler``.  Training losses, EMA
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
        # validation-side settings (diffsep.py:47-62); only read by the score-loss forward below
        self.time_sampling_strategy = getattr(model_cfg, "time_sampling_strategy", "uniform")
        self.init_hack = getattr(model_cfg, "init_hack", False)
        self.init_hack_p = getattr(model_cfg, "init_hack_p", 1.0 / self.sde.N)
        self.t_rev_init = getattr(model_cfg, "t_rev_init", 0.03)
        self.train_source_order = getattr(model_cfg, "train_source_order", "random")

    def forward(self, xt, time, mix):
        return self.score_model(xt, time, mix)

    def note_constant_mixture(self, mix):
        """Sampler hook (sdes/__init__.py): `mix` is passed unchanged to every evaluation of the coming run."""
        hook = getattr(self.score_model, "note_constant_mixture", None)
        if hook is not None:
            hook(mix)

    # ---- score-matching loss, FORWARD VALUE only (diffsep.py:100-181, 296-358, 476-487) ----------------
    # What validation_step logs as val/score_loss.  No backward kernels exist in this library: these run under no_grad.
    def sample_time(self, x):
        """diffsep.py:100-111."""
        if self.time_sampling_strategy == "uniform":
            return x.new_zeros(x.shape[0]).uniform_(self.t_eps, self.t_max)
        if self.time_sampling_strategy == "varprop":
            return self.sde.sample_time_varprop(x.shape[0], t_eps=self.t_eps, device=x.device)
        raise NotImplementedError(f"No sampling strategy {self.time_sampling_strategy}")

    def _draw(self, like, noise):
        """The normal draw of one loss evaluation: `noise` (a tensor, parity runs) or the library's Philox stream."""
        if noise is not None:
            return noise.to(device=like.device, dtype=torch.float32).contiguous()
        from . import ops
        z = torch.empty_like(like, dtype=torch.float32)
        ops.philox_normal(z, int(torch.randint(0, 2 ** 62, (1,)).item()), 0)
        return z

    def sample_prior(self, mix, target, time=None, noise=None):
        """diffsep.py:113-181: (x_t, time, L, z).  The default branch (init_hack not in 1..4, which includes the
        configured init_hack = 5) is one fused launch (ops.pc_perturb); the four "init hacks" that redefine the mean or
        the noise near t = T are the reference's small (B, n, n) algebra around the same kernel's mean output."""
        from . import ops
        time = self.sample_time(target) if time is None else time
        hack = self.init_hack if self.init_hack in (1, 2, 3, 4) else 0
        z = self._draw(target, noise)
        sel4 = None
        if hack == 4:
            sel4 = torch.rand_like(time) < 1 / self.sde.N
            time = torch.where(sel4, time.new_ones(time.shape) * self.sde.T, time)
        if hasattr(self.sde, "send_to"):
            self.sde.send_to(target.device)
        L = self.sde._std(time)
        x_t = target.to(torch.float32).contiguous().clone()
        mean = torch.empty_like(x_t) if hack else None
        ops.pc_perturb(self.sde.fused_params(n_src=target.shape[1]), x_t, time.to(torch.float32).contiguous(),
                       y=mix.to(torch.float32).contiguous(), noise=z, x_mean=mean)
        if not hack:
            return x_t, time, L, z
        true_mix = torch.broadcast_to(mix, target.shape) / target.shape[1]
        if hack in (1, 4):
            sel = sel4 if hack == 4 else ~(time < self.sde.T - self.t_rev_init)
            sel = torch.broadcast_to(sel[:, None, None], z.shape)
            z = torch.where(sel, z + self.sde.mult_std_inv(L, true_mix - mean), z)
            return mean + self.sde.mult_std(L, z), time, L, z
        T, Tm = self.sde.T, self.sde.T - self.t_rev_init
        beta = torch.clamp((time - Tm) / (T - Tm), min=0.0, max=1.0)[:, None, None]
        x_t = true_mix * beta + mean * (1.0 - beta) + self.sde.mult_std(L, z)
        if hack == 3:
            z = self.sde.mult_std_inv(L, x_t - mean)
        return x_t, time, L, z

    def _loss_reduce(self, per_item):
        """config.model.loss is torch.nn.MSELoss in every reference config; reduction "none" (init_hack 5..7,
        diffsep.py:64-69) keeps (B,), the default "mean" gives the batch mean."""
        loss_cfg = _to_plain(getattr(self.config.model, "loss", None)) or {}
        reduction = "none" if self.init_hack in (5, 6, 7) else loss_cfg.get("reduction", "mean")
        return per_item if reduction == "none" else per_item.mean()

    @torch.no_grad()
    def compute_score_loss(self, mix, target, time=None, noise=None):
        """diffsep.py:344-358: E (L s(x_t, t) + z)^2; perturbation, network and ONE reduction kernel (ops.score_loss)."""
        from . import ops
        x_t, time, L, z = self.sample_prior(mix, target, time=time, noise=noise)
        pred = self(x_t, time, mix).to(torch.float32).contiguous()
        per_item = ops.score_loss(self.sde.fused_params(n_src=target.shape[1]), pred, time.to(torch.float32).contiguous(),
                                  noise=z.contiguous())  # linalg.solve (init_hack 1 / 3 / 4) returns column-major batches
        return self._loss_reduce(per_item)

    @torch.no_grad()
    def compute_score_loss_init_hack_pit(self, mix, target, noise=None):
        """diffsep.py:305-338: t = T, the state is the true mixture plus L z0 whatever the source order, and the loss is
        the minimum over source permutations of the noise each permutation would explain."""
        import itertools
        from . import ops
        time = mix.new_ones(mix.shape[0]) * self.sde.T
        true_mix = torch.broadcast_to(mix, target.shape) / target.shape[1]
        z0 = self._draw(target, noise)
        p = self.sde.fused_params(n_src=target.shape[1])
        losses = []
        for perm in itertools.permutations(range(target.shape[1])):
            mean, L = self.sde.marginal_prob(target[:, perm, :], time, mix)
            z = (z0 + self.sde.mult_std_inv(L, true_mix - mean)).contiguous()
            x_t = mean + self.sde.mult_std(L, z)
            pred = self(x_t, time, mix).to(torch.float32).contiguous()
            losses.append(ops.score_loss(p, pred, time, noise=z))
        return torch.stack(losses, dim=1).min(dim=1).values

    @torch.no_grad()
    def compute_score_loss_with_pit(self, mix, target, time=None, noise=None, select=None):
        """diffsep.py:183-258: x_t is built around the mean of ONE randomly chosen source order; where the other orders'
        means are within `mmnr_thresh_pit` dB of the noise, the loss is the minimum over the orders (each order explains
        x_t with its own noise), elsewhere the plain loss.  Returns the losses of the PIT items followed by those of the
        others, like the reference.  The network output does not depend on the order, so it is evaluated once per item
        (the reference replicates the same input n_perm times)."""
        import itertools
        from . import ops
        time = self.sample_time(target) if time is None else time
        perms = list(itertools.permutations(range(target.shape[1])))
        means, L = [], None
        for p in perms:
            mean, L = self.sde.marginal_prob(target[:, p, :], time, mix)
            means.append(mean)
        means = torch.stack(means, dim=1)  # (batch, perm, src, samples)
        z = self._draw(target, noise)
        Lz = self.sde.mult_std(L, z)
        mean_select = utils.select_elem_at_random(means, dim=1, select=select)
        xt = (mean_select + Lz[:, None, ...])[:, 0]
        err = means - mean_select
        n_elems = (means.shape[1] - 1) * means.shape[2] * means.shape[3]
        err_pow = err.square().sum(dim=(1, 2, 3)) / n_elems
        noise_pow = Lz.square().mean(dim=(1, 2))
        mmnr = 10.0 * torch.log10(err_pow / noise_pow.clamp(min=1e-5))
        select_pit = mmnr < self.config.model.mmnr_thresh_pit
        pred = self(xt.contiguous(), time, mix).to(torch.float32).contiguous()
        par = self.sde.fused_params(n_src=target.shape[1])
        tc = time.to(torch.float32).contiguous()
        plain = ops.score_loss(par, pred, tc, noise=z.contiguous())
        per_perm = torch.stack([ops.score_loss(par, pred, tc, noise=(z + self.sde.mult_std_inv(L, err[:, k])).contiguous())
                                for k in range(len(perms))], dim=1)
        return torch.cat([per_perm[select_pit].min(dim=-1).values, plain[~select_pit]])

    @torch.no_grad()
    def compute_score_loss_with_pit_allthetime(self, mix, target, time=None, noise=None, shuffle=True):
        """diffsep.py:260-303: x_t around a randomly ordered target; the loss is the minimum over the source orders of the
        noise each order would need (z_p = z_0 + L^{-1}(mean_0 - mean_p)).  One network evaluation (x_t does not
        depend on p; the reference recomputes it per order)."""
        import itertools
        from . import ops
        time = self.sample_time(target) if time is None else time
        if shuffle:
            target = utils.shuffle_sources(target)
        mean_0, L = self.sde.marginal_prob(target, time, mix)
        z0 = self._draw(target, noise)
        x_t = mean_0 + self.sde.mult_std(L, z0)
        pred = self(x_t, time, mix).to(torch.float32).contiguous()
        par = self.sde.fused_params(n_src=target.shape[1])
        tc = time.to(torch.float32).contiguous()
        losses = []
        for p in itertools.permutations(range(target.shape[1])):
            mean_p, _ = self.sde.marginal_prob(target[:, p, :], time, mix)
            z_p = (z0 + self.sde.mult_std_inv(L, mean_0 - mean_p)).contiguous()
            losses.append(ops.score_loss(par, pred, tc, noise=z_p))
        return torch.stack(losses, dim=0).min(dim=0).values

    def _train_step_init(self, mix, target, rest_fn, pit=None):
        """diffsep.py:363-420: a random subset (probability init_hack_p) takes the t = T permutation-invariant loss, the
        rest `rest_fn`; the mean over all items."""
        if pit is None:
            pit = mix.new_zeros(mix.shape[0]).uniform_() < self.init_hack_p
        losses = []
        if pit.sum() > 0:
            losses.append(self.compute_score_loss_init_hack_pit(mix[pit], target[pit]))
        if pit.sum() != mix.shape[0]:
            losses.append(rest_fn(mix[~pit], target[~pit]))
        return torch.cat(losses).mean()

    @torch.no_grad()
    def train_step_init_6(self, mix, target, pit=None):
        return self._train_step_init(mix, target, lambda m, t: self.compute_score_loss_with_pit(m, utils.shuffle_sources(t)), pit)

    @torch.no_grad()
    def train_step_init_7(self, mix, target, pit=None):
        return self._train_step_init(mix, target, self.compute_score_loss_with_pit_allthetime, pit)

    @torch.no_grad()
    def train_step_init_5(self, mix, target, pit=None):
        """diffsep.py:363-382 (forward value): a random subset (probability init_hack_p) takes the permutation-invariant
        t = T loss, the rest the plain loss on randomly ordered sources."""
        if pit is None:
            pit = mix.new_zeros(mix.shape[0]).uniform_() < self.init_hack_p
        losses = []
        if pit.sum() > 0:
            losses.append(self.compute_score_loss_init_hack_pit(mix[pit], target[pit]))
        if pit.sum() != mix.shape[0]:
            losses.append(self.compute_score_loss(mix[~pit], utils.shuffle_sources(target[~pit])))
        return torch.cat(losses).mean()

    @torch.no_grad()
    def validation_score_loss(self, batch):
        """The val/score_loss of validation_step (diffsep.py:476-487): init_hack 5 / 6 / 7 or the plain loss."""
        (mix, target), *stats = self.normalize_batch(batch)
        if self.init_hack == 7:
            return self.train_step_init_7(mix, target)
        if self.init_hack == 6:
            return self.train_step_init_6(mix, target)
        if self.init_hack == 5:
            return self.train_step_init_5(mix, target)
        return self.compute_score_loss(mix, target)

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


class DiffSepOU(DiffSepModel):
    """Inference side of the reference's second Lightning module (diffsep.py:729-1217): the same score network behind
    an ``OUVESDE`` (PC sampler with the "ald" corrector) or an ``SBVESDE`` -- then ``forward`` wraps the network in the
    EDM preconditioning (:813-828) and ``separate`` runs the Schroedinger-bridge sampler (:1204-1217)."""

    def __init__(self, config):
        super().__init__(config)
        model_cfg = self.config.model
        self.is_edm = isinstance(self.sde, sdes.SBVESDE)
        if self.is_edm:
            self.network_scaling = getattr(model_cfg, "network_scaling", "1/sigma")
            self.c_in = self.c_out = self.c_skip = getattr(model_cfg, "c", "edm")
            self.sigma_data = getattr(model_cfg, "sigma_data", 0.1)

    # -- coefficients (diffsep.py:1165-1202).  The reference's forward always passes a non-empty pad_dim, which selects
    # the first branch of each; the other (pad_dim=None) branches are different formulae and are kept as they are.
    def _c_in(self, t, pad_dim=None):
        if self.c_in == "1":
            return 1.0
        if self.c_in == "edm":
            sigma = self.sde._std(t)
            if pad_dim:
                return (self.sigma_data ** 2 / (sigma ** 2 + self.sigma_data ** 2))[pad_dim]
            return 1.0 / torch.sqrt(sigma ** 2 + self.sigma_data ** 2)
        raise ValueError("Invalid c_in type: {}".format(self.c_in))

    def _c_out(self, t, pad_dim=None):
        if self.c_out == "1":
            return 1.0
        if self.c_out == "sigma":
            return self.sde._std(t)
        if self.c_out == "1/sigma":
            return 1.0 / self.sde._std(t)[pad_dim] if pad_dim else 1.0 / self.sde._std(t)
        if self.c_out == "edm":
            sigma = self.sde._std(t)
            v = (sigma * self.sigma_data) / torch.sqrt(self.sigma_data ** 2 + sigma ** 2)
            return v[pad_dim] if pad_dim else v
        raise ValueError("Invalid c_out type: {}".format(self.c_out))

    def _c_skip(self, t, pad_dim=None):
        if self.c_skip == "0":
            return 0.0
        if self.c_skip == "edm":
            sigma = self.sde._std(t)
            if pad_dim:
                return (sigma ** 2 / (sigma ** 2 + self.sigma_data ** 2))[pad_dim]
            return self.sigma_data ** 2 / (sigma ** 2 + self.sigma_data ** 2)
        raise ValueError("Invalid c_skip type: {}".format(self.c_skip))

    def edm_table(self, times: torch.Tensor):
        """(len(times), 4) f32 CPU rows [c_in, c_skip, c_out, f_div] of ``forward`` at each time (what
        ``ops.sb_step`` folds in), or None when this model is not preconditioned."""
        if not self.is_edm:
            return None
        t = times.detach().to("cpu", torch.float32).reshape(-1)
        pad = (...,)
        one = torch.ones_like(t)
        if self.network_scaling == "1/sigma":
            div = self.sde._std(t)
        elif self.network_scaling == "1/t":
            div = t
        else:
            div = one
        return torch.stack([one * self._c_in(t, pad), one * self._c_skip(t, pad), one * self._c_out(t, pad), div], dim=1)

    def forward(self, xt, time, mix):
        if not self.is_edm:
            return self.score_model(xt, time, mix)
        from . import ops
        B = xt.shape[0]
        tab = self.edm_table(time.reshape(-1)[:1].expand(B) if time.numel() == 1 else time).to(xt.device)
        xt = xt.to(torch.float32).contiguous()
        mix = mix.to(torch.float32).contiguous()
        coef = torch.zeros(2, B, 8, device=xt.device)
        coef[0, :, 0] = 1.0  # start-up row: x_in = c_in x, y_in = c_in y
        coef[0, :, 5:7] = 1.0
        coef[0, :, 7] = tab[:, 0]
        coef[1, :, 1] = 1.0  # x' = est = c_skip x + c_out (F / f_div)
        coef[1, :, 4:7] = tab[:, 1:4]
        coef[1, :, 7] = 1.0
        x_in, y_in = torch.empty_like(xt), torch.empty_like(mix)
        ops.sb_step(xt, None, mix, coef[0], x_in=x_in, y_in=y_in)
        f = self.score_model(x_in, time, y_in).contiguous()
        out = torch.empty_like(xt)
        ops.sb_step(xt, f, mix, coef[1], est=out)
        return out

    # ---- score-matching loss, FORWARD VALUE only (diffsep.py:796-811, 832-899, 931-942) -------------------
    def sample_time(self, x):
        return x.new_zeros(x.shape[0]).uniform_(self.t_eps, self.t_max)

    def sample_prior(self, mix, target, time=None, noise=None):
        """diffsep.py:800-811: x_t = mean(target, mix, t) + std(t) z, returned with (time, sigma broadcastable, z).  One
        fused launch either way: the OU perturbation row of the PC coefficient table (OUVESDE) or ops.sb_step in its
        "mean, then noise" mode with SBVESDE's host-side closed forms as the coefficient row."""
        from . import ops
        time = self.sample_time(target) if time is None else time
        z = self._draw(target, noise)
        tgt = target.to(torch.float32).contiguous()
        mixc = mix.to(torch.float32).contiguous()
        std = self.sde._std(time)
        if self.is_edm:
            sigma_t, sigma_T, sigma_bart, alpha_t, alpha_T, alpha_bart = self.sde._sigmas_alphas(time)
            coef = torch.zeros(target.shape[0], 8, device=target.device)
            coef[:, 0] = alpha_t * sigma_bart ** 2 / (sigma_T ** 2 + self.sde.eps)    # w_xt (sdes.py:757)
            coef[:, 2] = alpha_bart * sigma_t ** 2 / (sigma_T ** 2 + self.sde.eps)    # w_yt (:758)
            coef[:, 3] = std
            coef[:, 5:] = 1.0
            x_t = torch.empty_like(tgt)
            ops.sb_step(tgt.reshape(tgt.shape[0], tgt.shape[1], -1), None, mixc, coef, x_out=x_t, noise=z, use_noise=2)
        else:
            x_t = tgt.clone()
            ops.pc_perturb(self.sde.fused_params(n_src=target.shape[1]), x_t, time.to(torch.float32).contiguous(), y=mixc,
                           noise=z)
        return x_t, time, std[(...,) + (None,) * (x_t.ndim - std.ndim)], z

    def _step(self, y, x, time=None, noise=None):
        x_t, t, sigma, z = self.sample_prior(y, x, time=time, noise=noise)
        return self(x_t, t, y), sigma, z

    def _per_source_loss(self, pred, sigma, z):
        """mean over the trailing axes of (sigma pred + z)^2 -> (B, n_src), the reference's
        `loss.mean(dim=tuple(range(2 - loss.ndim, 0)))` of an un-reduced MSELoss (diffsep.py:839-840)."""
        from . import ops
        B, C = pred.shape[:2]
        flat = pred.to(torch.float32).contiguous().reshape(B * C, 1, -1)
        sig = sigma.reshape(B).to(torch.float32).repeat_interleave(C).contiguous()
        return ops.score_loss_sigma(flat, sig, noise=z.contiguous().reshape(B * C, 1, -1)).reshape(B, C)

    @torch.no_grad()
    def compute_score_loss(self, y, x, time=None, noise=None):
        """diffsep.py:836-841; (B, n_src) for the un-reduced loss of init_hack 5, else the scalar mean."""
        pred, sigma, z = self._step(y, x, time=time, noise=noise)
        return self._loss_reduce(self._per_source_loss(pred, sigma, z))

    @torch.no_grad()
    def compute_score_loss_init_hack_pit(self, mix, target, noise=None):
        """diffsep.py:843-869 for the OU SDE.  With an SBVESDE the reference multiplies a (B,) std into a (B, C, T) tensor
        (`sigma = std if self.is_edm`, :854) and fails for B > 1; its SB config never reaches it (init_hack_p = 0)."""
        import itertools
        if self.is_edm:
            raise NotImplementedError("compute_score_loss_init_hack_pit with an SBVESDE: shape error in the reference "
                                      "(src/diffsep.py:854-858); the diffsep_sb config sets init_hack_p = 0")
        time = mix.new_ones(mix.shape[0]) * self.sde.T
        z0 = self._draw(target, noise)
        losses = []
        for perm in itertools.permutations(range(target.shape[1])):
            mean, std = self.sde.marginal_prob(target[:, perm, ...], time, mix)
            sigma = std[(...,) + (None,) * (mix.ndim - time.ndim)]
            z = z0 + (mix - mean) / sigma
            pred = self(mix + sigma * z0, time, mix)
            losses.append(self._per_source_loss(pred, sigma, z))
        return torch.stack(losses, dim=1).min(dim=1).values

    @torch.no_grad()
    def validation_score_loss(self, batch):
        """The val/score_loss of DiffSepOU.validation_step (diffsep.py:931-942)."""
        (mix, target), *stats = self.normalize_batch(batch)
        if self.init_hack == 5:
            return self.train_step_init_5(mix, target)
        return self.compute_score_loss(mix, target)

    def separate(self, mix, **kwargs):
        """diffsep.py:1204-1217: SB sampler for an SBVESDE, else reverse-diffusion + "ald" PC sampler; the result tuple
        is returned as the sampler gives it (not denormalised)."""
        (mix, _), *stats = self.normalize_batch((mix, None))
        sampler_kwargs = dict(_to_plain(self.config.model.sampler))
        sampler_kwargs.update(kwargs)
        if self.is_edm:
            sampler = self.get_sb_sampler(self.sde, mix, **sampler_kwargs)
        else:
            sampler = self.get_pc_sampler("reverse_diffusion", "ald", mix, **sampler_kwargs)
        return sampler()

    def get_sb_sampler(self, sde, y, sampler_type="ode", N=None, **kwargs):
        """diffsep.py:1159-1163 (the sde argument is only read for its N; the module's own SDE is copied)."""
        N = sde.N if N is None else N
        sde = self.sde.copy()
        sde.N = N
        return sdes.get_sb_sampler(sde, self, y=y, sampler_type=sampler_type, **kwargs)


def sb_default_config(nf=64, N=30, precision=None, fs=8000):
    """src/config/diffsep_sb/model/default.yaml as a plain dict: the hot-path network behind SBVESDE(k=2.6, c=0.4)."""
    cfg = default_config(nf=nf, n_src=2, N=N, precision=precision, fs=fs)
    cfg["model"]["sde"] = {"_target_": "sdes.SBVESDE", "ndim": 2, "k": 2.6, "c": 0.4, "eps": 1e-8, "N": N,
                           "sampler_type": "ode"}
    cfg["model"]["sampler"] = {"N": N}
    return cfg


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
