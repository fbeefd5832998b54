"""Shared builders for the parity tests (TEST INFRASTRUCTURE)."""
import torch

from ditsep_b200.diffsep import DiffSepModel, default_config
from ditsep_b200.utils.synth import synth_mixtures
from oracle import ncsnpp_ref as R
from oracle import sampler_ref as S


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-30))


def build_model(precision="fp32", device="cpu", nf=64, n_src=2, N=30, seed=0, init_scale=1.0):
    """DiffSepModel with the oracle's deterministic weights (same dict the reference was pinned with)."""
    cfg = R.ScoreConfig(num_sources=n_src, backbone=R.NCSNppConfig(nf=nf))
    sd = R.make_state_dict(cfg.backbone, seed=seed, init_scale=init_scale)
    model = DiffSepModel(default_config(nf=nf, n_src=n_src, N=N, precision=precision))
    model.score_model.load_state_dict(sd, strict=True)
    return model.to(device).eval(), sd, cfg


def score_inputs(B, T, seed=1234):
    """The inputs pin_against_reference.py used for the one-score-evaluation fixtures."""
    mix, src = synth_mixtures(B, 2, T, seed=seed)
    mixn, _, _ = S.normalize_batch_ref(mix)
    xt = 0.5 * mixn + 0.3 * S.NoiseTape(99).draw(0, (B, 2, T))
    t = torch.full((B,), 0.61)
    return xt, t, mixn


def sampler_inputs(B, T, seed=1234):
    mix, src = synth_mixtures(B, 2, T, seed=seed)
    mixn, _, _ = S.normalize_batch_ref(mix)
    return mixn, src
